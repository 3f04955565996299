from .misc import (AverageMeter, accuracy, clip_logit_scale_step, flush_meters,  # noqa: F401
                   reduce_update_many)
