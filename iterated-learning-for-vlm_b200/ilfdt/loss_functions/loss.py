"""Drop-in for prototype/loss_functions/loss.py:27-50."""
import torch
from torch.nn.modules.loss import _Loss

from .. import functional as Fn
from .. import linklink as link
from ..lazy import LazyLogits


class ClipInfoCELoss(_Loss):
    """forward(logits_per_image [b,B], logits_per_text [b,B]) -> (loss, labels).

    labels = arange(b) when B == b, else rank*b + arange(b) (loss.py:41-45);
    loss = (CE(logits_per_image, labels) + CE(logits_per_text, labels)) / 2 (loss.py:47-49).
    When both arguments are the LazyLogits of one model.forward, logits + both cross-entropies (and
    their backward) run in the fused kernels and the [b,B] matrices are never formed. Plain tensors take
    the stored-logits kernels (ilfdt_ce_rows_*)."""

    def __init__(self):
        super().__init__()

    def forward(self, logits_per_image, logits_per_text):
        bs, l_bs = logits_per_image.shape
        device = logits_per_image.device
        offset = 0 if l_bs == bs else link.get_rank() * bs
        labels = offset + torch.arange(0, bs, dtype=torch.long, device=device)
        labels._ilfdt_label_offset = offset   # lets ilfdt.utils.accuracy recognise these labels without reading them
        if (isinstance(logits_per_image, LazyLogits) and isinstance(logits_per_text, LazyLogits)
                and logits_per_image._stash is logits_per_text._stash
                and logits_per_image._which == 0 and logits_per_text._which == 1):
            return logits_per_image._stash.fused_loss(), labels
        li = logits_per_image.materialize() if isinstance(logits_per_image, LazyLogits) else logits_per_image
        lt = logits_per_text.materialize() if isinstance(logits_per_text, LazyLogits) else logits_per_text
        loss_i = Fn.CrossEntropyRowsFunction.apply(li, offset)
        loss_t = Fn.CrossEntropyRowsFunction.apply(lt, offset)
        return (loss_i + loss_t) / 2, labels
