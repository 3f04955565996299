"""Drop-in for prototype/utils/misc.py:464-477 (`accuracy`), the per-step consumer of logits_per_image at
example/clip_fdt/train_solver.py:421."""
import torch

from ..lazy import LazyLogits


def accuracy(output, target, topk=(1,)):
    """Computes the precision@k for the specified values of k -> list of [1] tensors (percent), as the reference.

    When `output` is the LazyLogits of logits_per_image, `target` are the labels the drop-in ClipInfoCELoss returned
    for the same step, and every k is 1 or 5 (the reference's only call), the hits are counted by the fused kernel
    (ilfdt_nce_topk_hits) straight from the embeddings: the [b,B] logits are never stored. Everything else takes
    the reference's own route (topk on the materialised matrix)."""
    batch_size = target.size(0)
    if (isinstance(output, LazyLogits) and output._which == 0 and all(k in (1, 5) for k in topk)
            and getattr(target, "_ilfdt_label_offset", None) == output._stash.label_offset()
            and batch_size == output._stash.b):
        counts = output._stash.topk_hits().float()
        return [counts[0 if k == 1 else 1].reshape(1).mul_(100.0 / batch_size) for k in topk]
    top = output.topk(max(topk), dim=1).indices                 # [b, maxk], best first
    hit = top.eq(target.view(-1, 1))
    return [hit[:, :k].any(dim=1).float().sum().reshape(1) * (100.0 / batch_size) for k in topk]
