"""Drop-in for prototype/utils/misc.py:464-477 (`accuracy`), the per-step consumer of logits_per_image at
example/clip_fdt/train_solver.py:421, and for the meters / parameter clip of the solver step whose
host reads stall the device."""
import torch
import torch.distributed as dist

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


# ---- the solver's per-step host stalls --------------------------------------------------
class AverageMeter(object):
    """prototype/utils/misc.py:22-56: same attributes (val, avg, sum, count / history) and update rule.
    `reduce_update` keeps the reference's semantics (all-reduce, then ONE host read). The solver calls it three
    times per step (train_solver.py:447-449: loss, top1, top5) = three all-reduces and three blocking `.item()`
    reads; `reduce_update_many` does the same bookkeeping with one all-reduce and at most one read."""

    def __init__(self, length=0):
        self.length = length
        self.reset()

    def reset(self):
        if self.length > 0:
            self.history = []
        else:
            self.count = 0
            self.sum = 0.0
        self.val = 0.0
        self.avg = 0.0

    def reduce_update(self, tensor, num=1):
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(tensor)
        self.update(tensor.item(), num=num)

    def update(self, val, num=1):
        if self.length > 0:
            assert num == 1
            self.history.append(val)
            if len(self.history) > self.length:
                del self.history[0]
            self.val = self.history[-1]
            self.avg = sum(self.history) / len(self.history)
        else:
            self.val = val
            self.sum += val * num
            self.count += num
            self.avg = self.sum / self.count


class _PendingRead(object):
    __slots__ = ("meters", "host", "event", "nums")

    def __init__(self, meters, host, event, nums):
        self.meters, self.host, self.event, self.nums = meters, host, event, nums


_pending = []


def reduce_update_many(meters, tensors, nums=None, defer=False):
    """meters[i].reduce_update(tensors[i], nums[i]) for all i with ONE all-reduce of the stacked scalars.

    defer=False: one host read (instead of len(meters)), values visible on return -- a drop-in for the three calls at
    train_solver.py:447-449. defer=True (CUDA tensors): the read is an asynchronous copy into pinned memory behind an
    event; the meters are updated by the next reduce_update_many / flush_meters call, so the step never waits for the
    device (the reference only prints the meters every 50 steps, train_solver.py:452)."""
    assert len(meters) == len(tensors)
    nums = [1] * len(meters) if nums is None else list(nums)
    flush_meters(block=False)
    packed = torch.stack([t.detach().reshape(()).float() for t in tensors])
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(packed)
    if defer and packed.is_cuda:
        host = torch.empty(packed.shape, dtype=packed.dtype, pin_memory=True)
        host.copy_(packed, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        _pending.append(_PendingRead(list(meters), host, ev, nums))
        return
    for m, v, n in zip(meters, packed.tolist(), nums):
        m.update(v, num=n)


def flush_meters(block=True):
    """Apply deferred meter updates whose copy has landed (all of them when block=True), oldest first."""
    while _pending:
        p = _pending[0]
        if not block and not p.event.query():
            break
        p.event.synchronize()
        for m, v, n in zip(p.meters, p.host.tolist(), p.nums):
            m.update(v, num=n)
        _pending.pop(0)


def clip_logit_scale_step(param, before, value):
    """grad_clip.type == 'logit_scale_param' (train_solver.py:378-379,384-393): after the optimizer step the
    parameter may have moved by at most `value` from where it was before. The reference reads both values on the
    host (two `.item()` stalls per step) and rebuilds the tensor; here `before` is a device clone taken by the caller
    (`before = param.detach().clone()`) and the clamp runs on the device. Same result for every case of :388-393."""
    with torch.no_grad():
        b64 = before.detach().double()      # the reference forms before +- value in Python floats (fp64)
        param.data.copy_(torch.minimum(torch.maximum(param.data.double(), b64 - value), b64 + value))
    return param
