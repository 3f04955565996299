"""Lazy [b,B] logits (SURVEY.md 8b "fused fast path").

The reference's model.forward returns the two logit matrices and its ClipInfoCELoss reads them
(clip_fdt.py:421-428, loss.py:40-50). At global batch 32k they are 2 x b x B fp32 (8 GiB on one GPU),
written once and re-read three times. The drop-in forward returns `LazyLogits` handles instead: the
drop-in ClipInfoCELoss recognises them and runs the fused kernels on the stashed embeddings; any other
consumer (accuracy(), visualisation, indexing, torch functions) transparently materialises the matrix
with the logits kernel, with autograd intact.
"""
import dataclasses

import torch
import torch.distributed as dist

from . import functional as Fn


class ContrastiveStash:
    """What clip_fdt.py:411-419 produced: normalised local rows, gathered rows, clamped scale."""

    def __init__(self, x_img, x_txt, g_img, g_txt, scale, mode, gathered=False):
        # g_img / g_txt: [B,D] tensors, or functional.GatherHandle objects whose all-gather is still in flight
        self.x_img, self.x_txt, self.g_img, self.g_txt = x_img, x_txt, g_img, g_txt
        self.scale, self.mode = scale, mode
        # True when g_* are AllGather(x_*) of this very step (the models' forward): enables the backward that
        # needs no [B,D] gradient exchange (functional.InfoNCEFunction)
        self.gathered = gathered
        self._mat = [None, None]
        self._kept = {}

    @property
    def b(self):
        return self.x_img.shape[0]

    @property
    def B(self):
        g = self.g_img
        return (g.out if isinstance(g, Fn.GatherHandle) else g).shape[0]

    def _gathered_tensor(self, which):
        """The gathered rows as an autograd-connected tensor. With in-flight handles (the models' forward) the
        differentiable form is rebuilt with AllGather.apply, whose backward is the reference's reduce of the [B,D]
        gradient (clip_fdt.py:182-188): a materialised logits matrix back-propagates exactly as the reference's does."""
        g, x = (self.g_img, self.x_img) if which == 0 else (self.g_txt, self.x_txt)
        if not isinstance(g, Fn.GatherHandle):
            return g
        if torch.is_grad_enabled() and x.requires_grad:
            out = Fn.AllGather.apply(x)
            return out.view(-1, *out.shape[2:])
        return g.wait()

    def label_offset(self):
        """loss.py:41-45: arange(b) when the logits are square, else rank*b + arange(b)."""
        if self.B == self.b:
            return 0
        rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        return rank * self.b

    def materialize(self, which):
        if self._mat[which] is None:
            x, g = (self.x_img, self._gathered_tensor(1)) if which == 0 else (self.x_txt, self._gathered_tensor(0))
            self._mat[which] = Fn.LogitsFunction.apply(x, g, self.scale, self.mode)
        return self._mat[which]

    def fused_loss(self):
        return Fn.InfoNCEFunction.apply(self.x_img, self.x_txt, self.g_img, self.g_txt, self.scale,
                                        self.label_offset(), self.mode, self.gathered, self._kept)

    def topk_hits(self):
        """int32[2]: rows of logits_per_image whose label logit is the row max / has fewer than 5 entries above it
        (accuracy(logits_per_image, labels, topk=(1, 5)), train_solver.py:421) -- counted by the fused kernel from
        the embeddings; the [b,B] matrix is not formed."""
        from . import ops
        if self._mat[0] is not None:
            return ops.topk_hits(self._mat[0].detach().contiguous(), self.label_offset())
        if not self._kept:   # accuracy() before the loss: run the forward kernels for the label logits
            with torch.no_grad():
                Fn.InfoNCEFunction.apply(self.x_img.detach(), self.x_txt.detach(), self._gathered_tensor(0).detach(),
                                         self._gathered_tensor(1).detach(), self.scale.detach(), self.label_offset(),
                                         self.mode, False, self._kept)
        k = self._kept
        return ops.nce_topk_hits(k["x"], k["g"], k["scale"], k["label_offset"], k["diag"])


@dataclasses.dataclass(eq=False, repr=False, init=False)
class LazyLogits:
    """Stands in for logits_per_image (which=0) / logits_per_text (which=1).

    It is a dataclass on purpose: the reference wraps the model in DistributedDataParallel with
    find_unused_parameters=True (prototype/utils/torch_ddp_dist.py:65), and DDP discovers the autograd graph of a
    forward from the TENSORS IT CAN FIND in the returned object -- tuples, lists, dicts and dataclass fields
    (torch/nn/parallel/distributed.py:_find_tensors). `autograd_anchors` exposes the graph-connected tensors the
    lazy matrix is a function of (normalised local embeddings, scale); without it DDP would see no tensor at all,
    declare every parameter unused and fail on the real backward."""
    autograd_anchors: tuple = ()

    def __init__(self, stash, which):
        object.__setattr__(self, "_stash", stash)
        object.__setattr__(self, "_which", which)
        object.__setattr__(self, "autograd_anchors", tuple(
            t for t in (stash.x_img, stash.x_txt, stash.scale) if isinstance(t, torch.Tensor) and t.requires_grad))

    # -- cheap metadata that does not need the matrix --
    @property
    def shape(self):
        return torch.Size((self._stash.b, self._stash.B))

    def size(self, dim=None):
        return self.shape if dim is None else self.shape[dim]

    def dim(self):
        return 2

    def __len__(self):
        return self._stash.b

    @property
    def dtype(self):
        return torch.float32

    @property
    def device(self):
        return self._stash.x_img.device

    @property
    def is_materialized(self):
        return self._stash._mat[self._which] is not None

    def materialize(self):
        return self._stash.materialize(self._which)

    # -- everything else behaves like the real tensor --
    def __getattr__(self, name):
        return getattr(self.materialize(), name)

    def __getitem__(self, idx):
        return self.materialize()[idx]

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        def conv(a):
            if isinstance(a, LazyLogits):
                return a.materialize()
            if isinstance(a, (list, tuple)):
                return type(a)(conv(x) for x in a)
            return a

        kwargs = {k: conv(v) for k, v in (kwargs or {}).items()}
        return func(*conv(args), **kwargs)

    def _binary(self, other, op):
        other = other.materialize() if isinstance(other, LazyLogits) else other
        return op(self.materialize(), other)

    def __add__(self, o): return self._binary(o, torch.add)
    def __radd__(self, o): return self._binary(o, torch.add)
    def __sub__(self, o): return self._binary(o, torch.sub)
    def __mul__(self, o): return self._binary(o, torch.mul)
    def __rmul__(self, o): return self._binary(o, torch.mul)
    def __truediv__(self, o): return self._binary(o, torch.div)
    def __matmul__(self, o): return self._binary(o, torch.matmul)

    def __repr__(self):
        return "LazyLogits(which=%d, shape=%s, materialized=%s)" % (self._which, tuple(self.shape),
                                                                   self.is_materialized)


def contrastive_logits(x_img, x_txt, g_img, g_txt, scale, mode=None, gathered=False):
    """(logits_per_image, logits_per_text) = (scale * x_img @ g_txt.T, scale * x_txt @ g_img.T) as LazyLogits.

    The building block for heads that form several logits pairs from one set of embeddings -- DeCLIP-FDT's eight
    logits are four such pairs (two image views x masked / augmented captions, prototype/model/declip_fdt.py:389-399):
    each pair handed to ClipInfoCELoss runs the fused kernels, none of the [b,B] matrices is stored.
    gathered=True states that g_* are AllGather(x_*) of this step (see ContrastiveStash)."""
    stash = ContrastiveStash(x_img, x_txt, g_img, g_txt, scale, mode or Fn.get_compute_dtype(), gathered)
    return LazyLogits(stash, 0), LazyLogits(stash, 1)
