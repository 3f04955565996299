"""TEST INFRASTRUCTURE ONLY -- imports the UNMODIFIED reference from /root/reference.

This module exists to (a) validate `oracle/fdt_oracle.py` (the CPU restatement) against the live
reference and (b) generate the golden vectors under `tests/golden/` (see `oracle/make_golden.py`).
`/root/reference` does not exist on the GPU box, so nothing at run time (tests -m gpu, smoke(),
bench.py) may import this file; only the `-m "not gpu"` pinning tests do, and they skip when the
reference tree is absent.

What is neutralised, none of which carries hot-path arithmetic:
  * `easydict`, `ftfy`, `timm` are absent from this image -> stub modules in sys.modules
    (`prototype/model/clip_fdt.py:11`, `model/utils/text_utils/simple_tokenizer.py`, `model/swin/models/*`).
  * `ClipInfoCELoss` calls `.cuda()` on the labels (`prototype/loss_functions/loss.py:43,45`)
    -> `torch.Tensor.cuda` is patched to identity while the reference runs on CPU.
  * `prototype/model/sparsemax.py:6` module-global `device` -> forced to CPU.
  * `AllGather.forward` needs a process group even at W=1 (`clip_fdt.py:175`) -> gloo group.
  * encoders are replaced by fakes that return seeded synthetic features in the reference's return
    shapes (`visual_transformer.py:82-91`, `text_transformer.py:322-338`), so the unmodified
    `Clip_FDT.forward` / `CLIP.forward` is what runs.
"""
import contextlib
import os
import sys
import types

import torch
import torch.nn as nn

REFERENCE_ROOT = os.environ.get("ILFDT_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "prototype", "model"))


def _install_stubs():
    if "easydict" not in sys.modules:
        m = types.ModuleType("easydict")

        class EasyDict(dict):
            def __getattr__(self, k):
                try:
                    return self[k]
                except KeyError as e:
                    raise AttributeError(k) from e

            def __setattr__(self, k, v):
                self[k] = v

        m.EasyDict = EasyDict
        sys.modules["easydict"] = m
    if "ftfy" not in sys.modules:
        m = types.ModuleType("ftfy")
        m.fix_text = lambda s: s
        sys.modules["ftfy"] = m
    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")
        models = types.ModuleType("timm.models")
        layers = types.ModuleType("timm.models.layers")

        class DropPath(nn.Identity):
            def __init__(self, *a, **k):
                super().__init__()

        layers.DropPath = DropPath
        layers.to_2tuple = lambda x: (x, x) if not isinstance(x, tuple) else x
        layers.trunc_normal_ = lambda t, std=1.0, **k: nn.init.trunc_normal_(t, std=std)
        timm.models = models
        models.layers = layers
        sys.modules["timm"] = timm
        sys.modules["timm.models"] = models
        sys.modules["timm.models.layers"] = layers


_REF = {}


def load_reference():
    """Returns a dict with the reference's own classes (imported, not copied)."""
    if _REF:
        return _REF
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import prototype.model.sparsemax as ref_sparsemax
    import prototype.model.clip_fdt as ref_clip_fdt
    import prototype.model.clip as ref_clip
    import prototype.loss_functions.loss as ref_loss

    ref_sparsemax.device = torch.device("cpu")
    _REF.update(
        Sparsemax=ref_sparsemax.Sparsemax,
        Query_model=ref_clip_fdt.Query_model,
        Clip_FDT=ref_clip_fdt.Clip_FDT,
        AllGather=ref_clip_fdt.AllGather,
        CLIP=ref_clip.CLIP,
        ClipInfoCELoss=ref_loss.ClipInfoCELoss,
        sparsemax_module=ref_sparsemax,
    )
    return _REF


@contextlib.contextmanager
def cpu_cuda_identity():
    """`loss.py:43,45` call `.cuda()`; make that a no-op while running the reference on CPU."""
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        yield
    finally:
        torch.Tensor.cuda = orig


def ensure_process_group(rank=0, world_size=1, port=None):
    import torch.distributed as dist

    if dist.is_initialized():
        return
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(port or 29531))
    os.environ["RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world_size)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)


class FakeVisual(nn.Module):
    """Stands in for `visual_transformer_B32/B16`: returns (proj_cls, dense_feat, raw_cls)."""

    def __init__(self, patch_ft):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 1, 1)  # `Clip_FDT.dtype` reads visual.conv1.weight.dtype (clip_fdt.py:236-247)
        self.patch_ft = patch_ft

    def forward(self, image, return_dense=False, return_raw_feature=False):
        cls = self.patch_ft.mean(1)
        ret = [cls]
        if return_dense:
            ret.append(self.patch_ft)
        if return_raw_feature:
            ret.append(cls)
        return tuple(ret) if len(ret) > 1 else ret[0]


class FakeText(nn.Module):
    """Stands in for `text_transformers`: returns [proj, words, raw, pad_mask]."""

    def __init__(self, word_ft, pad_mask):
        super().__init__()
        self.text_projection = nn.Linear(1, 1)
        self.word_ft = word_ft
        self.pad_mask = pad_mask

    def forward(self, texts, return_dense=False, return_padmask=False, return_raw_feature=False, raw_text=True):
        cls = self.word_ft.mean(1)
        ret = [cls]
        if return_dense:
            ret.append(self.word_ft)
        if return_raw_feature:
            ret.append(cls)
        if return_padmask:
            ret.append(self.pad_mask)
        return ret if len(ret) > 1 else ret[0]


class FakeFeat(nn.Module):
    """Encoder stand-in for baseline CLIP: returns a fixed [b, D] feature tensor."""

    def __init__(self, feat):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 1, 1)
        self.text_projection = nn.Linear(1, 1)
        self.feat = feat

    def forward(self, x, *a, **k):
        return self.feat


def build_reference_fdt(patch_ft, word_ft, pad_mask, sd_num, sd_dim, temperature=1000.0,
                        att_func_type="sparsemax", pool_type="max"):
    ref = load_reference()
    model = ref["Clip_FDT"](
        FakeVisual(patch_ft), FakeText(word_ft, pad_mask), use_allgather=True,
        sd_num=sd_num, sd_dim=sd_dim, raw_img_ft_dim=patch_ft.shape[-1], raw_txt_ft_dim=word_ft.shape[-1],
        att_func_type=att_func_type, pool_type=pool_type, sd_temperature=temperature)
    return model
