"""ilfdt -- B200-native (sm_100a) hot path of iterated-learning FDT-CLIP training.

Public surface mirrors the reference's module layout:
    ilfdt.model.clip_fdt   Query_model, AllGather, Clip_FDT, clip_fdt_vitb32/16   (prototype/model/clip_fdt.py)
    ilfdt.model.clip       CLIP                                                   (prototype/model/clip.py)
    ilfdt.model.sparsemax  Sparsemax                                              (prototype/model/sparsemax.py)
    ilfdt.loss_functions   ClipInfoCELoss                                         (prototype/loss_functions/loss.py)
    ilfdt.linklink         get_rank / get_world_size / barrier ...                (prototype/linklink/__init__.py)
All arithmetic of the path runs in libilfdt.so (include/ilfdt.h); there is no CPU or PyTorch fallback.
"""
from .functional import (compute_dtype, get_compute_dtype, set_compute_dtype, set_fp32_tensor_cores,  # noqa: F401
                         set_fused_q_map)
from . import _lib  # noqa: F401

__all__ = ["compute_dtype", "get_compute_dtype", "set_compute_dtype", "set_fp32_tensor_cores", "set_fused_q_map"]
