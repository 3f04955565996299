"""ctypes binding of libilfdt.so (include/ilfdt.h). No fallback: if the library is missing or a call
fails, a RuntimeError is raised -- nothing on the product path ever drops to PyTorch/CPU math."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libilfdt.so")

OK = 0
F32, BF16 = 0, 1
POOL_MAX, POOL_SUM, POOL_MEAN = 0, 1, 2
POOL_TYPES = {"max": POOL_MAX, "sum": POOL_SUM, "mean": POOL_MEAN}

_p = ctypes.c_void_p
_i32 = ctypes.c_int32
_i64 = ctypes.c_int64
_u64 = ctypes.c_uint64
_u32 = ctypes.c_uint32
_f = ctypes.c_float

# name -> (restype, argtypes); mirrors include/ilfdt.h one to one
SIGNATURES = {
    "ilfdt_abi_version": (_i32, []),
    "ilfdt_last_error": (ctypes.c_char_p, []),
    "ilfdt_device_check": (_i32, [_i32]),
    "ilfdt_tune_set": (_i32, [_i32, _i64]),
    "ilfdt_tune_get": (_i64, [_i32]),
    "ilfdt_cast_bf16": (_i32, [_p, _p, _i64, _p]),
    "ilfdt_cast_bf16_rows": (_i32, [_p, _p, _i64, _i64, _p, _p]),
    "ilfdt_split3_bf16": (_i32, [_p, _p, _i64, _i64, _i32, _p, _p]),
    "ilfdt_mask_rows_bf16": (_i32, [_p, _p, _i64, _i64, _p, _p]),
    "ilfdt_transpose_bf16": (_i32, [_p, _i32, _i64, _i64, _p, _p]),
    "ilfdt_ground_fwd_ws_bytes": (_i64, [_i32, _i64, _i64, _i64, _i64]),
    "ilfdt_ground_fwd": (_i32, [_p, _p, _i32, _p, _i64, _i64, _i64, _i64, _f, _f, _i32, _p, _p, _p, _i64, _p]),
    "ilfdt_text_plan_ints": (_i64, [_i64]),
    "ilfdt_ground_fwd_text_supported": (_i32, [_i64]),
    "ilfdt_text_compact": (_i32, [_p, _i32, _p, _i64, _i64, _i64, _p, _p, _p, _p]),
    "ilfdt_text_compact_x3": (_i32, [_p, _p, _i64, _i64, _i64, _p, _p, _p, _p]),
    "ilfdt_ground_fwd_text": (_i32, [_p, _p, _p, _i64, _i64, _i64, _i64, _f, _f, _p, _p, _p]),
    "ilfdt_sparsemax_fwd": (_i32, [_p, _i64, _i64, _p, _p, _p, _p, _p]),
    "ilfdt_sparsemax_bwd": (_i32, [_p, _p, _i64, _i64, _p, _p]),
    "ilfdt_gemm_nt": (_i32, [_p, _p, _i32, _i64, _i64, _i64, _f, _i32, _p, _i64, _p]),
    "ilfdt_gemm_nt_splitk_ws_bytes": (_i64, [_i64, _i64, _i64]),
    "ilfdt_gemm_nt_splitk": (_i32, [_p, _p, _i64, _i64, _i64, _f, _i32, _p, _i64, _p, _i64, _p]),
    "ilfdt_gemm_nn_supported": (_i32, [_i64, _i64, _i64]),
    "ilfdt_gemm_nn": (_i32, [_p, _p, _i64, _i64, _i64, _f, _i32, _p, _i64, _p, _i64, _p]),
    "ilfdt_gemm_tn": (_i32, [_p, _p, _i64, _i64, _i64, _f, _i32, _p, _i64, _p]),
    "ilfdt_linear_bf16": (_i32, [_p, _p, _i64, _i64, _i64, _p, _p, _i64, _p]),
    "ilfdt_gemm_tn_splitk_ws_bytes": (_i64, [_i64, _i64, _i64]),
    "ilfdt_gemm_tn_splitk": (_i32, [_p, _p, _i64, _i64, _i64, _f, _i32, _p, _i64, _p, _i64, _p]),
    "ilfdt_gemm_f32_strided": (_i32, [_p, _i64, _i64, _p, _i64, _i64, _p, _i64, _i64, _i64, _i64, _i64, _f, _i32, _p]),
    "ilfdt_ground_bwd": (_i32, [_p, _p, _p, _p, _p, _i32, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _f, _f, _p, _i32, _p, _p]),
    "ilfdt_ground_bwd_w16": (_i32, [_p, _p, _p, _p, _p, _i32, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _f, _f, _p, _i32, _p, _p]),
    "ilfdt_att_fwd": (_i32, [_p, _i64, _i64, _i32, _p, _p, _p]),
    "ilfdt_att_bwd": (_i32, [_p, _p, _p, _i64, _i64, _i32, _p, _p]),
    "ilfdt_rownorm_fwd": (_i32, [_p, _p, _i64, _i64, _p, _p]),
    "ilfdt_rownorm_bwd": (_i32, [_p, _p, _p, _i64, _i64, _p, _p, _p]),
    "ilfdt_token_sum": (_i32, [_p, _p, _i64, _i64, _i64, _p, _p]),
    "ilfdt_token_bcast": (_i32, [_p, _p, _i64, _i64, _i64, _f, _p, _p]),
    "ilfdt_l2norm_fwd": (_i32, [_p, _i64, _i64, _f, _p, _p, _p, _p]),
    "ilfdt_l2norm_bwd": (_i32, [_p, _p, _p, _i64, _i64, _f, _p, _p]),
    "ilfdt_logits": (_i32, [_p, _p, _i32, _i64, _i64, _i64, _p, _p, _p]),
    "ilfdt_nce_ws_bytes": (_i64, [_i32, _i64, _i64, _i64]),
    "ilfdt_nce_fwd": (_i32, [_p, _p, _p, _p, _i32, _i64, _i64, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _i64, _p]),
    "ilfdt_nce_lse": (_i32, [_p, _p, _i32, _i64, _i64, _i64, _p, _i64, _p, _p, _p, _i64, _p]),
    "ilfdt_nce_loss": (_i32, [_p, _p, _p, _p, _i64, _p, _p]),
    "ilfdt_nce_bwd": (_i32, [_p, _p, _p, _p, _i32, _i64, _i64, _i64, _p, _i64, _p, _p, _p,
                             _p, _p, _p, _p, _p, _p, _i64, _p]),
    "ilfdt_nce_bwd_gathered": (_i32, [_p, _p, _p, _p, _i32, _i64, _i64, _i64, _p, _i64, _p, _p, _p,
                                      _p, _p, _p, _p, _i64, _p]),
    "ilfdt_ce_rows_fwd": (_i32, [_p, _i64, _i64, _i64, _p, _p, _p]),
    "ilfdt_ce_rows_bwd": (_i32, [_p, _i64, _i64, _i64, _p, _p, _f, _p, _p]),
    "ilfdt_topk_hits": (_i32, [_p, _i64, _i64, _i64, _p, _p]),
    "ilfdt_nce_topk_hits": (_i32, [_p, _p, _i32, _i64, _i64, _i64, _p, _i64, _p, _p, _p, _i64, _p]),
    "ilfdt_ln_fwd": (_i32, [_p, _i32, _p, _p, _p, _i64, _i64, _f, _i32, _p, _p, _p, _p, _p]),
    "ilfdt_ln_bwd_ws_bytes": (_i64, [_i64, _i64]),
    "ilfdt_ln_bwd": (_i32, [_p, _i32, _p, _i32, _p, _p, _p, _i64, _i64, _i32, _p, _i32, _p, _p, _p, _p, _i64, _p]),
    "ilfdt_bias_cast": (_i32, [_p, _p, _i64, _i64, _p, _i32, _p]),
    "ilfdt_colsum_ws_bytes": (_i64, [_i64, _i64]),
    "ilfdt_colsum": (_i32, [_p, _i32, _i64, _i64, _p, _p, _i64, _p]),
    "ilfdt_colsum_static": (_i32, [_p, _i32, _i64, _i64, _p, _p]),
    "ilfdt_randn_fill": (_i32, [_p, _i64, _u64, _u32, _p]),
    "ilfdt_copy": (_i32, [_p, _p, _i64, _p]),
}

TUNE = {"dsd_split_rows": 0, "dsd_tile": 1, "dsd_stream": 2, "ground_2cta": 3, "gemm_wide2": 4, "nce_2cta": 5,
        "sparse_bwd_nnz": 6, "dsd_pipe": 7, "dq_pipe": 8, "f32_x3": 9, "dq_deep": 10, "text_merged_max_b": 11}

_lib = None


def load():
    """Loads libilfdt.so (once). Raises if it has not been built: there is no other code path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libilfdt.so not found at %s -- build it with `python iterated-learning-for-vlm_b200/build.py` "
            "(or __graft_entry__.build()); ilfdt has no PyTorch/CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.ilfdt_abi_version() != 4:
        raise RuntimeError("libilfdt ABI version mismatch")
    _lib = lib
    return lib


class IlfdtError(RuntimeError):
    pass


def check(rc):
    if rc != OK:
        msg = load().ilfdt_last_error()
        raise IlfdtError("libilfdt call failed (%d): %s" % (rc, msg.decode() if msg else "?"))
