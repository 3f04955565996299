"""Tensor-level wrappers over the C ABI: allocate outputs with torch (device memory + stream are the
only things PyTorch provides here), pass raw pointers, raise on error. No arithmetic happens in
Python and nothing falls back to torch ops."""
import math

import torch

from . import _lib
from ._lib import BF16, F32, POOL_TYPES, check

_DT = {torch.float32: F32, torch.bfloat16: BF16}


def _stream():
    return torch.cuda.current_stream().cuda_stream


# ---- launch accounting / per-op device timing (bench.py) -------------------------------------------
LAUNCHES = 0          # kernels launched through this module since it was last zeroed
PROFILE_ON = False    # when True every C call is bracketed by CUDA events on the launching stream
PROFILE = {}          # op name -> [(start_event, end_event), ...]
_KERNELS_PER_CALL = {"ilfdt_cast_bf16": 1, "ilfdt_split3_bf16": 1, "ilfdt_cast_bf16_rows": 1, "ilfdt_mask_rows_bf16": 1, "ilfdt_transpose_bf16": 1, "ilfdt_ground_fwd": 1, "ilfdt_ground_fwd_text": 5, "ilfdt_text_compact": 3, "ilfdt_text_compact_x3": 3, "ilfdt_sparsemax_fwd": 1,
                     "ilfdt_sparsemax_bwd": 1, "ilfdt_gemm_nt": 1, "ilfdt_gemm_nt_splitk": 2, "ilfdt_gemm_nn": 2, "ilfdt_gemm_f32_strided": 1, "ilfdt_ground_bwd": 2, "ilfdt_ground_bwd_w16": 2,   # + 2 regime launches that return at once
                     "ilfdt_l2norm_fwd": 1, "ilfdt_l2norm_bwd": 1, "ilfdt_logits": 1, "ilfdt_nce_fwd": 5, "ilfdt_nce_lse": 2, "ilfdt_nce_loss": 1,
                     "ilfdt_nce_bwd": 4, "ilfdt_nce_bwd_gathered": 2, "ilfdt_ce_rows_fwd": 1, "ilfdt_ce_rows_bwd": 1, "ilfdt_topk_hits": 1,
                     "ilfdt_randn_fill": 1, "ilfdt_copy": 0,
                     "ilfdt_gemm_tn_splitk": 2, "ilfdt_linear_bf16": 1, "ilfdt_ln_fwd": 1, "ilfdt_ln_bwd": 4, "ilfdt_bias_cast": 1, "ilfdt_colsum": 2, "ilfdt_colsum_static": 2}


def _call(name, *args):
    global LAUNCHES
    fn = getattr(_lib.load(), name)
    if PROFILE_ON:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = fn(*args)
        e1.record()
        PROFILE.setdefault(name[len("ilfdt_"):], []).append((e0, e1))
    else:
        rc = fn(*args)
    check(rc)
    LAUNCHES += _KERNELS_PER_CALL.get(name, 1)


def profile_summary():
    """{op: (calls, total device ms)} for everything recorded while PROFILE_ON (synchronises)."""
    torch.cuda.synchronize()
    return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in PROFILE.items()}


class tune:
    """with ops.tune(dsd_split_rows=0): ... -- sets ilfdt_tune_* knobs (include/ilfdt.h) and restores them on exit."""

    def __init__(self, **kv):
        self.kv = kv

    def __enter__(self):
        lib = _lib.load()
        self.prev = {k: lib.ilfdt_tune_get(_lib.TUNE[k]) for k in self.kv}
        for k, v in self.kv.items():
            check(lib.ilfdt_tune_set(_lib.TUNE[k], int(v)))
        return self

    def __exit__(self, *a):
        lib = _lib.load()
        for k, v in self.prev.items():
            lib.ilfdt_tune_set(_lib.TUNE[k], int(v))


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _req(t, dtype=None, name="tensor"):
    if not t.is_cuda:
        raise RuntimeError("ilfdt: %s must be a CUDA tensor (there is no CPU path)" % name)
    if not t.is_contiguous():
        raise RuntimeError("ilfdt: %s must be contiguous" % name)
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError("ilfdt: %s must be %s, got %s" % (name, dtype, t.dtype))
    return t


def cast_bf16(x):
    _req(x, torch.float32, "x")
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    _call("ilfdt_cast_bf16", _ptr(x), _ptr(out), x.numel(), _stream())
    return out


def cast_bf16_rows(x, rowmask):
    """x [..., C] f32 or bf16 -> bf16 with the rows whose mask entry is non-zero (-inf = pad) zeroed."""
    if x.dtype == torch.bfloat16:
        _req(x, torch.bfloat16, "x")
        if rowmask is None:
            return x
        _req(rowmask, torch.float32, "rowmask")
        cols = x.shape[-1]
        rows = x.numel() // cols
        assert rowmask.numel() == rows
        out = torch.empty_like(x)
        _call("ilfdt_mask_rows_bf16", _ptr(x), _ptr(rowmask), rows, cols, _ptr(out), _stream())
        return out
    _req(x, torch.float32, "x")
    if rowmask is None:
        return cast_bf16(x)
    _req(rowmask, torch.float32, "rowmask")
    cols = x.shape[-1]
    rows = x.numel() // cols
    assert rowmask.numel() == rows
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    _call("ilfdt_cast_bf16_rows", _ptr(x), _ptr(rowmask), rows, cols, _ptr(out), _stream())
    return out


def split3_bf16(x, side, rowmask=None):
    """x [..., C] f32 -> bf16 [..., 6C]: the three-term bf16 split laid out for ONE fp32-grade tensor-core contraction
    (side 0 = A operand, 1 = B operand; see ilfdt_split3_bf16). Rows whose rowmask entry is non-zero are zeroed."""
    _req(x, torch.float32, "x")
    cols = x.shape[-1]
    rows = x.numel() // cols
    if rowmask is not None:
        _req(rowmask, torch.float32, "rowmask")
        assert rowmask.numel() == rows
    out = torch.empty(tuple(x.shape[:-1]) + (6 * cols,), dtype=torch.bfloat16, device=x.device)
    _call("ilfdt_split3_bf16", _ptr(x), _ptr(rowmask), rows, cols, int(side), _ptr(out), _stream())
    return out


def transpose_bf16(x):
    """x [R, C] (f32 or bf16) -> bf16 [C, R]"""
    _req(x, None, "x")
    assert x.dim() == 2 and x.dtype in _DT
    out = torch.empty(x.shape[1], x.shape[0], dtype=torch.bfloat16, device=x.device)
    _call("ilfdt_transpose_bf16", _ptr(x), _DT[x.dtype], x.shape[0], x.shape[1], _ptr(out), _stream())
    return out


def ground_fwd(q, sd, mask, temperature, pool_type="max", sqrt_d=None):
    """q [b,N,D], sd [K,D] (both f32 or both bf16), mask [b,N] f32 (0 / -inf) or None
    -> pooled [b,K] f32, arg [b,K] int16 (max pooling) or None. sqrt_d: the score scale when D is not the model width
    (the 6x-wide operands of the 3 x bf16 contraction)."""
    lib = _lib.load()
    _req(q, None, "q"); _req(sd, q.dtype, "sd")
    b, n_tok, D = q.shape
    K = sd.shape[0]
    assert sd.shape[1] == D
    if mask is not None:
        _req(mask, torch.float32, "mask")
        assert tuple(mask.shape) == (b, n_tok)
    dt = _DT[q.dtype]
    pt = POOL_TYPES[pool_type]
    pooled = torch.empty(b, K, dtype=torch.float32, device=q.device)
    arg = torch.empty(b, K, dtype=torch.int16, device=q.device) if pool_type == "max" else None
    ws_bytes = lib.ilfdt_ground_fwd_ws_bytes(dt, b, n_tok, K, D) if pool_type == "max" else 0
    ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device=q.device) if ws_bytes else None
    _call("ilfdt_ground_fwd", _ptr(q), _ptr(sd), dt, _ptr(mask), b, n_tok, K, D,
          float(math.sqrt(D) if sqrt_d is None else sqrt_d), float(temperature), pt, _ptr(pooled), _ptr(arg), _ptr(ws),
          ws_bytes, _stream())
    return pooled, arg


class TextPlan:
    """Device-side plan of ilfdt_text_compact: buckets, slot_of / row_base / sample_of (all int32 views of one buffer)."""

    def __init__(self, buf, b):
        self.buf, self.b = buf, b
        self.buckets = buf[:64].view(16, 4)          # {samples, first row, first slot, rows per sample}
        self.slot_of = buf[64:64 + b]
        self.row_base = buf[64 + b:64 + 2 * b]
        self.sample_of = buf[64 + 2 * b:64 + 3 * b]


def text_supported(n_tok):
    return bool(_lib.load().ilfdt_ground_fwd_text_supported(int(n_tok)))


def text_compact(q, mask):
    """q [b,N,D] f32 or bf16, mask [b,N] f32 (0 valid / non-zero pad) -> (compacted bf16 operand [b*N, D] of which only
    the planned rows are written, TextPlan). Valid rows + one pad per sample survive, bucketed by 16-row groups."""
    lib = _lib.load()
    _req(q, None, "q"); _req(mask, torch.float32, "mask")
    assert q.dtype in _DT
    b, n_tok, D = q.shape
    assert tuple(mask.shape) == (b, n_tok)
    plan = torch.empty(lib.ilfdt_text_plan_ints(b), dtype=torch.int32, device=q.device)
    groups = torch.empty(max(b, 1), dtype=torch.uint8, device=q.device)
    out = torch.empty(b * n_tok, D, dtype=torch.bfloat16, device=q.device)
    _call("ilfdt_text_compact", _ptr(q), _DT[q.dtype], _ptr(mask), b, n_tok, D, _ptr(groups), _ptr(plan), _ptr(out),
          _stream())
    return out, TextPlan(plan, b)


def text_compact_x3(q, mask):
    """q [b,N,D] f32, mask [b,N] -> (compacted side-1 split3 operand [b*N, 6D] bf16 of which only the planned rows are
    written, TextPlan): text_compact for the fp32-grade 3 x bf16 score contraction."""
    lib = _lib.load()
    _req(q, torch.float32, "q"); _req(mask, torch.float32, "mask")
    b, n_tok, D = q.shape
    assert tuple(mask.shape) == (b, n_tok)
    plan = torch.empty(lib.ilfdt_text_plan_ints(b), dtype=torch.int32, device=q.device)
    groups = torch.empty(max(b, 1), dtype=torch.uint8, device=q.device)
    out = torch.empty(b * n_tok, 6 * D, dtype=torch.bfloat16, device=q.device)
    _call("ilfdt_text_compact_x3", _ptr(q), _ptr(mask), b, n_tok, D, _ptr(groups), _ptr(plan), _ptr(out), _stream())
    return out, TextPlan(plan, b)


def ground_fwd_text(q_compact, sd, plan, n_tok, temperature, sqrt_d=None):
    """compacted bf16 operand + plan, bf16 sd [K,D] -> pooled [b,K] f32, arg [b,K] int16 (max pooling). sqrt_d: the score
    scale when D is not the model width (the 6x-wide operands of the 3 x bf16 contraction)."""
    _req(q_compact, torch.bfloat16, "q_compact"); _req(sd, torch.bfloat16, "sd")
    b = plan.b
    K, D = sd.shape
    pooled = torch.empty(b, K, dtype=torch.float32, device=sd.device)
    arg = torch.empty(b, K, dtype=torch.int16, device=sd.device)
    _call("ilfdt_ground_fwd_text", _ptr(q_compact), _ptr(sd), _ptr(plan.buf), b, n_tok, K, D,
          float(math.sqrt(D) if sqrt_d is None else sqrt_d), float(temperature), _ptr(pooled), _ptr(arg), _stream())
    lib = _lib.load()
    global LAUNCHES
    if lib.ilfdt_tune_get(_lib.TUNE["ground_2cta"]) and b <= lib.ilfdt_tune_get(_lib.TUNE["text_merged_max_b"]) \
            and D <= 512 and n_tok in (16, 32, 48, 64, 77):
        LAUNCHES -= 4          # small batches: the tiles of all length buckets in ONE launch (ground2_text_kernel)
    else:
        LAUNCHES += -5 + (n_tok + 15) // 16
    return pooled, arg


def sparsemax_fwd(z, want_bf16=False, want_f32=True):
    """-> (w f32 | None, w bf16 | None, tau, nnz); want_f32=False (needs want_bf16) keeps only the bf16 copy."""
    _req(z, torch.float32, "z")
    rows, K = z.shape
    if not (want_f32 or want_bf16):
        raise ValueError("ilfdt: sparsemax_fwd needs at least one output")
    w = torch.empty_like(z) if want_f32 else None
    wb = torch.empty(rows, K, dtype=torch.bfloat16, device=z.device) if want_bf16 else None
    tau = torch.empty(rows, dtype=torch.float32, device=z.device)
    nnz = torch.empty(rows, dtype=torch.int32, device=z.device)
    _call("ilfdt_sparsemax_fwd", _ptr(z), rows, K, _ptr(w), _ptr(wb), _ptr(tau), _ptr(nnz), _stream())
    return w, wb, tau, nnz


def sparsemax_bwd(w, g):
    _req(w, torch.float32, "w"); _req(g, torch.float32, "g")
    dz = torch.empty_like(w)
    _call("ilfdt_sparsemax_bwd", _ptr(w), _ptr(g), w.shape[0], w.shape[1], _ptr(dz), _stream())
    return dz


def gemm_nt(A, B, alpha=1.0, out=None, accumulate=False):
    """out[N, M] (+)= alpha * B[N,Kr] . A[M,Kr]^T   (both operands reduction-contiguous; f32 or bf16)."""
    _req(A, None, "A"); _req(B, A.dtype, "B")
    M, Kr = A.shape
    N = B.shape[0]
    assert B.shape[1] == Kr
    if out is None:
        assert not accumulate
        out = torch.empty(N, M, dtype=torch.float32, device=A.device)
    _req(out, torch.float32, "out")
    if A.dtype == torch.bfloat16:
        nws = _lib.load().ilfdt_gemm_nt_splitk_ws_bytes(M, N, Kr)   # > 0: few rows, long reduction (pooling at small batch)
        if nws > 0:
            ws = torch.empty(nws, dtype=torch.uint8, device=A.device)
            _call("ilfdt_gemm_nt_splitk", _ptr(A), _ptr(B), M, N, Kr, float(alpha), int(accumulate), _ptr(out), out.shape[1],
                  _ptr(ws), nws, _stream())
            return out
    _call("ilfdt_gemm_nt", _ptr(A), _ptr(B), _DT[A.dtype], M, N, Kr, float(alpha), int(accumulate),
                                    _ptr(out), out.shape[1], _stream())
    return out


def gemm_nn_supported(M, N, Kr):
    return _lib.load().ilfdt_gemm_nn_supported(int(M), int(N), int(Kr)) != 0


def gemm_nn(A, B, alpha=1.0, out=None, accumulate=False):
    """out[N, M] (+)= alpha * B[N,Kr] . A[Kr,M]   (bf16; A with the reduction index OUTERMOST -- the codebook sd [K, D] as it
    is, no transposed copy: the pooling of clip_fdt.py:154). Shapes: gemm_nn_supported(M, N, Kr)."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B")
    Kr, M = A.shape
    N = B.shape[0]
    assert B.shape[1] == Kr
    if out is None:
        assert not accumulate
        out = torch.empty(N, M, dtype=torch.float32, device=A.device)
    _req(out, torch.float32, "out")
    nws = _lib.load().ilfdt_gemm_nt_splitk_ws_bytes(M, N, Kr)
    ws = torch.empty(nws, dtype=torch.uint8, device=A.device) if nws > 0 else None
    _call("ilfdt_gemm_nn", _ptr(A), _ptr(B), M, N, Kr, float(alpha), int(accumulate), _ptr(out), out.shape[1], _ptr(ws), nws,
          _stream())
    if nws == 0:
        global LAUNCHES
        LAUNCHES -= 1          # no slab reduction behind the unsplit launch
    return out


def gemm_tn(A, B, alpha=1.0, out=None, accumulate=False):
    """out[N, M] (+)= alpha * B[Kr,N]^T . A[Kr,M]   (bf16 operands with the reduction index outermost)."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B")
    Kr, M = A.shape
    N = B.shape[1]
    assert B.shape[0] == Kr
    if out is None:
        assert not accumulate
        out = torch.empty(N, M, dtype=torch.float32, device=A.device)
    _req(out, torch.float32, "out")
    _call("ilfdt_gemm_tn", _ptr(A), _ptr(B), M, N, Kr, float(alpha), int(accumulate), _ptr(out), out.shape[1], _stream())
    return out


def linear_bf16(W, X, bias=None):
    """nn.Linear on bf16 operands: X [R,Kr] . W [M,Kr]^T + bias [M] -> bf16 [R,M] (bias add + rounding in the GEMM's epilogue)."""
    _req(W, torch.bfloat16, "W"); _req(X, torch.bfloat16, "X")
    M, Kr = W.shape
    R = X.shape[0]
    assert X.shape[1] == Kr
    if bias is not None:
        _req(bias, torch.float32, "bias")
    out = torch.empty(R, M, dtype=torch.bfloat16, device=X.device)
    _call("ilfdt_linear_bf16", _ptr(W), _ptr(X), M, R, Kr, _ptr(bias), _ptr(out), M, _stream())
    return out


def gemm_tn_splitk(A, B, alpha=1.0, out=None, accumulate=False):
    """gemm_tn with the reduction split over the SMs (weight-gradient shapes); same contract."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B")
    Kr, M = A.shape
    N = B.shape[1]
    assert B.shape[0] == Kr
    if out is None:
        assert not accumulate
        out = torch.empty(N, M, dtype=torch.float32, device=A.device)
    _req(out, torch.float32, "out")
    nws = _lib.load().ilfdt_gemm_tn_splitk_ws_bytes(M, N, Kr)
    ws = torch.empty(max(nws, 4), dtype=torch.uint8, device=A.device)
    _call("ilfdt_gemm_tn_splitk", _ptr(A), _ptr(B), M, N, Kr, float(alpha), int(accumulate), _ptr(out), out.shape[1],
          _ptr(ws), nws, _stream())
    return out


def gemm_f32(A, B, a_t=False, b_t=False, alpha=1.0, out=None, accumulate=False):
    """fp32 C = alpha * op(A) @ op(B) for 2-D row-major tensors (op = transpose if *_t)."""
    _req(A, torch.float32, "A"); _req(B, torch.float32, "B")
    M, Kr = (A.shape[1], A.shape[0]) if a_t else A.shape
    Kb, N = (B.shape[1], B.shape[0]) if b_t else B.shape
    assert Kr == Kb
    a_rs, a_cs = (1, A.shape[1]) if a_t else (A.shape[1], 1)      # A(m,k)
    b_rs, b_cs = (B.shape[1], 1) if b_t else (1, B.shape[1])      # B(n,k) = op(B)[k,n]
    if out is None:
        assert not accumulate
        out = torch.empty(M, N, dtype=torch.float32, device=A.device)
    _req(out, torch.float32, "out")
    _call("ilfdt_gemm_f32_strided", _ptr(A), a_rs, a_cs, _ptr(B), b_rs, b_cs, _ptr(out), N, 1, M, N, Kr,
                                             float(alpha), int(accumulate), _stream())
    return out


def ground_bwd(g, w, arg, q, sd, mask, temperature, dsd, dq_dtype=torch.float32, row_base=None, n_tok=None, nnz=None):
    """g [b,K] is scratch (overwritten). Returns dq [b,N,D] (dq_dtype); accumulates into dsd [K,D] f32.
    row_base (int32 [b]) + n_tok: q is the compacted text operand [rows, D] of text_compact.
    nnz (int32 [b], from sparsemax_fwd): lets the library serve a launch whose mean support is a few tens of codes with
    the sparse-regime kernel (decided on the device)."""
    _req(g, torch.float32, "g"); _req(arg, torch.int16, "arg")
    fn = "ilfdt_ground_bwd"
    if w is not None and w.dtype == torch.bfloat16:   # support read from the bf16 copy of the weights
        _req(w, torch.bfloat16, "w")
        fn = "ilfdt_ground_bwd_w16"
    elif w is not None:   # None: g already is dLoss/dpooled (softmax / sigmoid attention)
        _req(w, torch.float32, "w")
    _req(q, None, "q"); _req(sd, q.dtype, "sd"); _req(dsd, torch.float32, "dsd")
    if row_base is None:
        b, n_tok, D = q.shape
    else:
        _req(row_base, torch.int32, "row_base")
        b, D = g.shape[0], q.shape[-1]
    K = sd.shape[0]
    dq = torch.empty(b, n_tok, D, dtype=dq_dtype, device=q.device)
    regime = None
    if nnz is not None:
        _req(nnz, torch.int32, "nnz")
        regime = torch.empty(1, dtype=torch.int32, device=q.device)
    _call(fn, _ptr(g), _ptr(w), _ptr(arg), _ptr(q), _ptr(sd), _DT[q.dtype], _ptr(mask), _ptr(row_base), _ptr(nnz),
          _ptr(regime), b, n_tok, K, D, float(math.sqrt(D)), float(temperature), _ptr(dq), _DT[dq_dtype], _ptr(dsd), _stream())
    return dq


ATT_KINDS = {"softmax": 1, "sigmoid": 2}


def att_fwd(z, kind):
    """softmax / sigmoid over the last dim of z [rows,K] -> (w, rowsum [rows])"""
    _req(z, torch.float32, "z")
    rows, K = z.shape
    w = torch.empty_like(z)
    rowsum = torch.empty(rows, dtype=torch.float32, device=z.device)
    _call("ilfdt_att_fwd", _ptr(z), rows, K, ATT_KINDS[kind], _ptr(w), _ptr(rowsum), _stream())
    return w, rowsum


def att_bwd(w, g, kind, rowbias=None):
    _req(w, torch.float32, "w"); _req(g, torch.float32, "g")
    dz = torch.empty_like(w)
    _call("ilfdt_att_bwd", _ptr(w), _ptr(g), _ptr(rowbias), w.shape[0], w.shape[1], ATT_KINDS[kind], _ptr(dz),
          _stream())
    return dz


def rownorm_fwd(u, s):
    _req(u, torch.float32, "u"); _req(s, torch.float32, "s")
    f = torch.empty_like(u)
    _call("ilfdt_rownorm_fwd", _ptr(u), _ptr(s), u.shape[0], u.shape[1], _ptr(f), _stream())
    return f


def rownorm_bwd(df, f, s):
    _req(df, torch.float32, "df"); _req(f, torch.float32, "f")
    du = torch.empty_like(df)
    ds = torch.empty_like(s)
    _call("ilfdt_rownorm_bwd", _ptr(df), _ptr(f), _ptr(s), df.shape[0], df.shape[1], _ptr(du), _ptr(ds), _stream())
    return du, ds


def token_sum(q, mask):
    _req(q, torch.float32, "q")
    b, n_tok, D = q.shape
    out = torch.empty(b, D, dtype=torch.float32, device=q.device)
    _call("ilfdt_token_sum", _ptr(q), _ptr(mask), b, n_tok, D, _ptr(out), _stream())
    return out


def token_bcast(v, mask, n_tok, scale):
    _req(v, torch.float32, "v")
    b, D = v.shape
    out = torch.empty(b, n_tok, D, dtype=torch.float32, device=v.device)
    _call("ilfdt_token_bcast", _ptr(v), _ptr(mask), b, n_tok, D, float(scale), _ptr(out), _stream())
    return out


def l2norm_fwd(f, eps, want_bf16=False):
    _req(f, torch.float32, "f")
    rows, D = f.shape
    x = torch.empty_like(f)
    xb = torch.empty(rows, D, dtype=torch.bfloat16, device=f.device) if want_bf16 else None
    norm = torch.empty(rows, dtype=torch.float32, device=f.device)
    _call("ilfdt_l2norm_fwd", _ptr(f), rows, D, float(eps), _ptr(x), _ptr(xb), _ptr(norm), _stream())
    return x, xb, norm


def l2norm_bwd(x, norm, g, eps):
    _req(x, torch.float32, "x"); _req(g, torch.float32, "g")
    df = torch.empty_like(x)
    _call("ilfdt_l2norm_bwd", _ptr(x), _ptr(norm), _ptr(g), x.shape[0], x.shape[1], float(eps), _ptr(df),
                                       _stream())
    return df


def logits(x, g, scale):
    """x [b,D], g [B,D] (f32 or bf16), scale: f32 device tensor with 1 element -> [b,B] f32"""
    _req(x, None, "x"); _req(g, x.dtype, "g"); _req(scale, torch.float32, "scale")
    b, D = x.shape
    B = g.shape[0]
    out = torch.empty(b, B, dtype=torch.float32, device=x.device)
    _call("ilfdt_logits", _ptr(x), _ptr(g), _DT[x.dtype], b, B, D, _ptr(scale), _ptr(out), _stream())
    return out


def _nce_ws(dt, b, B, D, device):
    n = _lib.load().ilfdt_nce_ws_bytes(dt, b, B, D)
    return torch.empty(max(n, 1), dtype=torch.uint8, device=device), n


def nce_fwd(x_img, x_txt, g_img, g_txt, scale, label_offset):
    for t in (x_img, x_txt, g_img, g_txt):
        _req(t, x_img.dtype, "embedding")
    _req(scale, torch.float32, "scale")
    b, D = x_img.shape
    B = g_img.shape[0]
    dev = x_img.device
    dt = _DT[x_img.dtype]
    lse_i, lse_t, diag_i, diag_t = (torch.empty(b, dtype=torch.float32, device=dev) for _ in range(4))
    loss = torch.empty((), dtype=torch.float32, device=dev)
    ws, n = _nce_ws(dt, b, B, D, dev)
    _call("ilfdt_nce_fwd", _ptr(x_img), _ptr(x_txt), _ptr(g_img), _ptr(g_txt), dt, b, B, D, _ptr(scale),
                                    int(label_offset), _ptr(lse_i), _ptr(lse_t), _ptr(diag_i), _ptr(diag_t),
                                    _ptr(loss), _ptr(ws), n, _stream())
    return loss, lse_i, lse_t, diag_i, diag_t


def nce_lse(x, g, scale, label_offset):
    """One direction of nce_fwd: (lse [b], diag [b]) of the rows scale * x . g^T, labels label_offset + row."""
    _req(x, None, "x"); _req(g, x.dtype, "g"); _req(scale, torch.float32, "scale")
    b, D = x.shape
    B = g.shape[0]
    dt = _DT[x.dtype]
    lse = torch.empty(b, dtype=torch.float32, device=x.device)
    diag = torch.empty_like(lse)
    ws, n = _nce_ws(dt, b, B, D, x.device)
    _call("ilfdt_nce_lse", _ptr(x), _ptr(g), dt, b, B, D, _ptr(scale), int(label_offset), _ptr(lse), _ptr(diag), _ptr(ws), n,
          _stream())
    return lse, diag


def nce_loss(lse_i, diag_i, lse_t, diag_t):
    for t in (lse_i, diag_i, lse_t, diag_t):
        _req(t, torch.float32, "lse / diag")
    loss = torch.empty((), dtype=torch.float32, device=lse_i.device)
    _call("ilfdt_nce_loss", _ptr(lse_i), _ptr(diag_i), _ptr(lse_t), _ptr(diag_t), lse_i.numel(), _ptr(loss), _stream())
    return loss


def nce_bwd(x_img, x_txt, g_img, g_txt, scale, label_offset, lse_i, lse_t, grad_out):
    """Returns dx_img, dx_txt [b,D], dg_img, dg_txt [B,D], dscale [] (all f32)."""
    b, D = x_img.shape
    B = g_img.shape[0]
    dev = x_img.device
    dt = _DT[x_img.dtype]
    _req(grad_out, torch.float32, "grad_out")
    dx_img = torch.empty(b, D, dtype=torch.float32, device=dev)
    dx_txt = torch.empty_like(dx_img)
    dg_img = torch.empty(B, D, dtype=torch.float32, device=dev)
    dg_txt = torch.empty_like(dg_img)
    dscale = torch.empty((), dtype=torch.float32, device=dev)
    ws, n = _nce_ws(dt, b, B, D, dev)
    _call("ilfdt_nce_bwd", _ptr(x_img), _ptr(x_txt), _ptr(g_img), _ptr(g_txt), dt, b, B, D, _ptr(scale),
          int(label_offset), _ptr(lse_i), _ptr(lse_t), _ptr(grad_out), _ptr(dx_img), _ptr(dx_txt),
          _ptr(dg_img), _ptr(dg_txt), _ptr(dscale), _ptr(ws), n, _stream())
    return dx_img, dx_txt, dg_img, dg_txt, dscale


def nce_bwd_gathered(x_img, x_txt, g_img, g_txt, scale, label_offset, lse_i_all, lse_t_all, grad_out_all):
    """g_* = all-gather of x_*; lse_*_all [B], grad_out_all [B/b] gathered over ranks.
    Returns the local rows' total gradient dx_img, dx_txt [b,D] and the rank-local dscale []."""
    b, D = x_img.shape
    B = g_img.shape[0]
    dev = x_img.device
    dt = _DT[x_img.dtype]
    for t, nm, cnt in ((lse_i_all, "lse_i_all", B), (lse_t_all, "lse_t_all", B), (grad_out_all, "grad_out_all", B // b)):
        _req(t, torch.float32, nm)
        assert t.numel() == cnt, nm
    dx_img = torch.empty(b, D, dtype=torch.float32, device=dev)
    dx_txt = torch.empty_like(dx_img)
    dscale = torch.empty((), dtype=torch.float32, device=dev)
    ws, n = _nce_ws(dt, b, B, D, dev)
    _call("ilfdt_nce_bwd_gathered", _ptr(x_img), _ptr(x_txt), _ptr(g_img), _ptr(g_txt), dt, b, B, D, _ptr(scale),
          int(label_offset), _ptr(lse_i_all), _ptr(lse_t_all), _ptr(grad_out_all), _ptr(dx_img), _ptr(dx_txt),
          _ptr(dscale), _ptr(ws), n, _stream())
    return dx_img, dx_txt, dscale


def ce_rows_fwd(logits_f32, label_offset):
    _req(logits_f32, torch.float32, "logits")
    b, B = logits_f32.shape
    lse = torch.empty(b, dtype=torch.float32, device=logits_f32.device)
    diag = torch.empty_like(lse)
    _call("ilfdt_ce_rows_fwd", _ptr(logits_f32), b, B, int(label_offset), _ptr(lse), _ptr(diag), _stream())
    return lse, diag


def ce_rows_bwd(logits_f32, label_offset, lse, grad_out, coef):
    b, B = logits_f32.shape
    out = torch.empty_like(logits_f32)
    _call("ilfdt_ce_rows_bwd", _ptr(logits_f32), b, B, int(label_offset), _ptr(lse), _ptr(grad_out),
                                        float(coef), _ptr(out), _stream())
    return out


def topk_hits(logits_f32, label_offset):
    _req(logits_f32, torch.float32, "logits")
    counts = torch.empty(2, dtype=torch.int32, device=logits_f32.device)
    _call("ilfdt_topk_hits", _ptr(logits_f32), logits_f32.shape[0], logits_f32.shape[1], int(label_offset),
                                      _ptr(counts), _stream())
    return counts


def nce_topk_hits(x, g, scale, label_offset, diag):
    """(top-1 hits, top-5 hits) int32[2] of logits = scale * x . g^T against labels label_offset + row, without the
    [b,B] matrix (prototype/utils/misc.py:464-477 as called at train_solver.py:421).
    diag = the rows' label logits (nce_fwd's diag_i for x = x_img, g = g_txt)."""
    _req(x, None, "x"); _req(g, x.dtype, "g"); _req(scale, torch.float32, "scale"); _req(diag, torch.float32, "diag")
    b, D = x.shape
    B = g.shape[0]
    dt = _DT[x.dtype]
    counts = torch.empty(2, dtype=torch.int32, device=x.device)
    ws, n = _nce_ws(dt, b, B, D, x.device)
    _call("ilfdt_nce_topk_hits", _ptr(x), _ptr(g), dt, b, B, D, _ptr(scale), int(label_offset), _ptr(diag), _ptr(counts),
          _ptr(ws), n, _stream())
    return counts


def randn_fill_(t, seed, generation):
    _req(t, torch.float32, "t")
    _call("ilfdt_randn_fill", _ptr(t), t.numel(), int(seed) & (2 ** 64 - 1), int(generation) & 0xFFFFFFFF,
                                       _stream())
    return t


def copy_(dst, src):
    _req(dst, None, "dst"); _req(src, dst.dtype, "src")
    assert dst.numel() == src.numel()
    _call("ilfdt_copy", _ptr(dst), _ptr(src), dst.numel() * dst.element_size(), _stream())
    return dst


# ---- q_map producer rows (clip_fdt.py:86-92) -----------------------------------------
def ln_fwd(src, gamma, beta, eps, bias=None, gelu=None, want_pre=None):
    """y = LayerNorm(act(src + bias)) as bf16 [R,C] with mean, rstd [R]. gelu (default: bias is given): act = exact GELU;
    src is then either the bf16 pre-activation itself (no bias) or the fp32 product without its bias, in which case the
    bf16 pre-activation src + bias is returned as well."""
    _req(src, None, "src"); _req(gamma, torch.float32, "gamma"); _req(beta, torch.float32, "beta")
    gelu = (bias is not None) if gelu is None else gelu
    want_pre = (bias is not None) if want_pre is None else want_pre
    R, C = src.shape
    y = torch.empty(R, C, dtype=torch.bfloat16, device=src.device)
    mean = torch.empty(R, dtype=torch.float32, device=src.device)
    rstd = torch.empty_like(mean)
    pre = None
    if bias is not None:
        _req(bias, torch.float32, "bias")
    if want_pre:
        pre = torch.empty_like(y)
    _call("ilfdt_ln_fwd", _ptr(src), _DT[src.dtype], _ptr(bias), _ptr(gamma), _ptr(beta), R, C, float(eps),
          int(gelu), _ptr(y), _ptr(pre), _ptr(mean), _ptr(rstd), _stream())
    return y, pre, mean, rstd


def ln_bwd(dy, src, gamma, mean, rstd, gelu):
    """-> dsrc (bf16 through the GELU, else f32), dgamma, dbeta, dbias (GELU form only, else None)."""
    _req(dy, None, "dy"); _req(src, None, "src"); _req(gamma, torch.float32, "gamma")
    R, C = dy.shape
    assert src.shape == dy.shape
    dsrc = torch.empty(R, C, dtype=torch.bfloat16 if gelu else torch.float32, device=dy.device)
    dgamma = torch.empty(C, dtype=torch.float32, device=dy.device)
    dbeta = torch.empty_like(dgamma)
    dbias = torch.empty_like(dgamma) if gelu else None
    nws = _lib.load().ilfdt_ln_bwd_ws_bytes(R, C)
    ws = torch.empty(nws, dtype=torch.uint8, device=dy.device)
    _call("ilfdt_ln_bwd", _ptr(dy), _DT[dy.dtype], _ptr(src), _DT[src.dtype], _ptr(gamma), _ptr(mean), _ptr(rstd), R, C, int(gelu),
          _ptr(dsrc), _DT[dsrc.dtype], _ptr(dgamma), _ptr(dbeta), _ptr(dbias), _ptr(ws), nws, _stream())
    return dsrc, dgamma, dbeta, dbias


def bias_cast(src, bias, out_dtype=torch.bfloat16):
    _req(src, torch.float32, "src"); _req(bias, torch.float32, "bias")
    R, C = src.shape
    out = torch.empty(R, C, dtype=out_dtype, device=src.device)
    _call("ilfdt_bias_cast", _ptr(src), _ptr(bias), R, C, _ptr(out), _DT[out_dtype], _stream())
    return out


def colsum(src):
    _req(src, None, "src")
    R, C = src.shape
    out = torch.empty(C, dtype=torch.float32, device=src.device)
    nws = _lib.load().ilfdt_colsum_ws_bytes(R, C)
    ws = torch.empty(nws, dtype=torch.uint8, device=src.device)
    _call("ilfdt_colsum", _ptr(src), _DT[src.dtype], R, C, _ptr(out), _ptr(ws), nws, _stream())
    return out


def colsum_static(src, out):
    """ilfdt_colsum_static (include/ilfdt.h): out[c] = sum_r src[r,c] into the caller's `out`, partial sums in the library's
    static device scratch -- nothing is allocated here or in the library. Calls that overlap on the GPU race on that scratch."""
    _req(src, None, "src")
    _req(out, torch.float32, "out")
    R, C = src.shape
    if out.numel() != C:
        raise ValueError("colsum_static: out must hold %d floats" % C)
    _call("ilfdt_colsum_static", _ptr(src), _DT[src.dtype], R, C, _ptr(out), _stream())
    return out
