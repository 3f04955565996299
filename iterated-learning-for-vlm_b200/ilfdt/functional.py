"""torch.autograd.Function wrappers: each forward/backward is a handful of libilfdt calls. These are
the seams between the reference-shaped nn.Modules (ilfdt.model, ilfdt.loss_functions) and the
kernels. Reference lines cited are relative to the reference repo root."""
import math
import os
import warnings

import torch
import torch.distributed as dist

from . import ops

# "bf16": tcgen05 tensor-core kernels, fp32 accumulation (<= 2e-2 rel vs the fp32 reference)
# "fp32": SIMT fp32 kernels (<= 1e-5 rel) -- the parity mode
_COMPUTE_DTYPE = os.environ.get("ILFDT_COMPUTE_DTYPE", "bf16")


def set_compute_dtype(name):
    global _COMPUTE_DTYPE
    if name not in ("bf16", "fp32"):
        raise ValueError("compute dtype must be 'bf16' or 'fp32'")
    _COMPUTE_DTYPE = name


def get_compute_dtype():
    return _COMPUTE_DTYPE


class compute_dtype:
    """Context manager: with ilfdt.compute_dtype('fp32'): ..."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        self.prev = get_compute_dtype()
        set_compute_dtype(self.name)

    def __exit__(self, *a):
        set_compute_dtype(self.prev)


# fp32 mode: run the contractions whose reduction is the model width (the score block q.sd^T and g = df.sd^T) as ONE bf16
# tensor-core contraction over a 3-term split of both operands (ops.split3_bf16: six partial products, smallest first,
# fp32 accumulation in tensor memory). Measured against fp64: 7.4e-7 relative (the SIMT fp32 kernel: 1.1e-6) at 4.5 x its
# speed. Long reductions (pooling over K, dX over B) stay on the SIMT kernels: the tensor core adds each 16-term partial
# into its fp32 accumulator with truncation, which is only harmless for a few dozen leading-term steps.
FP32_TENSOR_CORES = os.environ.get("ILFDT_FP32_TENSOR_CORES", "1") != "0"


def set_fp32_tensor_cores(on):
    """fp32 mode: True (default) = contractions over the model width run on the tensor cores as a 3-term bf16 split;
    False = SIMT FFMA kernels everywhere."""
    global FP32_TENSOR_CORES
    from . import _lib
    FP32_TENSOR_CORES = bool(on)
    _lib.check(_lib.load().ilfdt_tune_set(_lib.TUNE["f32_x3"], int(bool(on))))


def _x3_ok(q, sd):
    return FP32_TENSOR_CORES and q.dtype == torch.float32 and q.shape[2] % 8 == 0 and q.shape[2] <= 1024 \
        and q.shape[1] <= 256 and q.shape[0] * q.shape[1] < 2 ** 31 // 8


# padded text in bf16 mode: contract only the rows that can matter (csrc/text_buckets.cu); 0 = full-width operand
COMPACT_TEXT = os.environ.get("ILFDT_COMPACT_TEXT", "1") != "0"
LEAN_WEIGHTS = os.environ.get("ILFDT_LEAN_WEIGHTS", "1") != "0"   # need_weights=False skips the fp32 att_weight

# bf16 copies of the codebook ([K,D] and its [D,K] transpose). They are recomputed on every call (two small
# kernels) unless the caller opens an explicit `codebook_scope()` around calls that provably see the same
# codebook values -- Clip_FDT.forward does, around its two Query_models. There is deliberately NO cache keyed on
# (data_ptr, _version): writes through `.data`, raw pointers (reset_codebook, keep_codebook_value) or a CUDA-graph
# replay do not bump `_version`, and a stale bf16 codebook is a silent wrong answer.
_sd_scope = None


class codebook_scope:
    """with codebook_scope(): ... -- calls inside share one bf16 copy per codebook tensor. The scope must not span
    any write to the codebook; it is re-entrant (an inner scope reuses the outer one) and is closed on exit, so the
    casts are issued (and captured, under CUDA-graph capture) once per step."""

    def __enter__(self):
        global _sd_scope
        self.outer = _sd_scope
        if _sd_scope is None:
            _sd_scope = {}
        return self

    def __exit__(self, *a):
        global _sd_scope
        _sd_scope = self.outer


def _sd_bf16(sd):
    key = (sd.data_ptr(), tuple(sd.shape), sd.device)
    if _sd_scope is not None:
        hit = _sd_scope.get(key)
        if hit is not None:
            return hit
    sd_b = ops.cast_bf16(sd.detach())
    sd_bt = _LazyTranspose(sd_b)
    if _sd_scope is not None:
        _sd_scope[key] = (sd_b, sd_bt)
    return sd_b, sd_bt


class _LazyTranspose:
    """The [D, K] copy of the bf16 codebook, made only if a pooling shape needs it (ilfdt_gemm_nn reads sd [K, D] itself)."""

    def __init__(self, src):
        self.src, self.t = src, None

    def get(self):
        if self.t is None:
            self.t = ops.transpose_bf16(self.src)
        return self.t


def _pool_codes(w_b, sd_k, sd_kt):
    """f[b,D] = w[b,K] . sd[K,D]  (clip_fdt.py:154) on the tensor cores."""
    if ops.gemm_nn_supported(sd_k.shape[1], w_b.shape[0], sd_k.shape[0]):
        return ops.gemm_nn(sd_k, w_b)
    return ops.gemm_nt(sd_kt.get(), w_b)


_warned = set()


def _warn_once(msg):
    if msg not in _warned:
        _warned.add(msg)
        warnings.warn(msg, RuntimeWarning, stacklevel=3)


def _score_pool(q, sd_c, mask_f, temperature, mode):
    """q . sd^T -> token max-pool, on the route `mode` and the shapes select: (pooled [b,K], arg [b,K], and the operands
    the backward re-reads: q_k, sd_k, sd_kt, row_base). The one place that picks the route -- `grounding_choices` below
    (the parity tests' view of the kernels' discrete choices) goes through it too."""
    row_base = None
    if mode == "bf16":
        sd_k, sd_kt = _sd_bf16(sd_c)
        # (the backward of the compacted operand runs the v3 / v4 scatter kernels: D in {256, 512}, K % 32 == 0)
        if mask_f is not None and COMPACT_TEXT and ops.text_supported(q.shape[1]) and sd_c.shape[0] % 32 == 0 \
                and q.shape[2] in (256, 512):
            # padded text: valid rows + ONE pad per sample can change the max over tokens (every pad scores
            # exactly 0, clip_fdt.py:125-130); the cast writes that compacted, bucketed operand directly
            q_k, plan = ops.text_compact(q, mask_f)
            row_base = plan.row_base
            pooled, arg = ops.ground_fwd_text(q_k, sd_k, plan, q.shape[1], temperature)
        else:
            # pad rows are zeroed in the bf16 operand: their tensor-core score is exactly the 0 the
            # reference's mask multiply yields (clip_fdt.py:125-130), so the fused epilogue needs no mask
            q_k = ops.cast_bf16_rows(q, mask_f)
            pooled, arg = ops.ground_fwd(q_k, sd_k, None, temperature, "max")
    else:
        q_k, sd_k, sd_kt = q, sd_c, None
        if _x3_ok(q, sd_c):
            # pad rows are zeroed in the split operand: their score is exactly the 0 of clip_fdt.py:125-130
            sd_x3 = ops.split3_bf16(sd_c, 0)
            if mask_f is not None and COMPACT_TEXT and ops.text_supported(q.shape[1]):
                # padded text: only the valid rows + one pad per sample are split and contracted (as in the bf16 mode;
                # the backward reads the fp32 q itself, so nothing else changes)
                q_x3, plan = ops.text_compact_x3(q, mask_f)
                pooled, arg = ops.ground_fwd_text(q_x3, sd_x3, plan, q.shape[1], temperature, sqrt_d=math.sqrt(q.shape[2]))
            else:
                rowmask = None if mask_f is None else mask_f.reshape(-1)
                pooled, arg = ops.ground_fwd(ops.split3_bf16(q, 1, rowmask), sd_x3, None, temperature, "max",
                                             sqrt_d=math.sqrt(q.shape[2]))
        else:
            pooled, arg = ops.ground_fwd(q_k, sd_k, mask_f, temperature, "max")
    return pooled, arg, q_k, sd_k, sd_kt, row_base


def grounding_choices(q, sd, mask, temperature, mode=None):
    """(arg [b,K] int64, pooled [b,K] fp32, support [b,K] bool): the discrete choices GroundingFunction.forward makes on
    these inputs -- same route, same kernels, same operands. Test / diagnostic helper (tests/test_fullsize_gpu.py,
    tools/check_multi_gpu.py): gradients are compared with the fp64 reference rule on the same piece of the piecewise-linear map."""
    mode = get_compute_dtype() if mode is None else mode
    q = q.detach().contiguous()
    if mode == "bf16" and not (q.shape[2] % 8 == 0 and q.shape[2] <= 512 and q.shape[1] <= 256 and sd.shape[0] % 8 == 0):
        mode, q = "fp32", q.float()
    mask_f = None if mask is None else mask.to(torch.float32).contiguous()
    pooled, arg = _score_pool(q, sd.detach().contiguous(), mask_f, temperature, mode)[:2]
    return arg.long(), pooled, ops.sparsemax_fwd(pooled)[0] > 0


class GroundingFunction(torch.autograd.Function):
    """Query_model.forward from q onward (clip_fdt.py:113-157), sparsemax + max pooling:
    (q [b,N,D], sd [K,D], mask [b,N] | None) -> (att_weight [b,K], att_ft [b,D]).
    need_weights=False (what Clip_FDT.forward asks for: it drops att_weight, clip_fdt.py:233-234): on the
    tensor-core route att_weight is returned as the non-differentiable bf16 copy the pooling GEMM consumes and
    the fp32 one is never written (4 bytes/entry less written in the forward, 6 less read in the backward)."""

    @staticmethod
    def forward(ctx, q, sd, mask, temperature, mode, need_weights=True):
        q = q.contiguous()
        sd_c = sd.contiguous()
        ctx.q_dtype = q.dtype
        ctx.set_materialize_grads(False)   # an unused att_weight output must not become a [b,K] tensor of zeros
        if q.dtype == torch.bfloat16 and mode != "bf16":
            raise RuntimeError("ilfdt: a bf16 q needs compute dtype 'bf16'")
        mask_f = None if mask is None else mask.to(torch.float32).contiguous()
        if mode == "bf16" and not (q.shape[2] % 8 == 0 and q.shape[2] <= 512 and q.shape[1] <= 256
                                   and sd.shape[0] % 8 == 0):
            # shapes outside the tcgen05 kernel's tiling take the fp32 CUDA kernels (still libilfdt, never PyTorch math)
            _warn_once("ilfdt: grounding shape q%s x codebook%s is outside the tensor-core tiling (D %% 8 == 0, D <= 512, "
                       "N <= 256, K %% 8 == 0): running the fp32 SIMT kernels" % (tuple(q.shape), tuple(sd.shape)))
            mode = "fp32"
            if q.dtype == torch.bfloat16:
                q = q.float()      # dq is cast back to bf16 at the end of backward
        pooled, arg, q_k, sd_k, sd_kt, row_base = _score_pool(q, sd_c, mask_f, temperature, mode)
        # the scatter kernel that reads the support from bf16 weights covers these shapes (ground_bwd.cu)
        # (token counts up to 96 run the v3 scatter kernel, up to 256 -- ViT-B/16's 196 patches -- the token-sorted one)
        lean = LEAN_WEIGHTS and (not need_weights) and mode == "bf16" and q.shape[2] in (256, 512) and q.shape[1] <= 256 \
            and sd.shape[0] % 4 == 0
        w, w_b, _tau, nnz = ops.sparsemax_fwd(pooled, want_bf16=(mode == "bf16"), want_f32=not lean)
        del pooled
        if mode == "bf16":
            f = _pool_codes(w_b, sd_k, sd_kt)               # f[b,D] = w[b,K] . sd[K,D]  (clip_fdt.py:154)
        else:
            f = ops.gemm_f32(w, sd_c)
        ctx.mode, ctx.temperature = mode, float(temperature)
        ctx.n_tok = q.shape[1]
        ctx.save_for_backward(q_k, sd_k, mask_f, w, arg, w_b, row_base, nnz)
        if lean:
            ctx.mark_non_differentiable(w_b)
            return w_b, f
        return w, f

    @staticmethod
    def backward(ctx, dw, df):
        q_k, sd_k, mask_f, w, arg, w_b, row_base, nnz = ctx.saved_tensors
        mode = ctx.mode
        if w is None:
            w = w_b          # lean route: the support is read from the bf16 weights
        b, K = w.shape
        D = sd_k.shape[1]
        dsd = torch.zeros(K, D, dtype=torch.float32, device=w.device)
        if df is not None:
            df = df.contiguous().float()
            if mode == "bf16":
                df_b = ops.cast_bf16(df)
                g = ops.gemm_nt(sd_k, df_b)                  # g[b,K] = df . sd^T
                # dsd[K,D] += w^T . df straight from the row-major w [b,K] and df [b,D] (MN-major tensor-core operands)
                ops.gemm_tn(df_b, w_b, out=dsd, accumulate=True)
            else:
                if FP32_TENSOR_CORES and D % 8 == 0 and D <= 1024:
                    g = ops.gemm_nt(ops.split3_bf16(sd_k, 0), ops.split3_bf16(df, 1))     # g = df . sd^T, reduction D
                else:
                    g = ops.gemm_f32(df, sd_k, b_t=True)
                ops.gemm_f32(w, df, a_t=True, out=dsd, accumulate=True)
            if dw is not None:
                g.add_(dw)
        else:
            g = dw.contiguous().float().clone()
        dq_dtype = torch.float32
        if ctx.q_dtype == torch.bfloat16 and mode == "bf16" and D in (256, 512, 768, 1024) and ctx.n_tok <= 256 \
                and K % 4 == 0:
            dq_dtype = torch.bfloat16      # written directly by the scatter kernel
        dq = ops.ground_bwd(g, w, arg, q_k, sd_k, mask_f, ctx.temperature, dsd, dq_dtype, row_base, ctx.n_tok, nnz)
        if dq.dtype != ctx.q_dtype:
            dq = ops.cast_bf16(dq)
        return dq, dsd, None, None, None, None


# q_map on the library's kernels: off by default -- the reference keeps it an fp32 nn.Sequential and
# the fp32 parity mode always does; ilfdt.set_fused_q_map(True) / ILFDT_FUSED_QMAP=1 turns it on for the bf16 mode
FUSED_QMAP = os.environ.get("ILFDT_FUSED_QMAP", "0") != "0"
QMAP_WIDTHS = (256, 512, 768, 1024)


def set_fused_q_map(on):
    global FUSED_QMAP
    FUSED_QMAP = bool(on)


def fused_q_map_enabled():
    return FUSED_QMAP and _COMPUTE_DTYPE == "bf16"


class QMapFunction(torch.autograd.Function):
    """Query_model.q_map (clip_fdt.py:86-92,111): LayerNorm(F) -> Linear(F,D) -> GELU -> LayerNorm(D) -> Linear(D,D),
    x [b,N,F] (f32 or bf16) -> q [b,N,D] in bf16, the operand the score kernel streams (no cast pass, half the bytes).
    Both Linear layers are tcgen05 contractions on bf16 operands with fp32 accumulation whose read-out adds the bias
    and rounds to bf16 (ilfdt_linear_bf16: the fp32 product never reaches HBM); LayerNorm and the exact GELU are fp32
    row kernels (csrc/qmap_rows.cu). Saved for the backward: the two bf16 LayerNorm outputs, the bf16 pre-GELU
    activation and the per-row mean / rstd -- 6 bytes per element of the hidden width. Weight gradients use the
    split-reduction contraction (ilfdt_gemm_tn_splitk), bias / affine gradients deterministic two-stage column sums."""

    @staticmethod
    def forward(ctx, x, ln1_w, ln1_b, w1, b1, ln2_w, ln2_b, w2, b2, eps1, eps2):
        D, F = w1.shape
        if F not in QMAP_WIDTHS or D not in QMAP_WIDTHS or tuple(w2.shape) != (D, D):
            raise NotImplementedError("ilfdt fused q_map: widths %s -> %s outside %s" % (F, D, (QMAP_WIDTHS,)))
        x2 = x.detach().reshape(-1, F).contiguous()
        f32 = lambda t: t.detach().float().contiguous()
        ln1_w, ln1_b, b1, ln2_w, ln2_b, b2 = map(f32, (ln1_w, ln1_b, b1, ln2_w, ln2_b, b2))
        xn, _, mean1, rstd1 = ops.ln_fwd(x2, ln1_w, ln1_b, eps1)
        u = ops.linear_bf16(ops.cast_bf16(f32(w1)), xn, b1)             # [R,D] pre-activation xn . W1^T + b1, bf16
        a2, _, mean2, rstd2 = ops.ln_fwd(u, ln2_w, ln2_b, eps2, gelu=True)
        q = ops.linear_bf16(ops.cast_bf16(f32(w2)), a2, b2)             # [R,D] = a2 . W2^T + b2, bf16
        ctx.save_for_backward(x2, xn, mean1, rstd1, ln1_w, w1, u, a2, mean2, rstd2, ln2_w, w2)
        ctx.x_shape, ctx.x_dtype = x.shape, x.dtype
        return q.view(*x.shape[:-1], D)

    @staticmethod
    def backward(ctx, dq):
        x2, xn, mean1, rstd1, ln1_w, w1, u, a2, mean2, rstd2, ln2_w, w2 = ctx.saved_tensors
        D, F = w1.shape
        dq2 = dq.reshape(-1, D).contiguous()
        db2 = ops.colsum(dq2)
        dq_b = dq2 if dq2.dtype == torch.bfloat16 else ops.cast_bf16(dq2.float())
        dw2 = ops.gemm_tn_splitk(a2, dq_b)                                     # [D_out, D_in] = dq^T . a2
        da2 = ops.linear_bf16(ops.transpose_bf16(w2.detach().float().contiguous()), dq_b)   # [R, D_in] = dq . W2, bf16
        du, dln2_w, dln2_b, db1 = ops.ln_bwd(da2, u, ln2_w, mean2, rstd2, gelu=True)
        del da2
        dw1 = ops.gemm_tn_splitk(xn, du)                                       # [D, F] = du^T . xn
        dxn = ops.linear_bf16(ops.transpose_bf16(w1.detach().float().contiguous()), du)     # [R, F] = du . W1, bf16
        dx, dln1_w, dln1_b, _ = ops.ln_bwd(dxn, x2, ln1_w, mean1, rstd1, gelu=False)
        del dxn
        if not ctx.needs_input_grad[0]:
            dx = None
        else:
            if ctx.x_dtype == torch.bfloat16:
                dx = ops.cast_bf16(dx)
            dx = dx.view(ctx.x_shape)
        return dx, dln1_w, dln1_b, dw1, db1, dln2_w, dln2_b, dw2, db2, None, None


def q_map_apply(q_map, x):
    """Runs a reference-shaped q_map nn.Sequential (LayerNorm, Linear, GELU, LayerNorm, Linear) through QMapFunction;
    parameter names and shapes are untouched, so checkpoints interchange."""
    ln1, fc1, _act, ln2, fc2 = q_map
    return QMapFunction.apply(x, ln1.weight, ln1.bias, fc1.weight, fc1.bias, ln2.weight, ln2.bias, fc2.weight, fc2.bias,
                              ln1.eps, ln2.eps)


class GroundingVariantFunction(torch.autograd.Function):
    """Query_model.forward from q onward for the non-default-tuned combinations of
    att_func_type in {softmax, sigmoid, sparsemax} x pool_type in {sum, mean, max} (clip_fdt.py:67-78,
    140-157). fp32 CUDA kernels throughout (these are epilogue variants, not the tuned path)."""

    @staticmethod
    def forward(ctx, q, sd, mask, temperature, att, pool):
        q = q.contiguous().float()
        sd_c = sd.contiguous().float()
        ctx.set_materialize_grads(False)
        mask_f = None if mask is None else mask.to(torch.float32).contiguous()
        pooled, arg = ops.ground_fwd(q, sd_c, mask_f, temperature, pool)
        rowsum = None
        if att == "sparsemax":
            w = ops.sparsemax_fwd(pooled)[0]
        else:
            w, rowsum = ops.att_fwd(pooled, att)
        u = ops.gemm_f32(w, sd_c)                                   # clip_fdt.py:154
        f = ops.rownorm_fwd(u, rowsum) if att == "sigmoid" else u   # :156-157
        ctx.att, ctx.pool, ctx.temperature = att, pool, float(temperature)
        ctx.save_for_backward(q, sd_c, mask_f, w, arg, rowsum, f)
        return w, f

    @staticmethod
    def backward(ctx, dw, df):
        q, sd_c, mask_f, w, arg, rowsum, f = ctx.saved_tensors
        att, pool = ctx.att, ctx.pool
        b, n_tok, D = q.shape
        K = sd_c.shape[0]
        dsd = torch.zeros(K, D, dtype=torch.float32, device=q.device)
        ds = None
        if df is not None:
            df = df.contiguous().float()
            du = df
            if att == "sigmoid":
                du, ds = ops.rownorm_bwd(df, f, rowsum)
            g = ops.gemm_f32(du, sd_c, b_t=True)
            ops.gemm_f32(w, du, a_t=True, out=dsd, accumulate=True)
            if dw is not None:
                g.add_(dw)
        else:
            g = dw.contiguous().float().clone()
        if att == "sparsemax":
            dz = ops.sparsemax_bwd(w, g)
        else:
            dz = ops.att_bwd(w, g, att, rowbias=ds if att == "sigmoid" else None)
        if pool == "max":
            dq = ops.ground_bwd(dz, None, arg, q, sd_c, mask_f, ctx.temperature, dsd)
        else:
            c = 1.0 / (math.sqrt(D) * ctx.temperature) / (n_tok if pool == "mean" else 1)
            dq = ops.token_bcast(ops.gemm_f32(dz, sd_c), mask_f, n_tok, c)
            ops.gemm_f32(dz, ops.token_sum(q, mask_f), a_t=True, alpha=c, out=dsd, accumulate=True)
        return dq, dsd, None, None, None, None


class SparsemaxFunction(torch.autograd.Function):
    """prototype/model/sparsemax.py:22-71 on a [rows, K] view."""

    @staticmethod
    def forward(ctx, z2d):
        w, _, _, _ = ops.sparsemax_fwd(z2d.contiguous().float())
        ctx.save_for_backward(w)
        return w

    @staticmethod
    def backward(ctx, g):
        (w,) = ctx.saved_tensors
        return ops.sparsemax_bwd(w, g.contiguous().float())


class L2NormFunction(torch.autograd.Function):
    """x = f / (||f|| + eps)   (clip_fdt.py:411-412; clip.py:133-134).
    want_lp=True additionally returns the bf16 copy the same kernel writes (non-differentiable): the operand of the
    tensor-core InfoNCE kernels and the payload of the all-gather (half the NVLink bytes, no cast pass afterwards)."""

    @staticmethod
    def forward(ctx, f, eps, want_lp=False):
        x, xb, norm = ops.l2norm_fwd(f.contiguous().float(), eps, want_bf16=bool(want_lp))
        ctx.eps = eps
        ctx.save_for_backward(x, norm)
        if want_lp:
            ctx.mark_non_differentiable(xb)
            return x, xb
        return x

    @staticmethod
    def backward(ctx, g, *unused):
        x, norm = ctx.saved_tensors
        return ops.l2norm_bwd(x, norm, g.contiguous().float(), ctx.eps), None, None


def lp_embeddings(mode, dim):
    """True when the contrastive head runs on the tensor cores for this width (bf16 operands, 16-byte TMA row pitch)."""
    return mode == "bf16" and dim % 8 == 0 and dim <= 512


class GatherHandle:
    """Rows of every rank, rank-major [W*b, D] (the forward of clip_fdt.py:164-188), gathered ASYNCHRONOUSLY: the NCCL
    all-gather is enqueued at construction on the process group's own stream and `wait()` makes the current stream
    depend on it. Clip_FDT.forward starts the image gather before the text grounding and the text gather before the
    logits_per_text half of the loss (which only needs the image rows), so neither sits on the critical path.
    No autograd through it: the fused loss forms the local rows' total gradient from all-gathered log-sum-exps
    (InfoNCEFunction, gathered form); callers that need d/dG use AllGather.apply."""

    def __init__(self, local):
        self.local = local.contiguous()
        self.work = None
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            w = dist.get_world_size()
            self.out = torch.empty((w * self.local.shape[0],) + tuple(self.local.shape[1:]), dtype=self.local.dtype,
                                   device=self.local.device)
            self.work = dist.all_gather_into_tensor(self.out, self.local, async_op=True)
        else:
            self.out = self.local

    def wait(self):
        if self.work is not None:
            self.work.wait()
            self.work = None
        return self.out


class AllGather(torch.autograd.Function):
    """clip_fdt.py:164-188. forward: all ranks' rows, rank-major [W, b, ...]; backward: the reference
    all-reduces the whole [W,b,...] gradient and slices its own rows -- identical to a
    reduce-scatter(SUM), which moves W times fewer bytes over NVLink."""

    @staticmethod
    def forward(ctx, tensor):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            ctx.world, ctx.rank = dist.get_world_size(), dist.get_rank()
            t = tensor.contiguous()
            out = torch.empty((ctx.world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
            dist.all_gather_into_tensor(out.view(-1, *t.shape[1:]) if t.dim() > 0 else out, t)
            return out
        ctx.world, ctx.rank = 1, 0
        return tensor.unsqueeze(0).clone()

    @staticmethod
    def backward(ctx, grad_output):
        if ctx.world == 1:
            return grad_output[0]
        g = grad_output.contiguous()
        if dist.get_backend() != "nccl":  # gloo (CPU tests) has no reduce_scatter: the reference's route
            g = g.clone()
            dist.all_reduce(g, op=dist.ReduceOp.SUM)
            return g[ctx.rank]
        out = torch.empty(g.shape[1:], dtype=g.dtype, device=g.device)
        dist.reduce_scatter_tensor(out, g.view(-1, *g.shape[2:]) if g.dim() > 2 else g.view(-1), op=dist.ReduceOp.SUM)
        return out


class LogitsFunction(torch.autograd.Function):
    """Materialised logits = scale * x @ g^T (clip_fdt.py:421-422); used when a caller really reads
    the [b,B] matrix (accuracy(), visualisation, or a loss other than ClipInfoCELoss)."""

    @staticmethod
    def forward(ctx, x, g, scale, mode):
        x, g = x.contiguous(), g.contiguous()
        sc = scale.detach().reshape(1).float().contiguous()
        if mode == "bf16":
            lp = lambda t: t if t.dtype == torch.bfloat16 else ops.cast_bf16(t.float())
            out = ops.logits(lp(x), lp(g), sc)
            x, g = x.float(), g.float()
        else:
            x, g = x.float(), g.float()
            out = ops.logits(x, g, sc)
        ctx.save_for_backward(x, g, sc, out)
        ctx.scale_shape = scale.shape
        return out

    @staticmethod
    def backward(ctx, dL):
        x, g, sc, out = ctx.saved_tensors
        dL = dL.contiguous().float()
        # fp32 SIMT contractions: this route is off the training hot path
        dLs = dL * sc
        dx = ops.gemm_f32(dLs, g)
        dg = ops.gemm_f32(dLs, x, a_t=True)
        dscale = ((dL * out).sum() / sc).reshape(ctx.scale_shape)
        return dx, dg, dscale, None


class InfoNCEFunction(torch.autograd.Function):
    """Fused logits + ClipInfoCELoss (clip_fdt.py:421-422 + loss_functions/loss.py:40-50):
    (x_img, x_txt [b,D]; g_img, g_txt [B,D]; scale [1]) -> loss []. No [b,B] tensor is created.

    gathered=True states that g_img/g_txt are AllGather(x_img)/AllGather(x_txt) (what Clip_FDT/CLIP.forward
    build, clip_fdt.py:418-419). The backward then all-gathers the per-row log-sum-exps (2b+1 floats per rank)
    and each rank forms the total gradient of its own rows in two launches; g_img/g_txt receive no gradient,
    so AllGather.backward (the reference's [B,D] all_reduce, clip_fdt.py:182-188) never runs. The result is
    the same sum the reference's dX + all_reduce(dG)[rank] produces."""

    @staticmethod
    def forward(ctx, x_img, x_txt, g_img, g_txt, scale, label_offset, mode, gathered=False, keep=None):
        sc = scale.detach().reshape(1).float().contiguous()
        D = x_img.shape[1]
        # the tensor-core kernels need 16-byte TMA row pitches: D % 8. Other shapes take the fp32 kernels
        # (still CUDA, still fused per row chunk).
        lp = lp_embeddings(mode, D)
        want = torch.bfloat16 if lp else torch.float32

        def operand(t):
            t = t.detach().contiguous()
            if t.dtype == want:
                return t
            return ops.cast_bf16(t.float()) if lp else t.float()

        if isinstance(g_img, GatherHandle):
            # gathers in flight (Clip_FDT / CLIP.forward): the handles carry the local rows in the operand dtype.
            # logits_per_text = x_txt . G_img^T needs only the image gather and runs while the text one completes
            assert gathered and isinstance(g_txt, GatherHandle)
            xl_i, xl_t = operand(g_img.local), operand(g_txt.local)
            G_i = operand(g_img.wait())
            lse_t, diag_t = ops.nce_lse(xl_t, G_i, sc, label_offset)
            G_t = operand(g_txt.wait())
            lse_i, diag_i = ops.nce_lse(xl_i, G_t, sc, label_offset)
            loss = ops.nce_loss(lse_i, diag_i, lse_t, diag_t)
            ts = [xl_i, xl_t, G_i, G_t]
        else:
            ts = [operand(t) for t in (x_img, x_txt, g_img, g_txt)]
            loss, lse_i, lse_t, diag_i, _ = ops.nce_fwd(ts[0], ts[1], ts[2], ts[3], sc, label_offset)
        if keep is not None:   # what a later top-k count on logits_per_image needs (lazy.ContrastiveStash.topk_hits)
            keep.update(x=ts[0], g=ts[3], scale=sc, diag=diag_i, label_offset=int(label_offset))
        ctx.save_for_backward(ts[0], ts[1], ts[2], ts[3], sc, lse_i, lse_t)
        ctx.label_offset, ctx.scale_shape, ctx.gathered = int(label_offset), scale.shape, bool(gathered)
        return loss

    @staticmethod
    def backward(ctx, go):
        x_img, x_txt, g_img, g_txt, sc, lse_i, lse_t = ctx.saved_tensors
        go = go.reshape(1).float().contiguous()
        if ctx.gathered:
            b, B = x_img.shape[0], g_img.shape[0]
            if B == b:
                lse_i_all, lse_t_all, go_all = lse_i, lse_t, go
            else:
                world = B // b
                mine = torch.cat([lse_i, lse_t, go])                      # 2b+1 floats
                every = torch.empty(world, 2 * b + 1, dtype=torch.float32, device=mine.device)
                dist.all_gather_into_tensor(every.view(-1), mine)
                lse_i_all = every[:, :b].contiguous().view(-1)
                lse_t_all = every[:, b:2 * b].contiguous().view(-1)
                go_all = every[:, 2 * b].contiguous()
            dx_img, dx_txt, dscale = ops.nce_bwd_gathered(x_img, x_txt, g_img, g_txt, sc, ctx.label_offset,
                                                          lse_i_all, lse_t_all, go_all)
            return dx_img, dx_txt, None, None, dscale.reshape(ctx.scale_shape), None, None, None, None
        dx_img, dx_txt, dg_img, dg_txt, dscale = ops.nce_bwd(x_img, x_txt, g_img, g_txt, sc, ctx.label_offset,
                                                             lse_i, lse_t, go)
        return dx_img, dx_txt, dg_img, dg_txt, dscale.reshape(ctx.scale_shape), None, None, None, None


class CrossEntropyRowsFunction(torch.autograd.Function):
    """mean_r(lse(row r) - row[label_offset + r]) over stored logits: F.cross_entropy of loss.py:47-48
    for callers that pass real [b,B] tensors."""

    @staticmethod
    def forward(ctx, logits, label_offset):
        lg = logits.contiguous().float()
        lse, diag = ops.ce_rows_fwd(lg, label_offset)
        ctx.save_for_backward(lg, lse)
        ctx.label_offset = int(label_offset)
        return (lse - diag).mean()

    @staticmethod
    def backward(ctx, go):
        lg, lse = ctx.saved_tensors
        return ops.ce_rows_bwd(lg, ctx.label_offset, lse, go.reshape(1).float().contiguous(), 1.0 / lg.shape[0]), None
