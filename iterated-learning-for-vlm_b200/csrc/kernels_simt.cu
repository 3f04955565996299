// fp32 (parity-mode) kernels built on the SIMT tile contraction, plus the small row-wise kernels
// shared by both modes (L2 normalise, log-sum-exp over stored logits, casts, Philox fill).
#include "gemm_simt.cuh"
#include "launch.h"

namespace ilfdt {

// ------------------------------------------------------------------------------------------------
// generic strided fp32 GEMM
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SG_THREADS, 2)
gemm_f32_kernel(StridedMat A, StridedMat B, float* __restrict__ C, int64_t c_rs, int64_t c_cs, int Kr,
                float alpha, const float* __restrict__ alpha_dev, int accumulate) {
  __shared__ SimtSmem sm;
  const int m0 = blockIdx.x * SG_BM, n0 = blockIdx.y * SG_BN;
  float acc[8][8];
  simt_mainloop(A, B, Kr, m0, n0, sm, acc);
  if (alpha_dev) alpha *= __ldg(alpha_dev);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int m = m0 + simt_row(i);
    if (m >= A.rows) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int n = n0 + simt_col(j);
      if (n >= B.rows) continue;
      float* c = C + (int64_t)m * c_rs + (int64_t)n * c_cs;
      float v = alpha * acc[i][j];
      *c = accumulate ? (*c + v) : v;
    }
  }
}

int32_t launch_gemm_f32(const float* A, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_rs,
                        int64_t b_cs, float* C, int64_t c_rs, int64_t c_cs, int64_t M, int64_t N,
                        int64_t Kr, float alpha, const float* alpha_dev, int accumulate, cudaStream_t st) {
  if (M == 0 || N == 0) return ILFDT_OK;
  ILFDT_REQUIRE(M < (1ll << 31) && N < (1ll << 31) && Kr < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE,
                "gemm_f32: dimension too large");
  ILFDT_REQUIRE(cdiv(N, SG_BN) <= 65535, ILFDT_E_UNSUPPORTED_SHAPE, "gemm_f32: N too large for grid.y");
  StridedMat a{A, a_rs, a_cs, (int)M}, b{B, b_rs, b_cs, (int)N};
  dim3 grid((unsigned)cdiv(M, SG_BM), (unsigned)cdiv(N, SG_BN));
  gemm_f32_kernel<<<grid, SG_THREADS, 0, st>>>(a, b, C, c_rs, c_cs, (int)Kr, alpha, alpha_dev, accumulate);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// ------------------------------------------------------------------------------------------------
// a1 fp32: q.sd^T tile -> token pooling, never storing the [b, n_tok, K] scores.
// Rows of the contraction are tokens (b*n_tok), columns are codes. Max-pool: each thread reduces
// its 8 token rows per code within a sample and publishes with a 64-bit atomicMax on
// (ordered(score) << 32 | 0xFFFF - token), so ties resolve to the lowest token index.
// Sum/mean-pool: atomicAdd of the fully scaled score.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SG_THREADS, 2)
ground_qk_f32_kernel(StridedMat Q, StridedMat SD, const float* __restrict__ mask, int n_tok, int K,
                     int D, float sqrt_d, float temperature, int pool_type,
                     unsigned long long* __restrict__ keys, float* __restrict__ pooled) {
  __shared__ SimtSmem sm;
  const int m0 = blockIdx.x * SG_BM, n0 = blockIdx.y * SG_BN;
  float acc[8][8];
  simt_mainloop(Q, SD, D, m0, n0, sm, acc);

  int tokrow[8], smp[8], tok[8];
  bool keep[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    tokrow[i] = m0 + simt_row(i);
    bool ok = tokrow[i] < Q.rows;
    smp[i] = ok ? tokrow[i] / n_tok : -1;
    tok[i] = ok ? tokrow[i] - smp[i] * n_tok : 0;
    keep[i] = ok && (mask == nullptr || __ldg(mask + tokrow[i]) == 0.f);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int code = n0 + simt_col(j);
    if (code >= K) continue;
    if (pool_type == ILFDT_POOL_MAX) {
      // candidate per token = keep ? dot : 0 (clip_fdt.py:125-130: pads contribute score 0); the
      // monotone scaling (/sqrt_d, /T) is applied after the max by the unpack kernel.
      int cur = -1;
      unsigned long long best = 0ull;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (smp[i] < 0) continue;
        float v = keep[i] ? acc[i][j] : 0.f;
        unsigned long long key =
            ((unsigned long long)float_to_ordered(v) << 32) | (unsigned long long)(0xFFFFu - (unsigned)tok[i]);
        if (smp[i] != cur) {
          if (cur >= 0) atomicMax(keys + (int64_t)cur * K + code, best);
          cur = smp[i];
          best = key;
        } else {
          best = key > best ? key : best;
        }
      }
      if (cur >= 0) atomicMax(keys + (int64_t)cur * K + code, best);
    } else {
      int cur = -1;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (smp[i] < 0) continue;
        float v = __fdiv_rn(__fdiv_rn(acc[i][j], sqrt_d) * (keep[i] ? 1.f : 0.f), temperature);
        if (smp[i] != cur) {
          if (cur >= 0) atomicAdd(pooled + (int64_t)cur * K + code, s);
          cur = smp[i];
          s = v;
        } else {
          s += v;
        }
      }
      if (cur >= 0) atomicAdd(pooled + (int64_t)cur * K + code, s);
    }
  }
}

__global__ void ground_unpack_kernel(const unsigned long long* __restrict__ keys, int64_t n, float sqrt_d,
                                     float temperature, float* __restrict__ pooled,
                                     int16_t* __restrict__ arg) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned long long k = keys[i];
  float v = ordered_to_float((uint32_t)(k >> 32));
  // same operation order as clip_fdt.py:123,137 (divide by sqrt(d), then by the temperature)
  pooled[i] = __fdiv_rn(__fdiv_rn(v, sqrt_d), temperature);
  if (arg) arg[i] = (int16_t)(0xFFFFu - (unsigned)(k & 0xFFFFull));
}

__global__ void scale_inplace_kernel(float* __restrict__ p, int64_t n, float s) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] *= s;
}

int32_t launch_ground_fwd_f32(const float* q, const float* sd, const float* mask, int64_t b, int64_t n_tok,
                              int64_t K, int64_t D, float sqrt_d, float temperature, int pool_type,
                              float* pooled, int16_t* arg, void* ws, int64_t ws_bytes, cudaStream_t st) {
  const int64_t rows = b * n_tok;
  ILFDT_REQUIRE(rows < (1ll << 31) && cdiv(K, SG_BN) <= 65535, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_fwd(f32): b*n_tok or K too large");
  ILFDT_REQUIRE(n_tok <= 0xFFFF, ILFDT_E_UNSUPPORTED_SHAPE, "ground_fwd: n_tok > 65535");
  StridedMat Q{q, D, 1, (int)rows}, S{sd, D, 1, (int)K};
  dim3 grid((unsigned)cdiv(rows, SG_BM), (unsigned)cdiv(K, SG_BN));
  unsigned long long* keys = nullptr;
  if (pool_type == ILFDT_POOL_MAX) {
    ILFDT_REQUIRE(ws != nullptr && ws_bytes >= b * K * 8, ILFDT_E_INVALID_ARG,
                  "ground_fwd(f32,max): workspace of %lld bytes required", (long long)(b * K * 8));
    keys = reinterpret_cast<unsigned long long*>(ws);
    ILFDT_CUDA_OK(cudaMemsetAsync(keys, 0, b * K * 8, st));
  } else {
    ILFDT_CUDA_OK(cudaMemsetAsync(pooled, 0, b * K * 4, st));
  }
  ground_qk_f32_kernel<<<grid, SG_THREADS, 0, st>>>(Q, S, mask, (int)n_tok, (int)K, (int)D, sqrt_d,
                                                    temperature, pool_type, keys, pooled);
  ILFDT_LAUNCH_OK();
  const int64_t n = b * K;
  if (pool_type == ILFDT_POOL_MAX) {
    ground_unpack_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(keys, n, sqrt_d, temperature, pooled, arg);
    ILFDT_LAUNCH_OK();
  } else if (pool_type == ILFDT_POOL_MEAN) {
    scale_inplace_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(pooled, n, 1.f / (float)n_tok);
    ILFDT_LAUNCH_OK();
  }
  return ILFDT_OK;
}

// ------------------------------------------------------------------------------------------------
// a5: L2 normalise rows. One warp per row (D = 512 -> 16 elements per lane).
// ------------------------------------------------------------------------------------------------
__global__ void l2norm_fwd_kernel(const float* __restrict__ f, int64_t rows, int D, float eps,
                                  float* __restrict__ x, __nv_bfloat16* __restrict__ xb,
                                  float* __restrict__ norm) {
  int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* fr = f + row * D;
  float ss = 0.f;
  for (int i = lane; i < D; i += 32) {
    float v = fr[i];
    ss = fmaf(v, v, ss);
  }
  ss = warp_sum(ss);
  float nrm = sqrtf(ss);
  float den = nrm + eps;
  for (int i = lane; i < D; i += 32) {
    float v = __fdiv_rn(fr[i], den);
    if (x) x[row * D + i] = v;
    if (xb) xb[row * D + i] = __float2bfloat16_rn(v);
  }
  if (lane == 0 && norm) norm[row] = nrm;
}

// x = f/(n+eps)  =>  df = g/(n+eps) - f * (f.g) / (n * (n+eps)^2) = (g - x * (x.g) * (n+eps)/n) / (n+eps)
__global__ void l2norm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ norm,
                                  const float* __restrict__ g, int64_t rows, int D, float eps,
                                  float* __restrict__ df) {
  int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float* xr = x + row * D;
  const float* gr = g + row * D;
  float dot = 0.f;
  for (int i = lane; i < D; i += 32) dot = fmaf(xr[i], gr[i], dot);
  dot = warp_sum(dot);
  float n = norm[row], den = n + eps;
  float c = (n > 0.f) ? dot * den / n : 0.f;
  for (int i = lane; i < D; i += 32) df[row * D + i] = (gr[i] - xr[i] * c) / den;
}

int32_t launch_l2norm_fwd(const float* f, int64_t rows, int64_t D, float eps, float* x, void* xb, float* norm,
                          cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  l2norm_fwd_kernel<<<(unsigned)cdiv(rows, 8), 256, 0, st>>>(f, rows, (int)D, eps, x,
                                                            reinterpret_cast<__nv_bfloat16*>(xb), norm);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_l2norm_bwd(const float* x, const float* norm, const float* g, int64_t rows, int64_t D, float eps,
                          float* df, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  l2norm_bwd_kernel<<<(unsigned)cdiv(rows, 8), 256, 0, st>>>(x, norm, g, rows, (int)D, eps, df);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// ------------------------------------------------------------------------------------------------
// a10 fp32 route: rows of stored logits -> lse, label logit;  then P = (softmax - onehot) * coef.
// One CTA (256 threads) per row, online (max, sum) in registers, block combine.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void lse_combine(float& m, float& s, float m2, float s2) {
  float mm = fmaxf(m, m2);
  if (mm == -INFINITY) { m = mm; s = 0.f; return; }
  s = s * __expf(m - mm) + s2 * __expf(m2 - mm);
  m = mm;
}

__global__ void __launch_bounds__(256)
lse_rows_kernel(const float* __restrict__ L, int64_t ld, int64_t cols, int64_t label0,
                float* __restrict__ lse, float* __restrict__ diag) {
  const int64_t row = blockIdx.x;
  const float* lr = L + row * ld;
  float m = -INFINITY, s = 0.f;
  for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) {
    float v = lr[c];
    if (v > m) { s = s * expf(m - v) + 1.f; m = v; }
    else s += expf(v - m);
  }
  __shared__ float sm_m[8], sm_s[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    lse_combine(m, s, m2, s2);
  }
  if ((threadIdx.x & 31) == 0) { sm_m[threadIdx.x >> 5] = m; sm_s[threadIdx.x >> 5] = s; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) lse_combine(m, s, sm_m[w], sm_s[w]);
    lse[row] = m + logf(s);
    diag[row] = lr[label0 + row];
  }
}

// In place: L[r,c] <- (exp(L[r,c] - lse[r]) - [c == label0 + r]) * go[0] * coef.
// d scale: the row's sum_c P[r,c] * L[r,c] / scale is formed WITHOUT its cancellation: since sum_c softmax = 1,
//   sum_c (softmax_c - [c == label]) * L_c  ==  sum_c softmax_c * (L_c - L_label),
// every term of the right-hand side is small where the left-hand side subtracts two O(|L|) numbers (the reference's
// autograd sits at 1e-7 of the fp64 value on config 1; the naive fp32 form sat at 1e-4). Threads sum in fp32, warps and
// rows in fp64; the per-row results go to row_acc[row] and are added in a fixed order by sum_rows_kernel (no atomics).
__device__ __forceinline__ void block_row_acc(float part, double* __restrict__ row_acc, int64_t row) {
  double dpart = warp_sum((double)part);
  __shared__ double sm[8];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = dpart;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w2 = 1; w2 < (int)(blockDim.x >> 5); ++w2) dpart += sm[w2];
    row_acc[row] = dpart;
  }
}

__global__ void __launch_bounds__(256)
softmax_grad_rows_kernel(const float* L, float* out, int64_t ld, int64_t cols, int64_t label0,
                         const float* __restrict__ lse, const float* __restrict__ go, float coef,
                         const float* __restrict__ scale_dev, double* __restrict__ row_acc) {
  const int64_t row = blockIdx.x;
  const float* lr = L + row * ld;
  float* orow = out + row * ld;
  const float l = lse[row];
  const float w = go[0] * coef;
  const float inv_scale = scale_dev ? 1.f / __ldg(scale_dev) : 0.f;
  const float v_lab = lr[label0 + row];
  __syncthreads();   // out may alias L: everyone has the label logit before anyone overwrites it
  float part = 0.f;
  for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) {
    float v = lr[c];
    float e = expf(v - l);
    part = fmaf(e * w, (v - v_lab) * inv_scale, part);
    orow[c] = (e - ((c == label0 + row) ? 1.f : 0.f)) * w;
  }
  if (row_acc == nullptr) return;
  block_row_acc(part, row_acc, row);
}

// Gathered form (ilfdt_nce_bwd_gathered, fp32 route), in place on a chunk of stored logits:
//   out[r,c] = wl * (exp(L[r,c] - lse_lane[r]) - d) + ws_c * (exp(L[r,c] - lse_stream[c]) - d),  d = [c == label0 + r]
// with wl = go_lane[0]*coef, ws_c = go_stream[c / rows_per_go]*coef. dscale accumulates the lane term only.
__global__ void __launch_bounds__(256)
softmax_grad_sym_rows_kernel(const float* L, float* out, int64_t ld, int64_t cols, int64_t label0,
                             const float* __restrict__ lse_lane, const float* __restrict__ lse_stream,
                             const float* __restrict__ go_lane, const float* __restrict__ go_stream,
                             int64_t rows_per_go, float coef, const float* __restrict__ scale_dev,
                             double* __restrict__ row_acc) {
  const int64_t row = blockIdx.x;
  const float* lr = L + row * ld;
  float* orow = out + row * ld;
  const float l = lse_lane[row];
  const float wl = go_lane[0] * coef;
  const float inv_scale = 1.f / __ldg(scale_dev);
  const float v_lab = lr[label0 + row];
  __syncthreads();
  float part = 0.f;
  for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) {
    const float v = lr[c];
    const float d = (c == label0 + row) ? 1.f : 0.f;
    const float e = expf(v - l);
    const float ps = (expf(v - lse_stream[c]) - d) * (go_stream[c / rows_per_go] * coef);
    part = fmaf(e * wl, (v - v_lab) * inv_scale, part);
    orow[c] = (e - d) * wl + ps;
  }
  block_row_acc(part, row_acc, row);
}

// out[0] = sum of n doubles in a fixed order (single CTA): the rows' d scale contributions
__global__ void __launch_bounds__(256) sum_rows_kernel(const double* __restrict__ acc, int64_t n, float* __restrict__ out) {
  double s = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += 256) s += acc[i];
  s = warp_sum(s);
  __shared__ double sm[8];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) s += sm[w];
    out[0] = (float)s;
  }
}
int32_t launch_sum_rows(const double* acc, int64_t n, float* out, cudaStream_t st) {
  sum_rows_kernel<<<1, 256, 0, st>>>(acc, n, out);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_softmax_grad_sym_rows(const float* L, float* out, int64_t ld, int64_t rows, int64_t cols, int64_t label0,
                                     const float* lse_lane, const float* lse_stream, const float* go_lane,
                                     const float* go_stream, int64_t rows_per_go, float coef, const float* scale_dev,
                                     double* row_acc, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  softmax_grad_sym_rows_kernel<<<(unsigned)rows, 256, 0, st>>>(L, out, ld, cols, label0, lse_lane, lse_stream, go_lane,
                                                              go_stream, rows_per_go, coef, scale_dev, row_acc);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_lse_rows(const float* L, int64_t ld, int64_t rows, int64_t cols, int64_t label0, float* lse,
                        float* diag, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  lse_rows_kernel<<<(unsigned)rows, 256, 0, st>>>(L, ld, cols, label0, lse, diag);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_softmax_grad_rows(const float* L, float* out, int64_t ld, int64_t rows, int64_t cols, int64_t label0,
                                 const float* lse, const float* go, float coef, const float* scale_dev,
                                 double* row_acc, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  softmax_grad_rows_kernel<<<(unsigned)rows, 256, 0, st>>>(L, out, ld, cols, label0, lse, go, coef, scale_dev,
                                                          row_acc);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// loss[0] = (mean(lse_i - diag_i) + mean(lse_t - diag_t)) / 2     (loss.py:47-49). Single CTA,
// fixed summation order -> deterministic.
__global__ void __launch_bounds__(1024)
nce_loss_kernel(const float* __restrict__ lse_i, const float* __restrict__ diag_i,
                const float* __restrict__ lse_t, const float* __restrict__ diag_t, int64_t b,
                float* __restrict__ loss) {
  double a = 0.0;
  for (int64_t r = threadIdx.x; r < b; r += blockDim.x)
    a += (double)(lse_i[r] - diag_i[r]) + (double)(lse_t[r] - diag_t[r]);
  a = warp_sum(a);
  __shared__ double sm[32];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) a += sm[w];
    loss[0] = (float)(a / (2.0 * (double)b));
  }
}
int32_t launch_nce_loss(const float* lse_i, const float* diag_i, const float* lse_t, const float* diag_t,
                        int64_t b, float* loss, cudaStream_t st) {
  nce_loss_kernel<<<1, 1024, 0, st>>>(lse_i, diag_i, lse_t, diag_t, b, loss);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// N1: top-1 / top-5 hits of the label logit per row (misc.py:464-477). One CTA per row.
__global__ void __launch_bounds__(256)
topk_hits_kernel(const float* __restrict__ L, int64_t cols, int64_t label0, int32_t* __restrict__ counts) {
  const int64_t row = blockIdx.x;
  const float* lr = L + row * cols;
  const float ref = lr[label0 + row];
  int gt = 0;
  for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) gt += (lr[c] > ref) ? 1 : 0;
  gt = warp_sum(gt);
  __shared__ int sm[8];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = gt;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) gt += sm[w];
    if (gt < 1) atomicAdd(counts + 0, 1);
    if (gt < 5) atomicAdd(counts + 1, 1);
  }
}
// accumulating form (counts not cleared): one row chunk of a streamed logits matrix
int32_t launch_topk_hits_acc(const float* L, int64_t rows, int64_t B, int64_t label0, int32_t* counts, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  topk_hits_kernel<<<(unsigned)rows, 256, 0, st>>>(L, B, label0, counts);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_topk_hits(const float* L, int64_t b, int64_t B, int64_t label0, int32_t* counts, cudaStream_t st) {
  ILFDT_CUDA_OK(cudaMemsetAsync(counts, 0, 8, st));
  if (b == 0) return ILFDT_OK;
  topk_hits_kernel<<<(unsigned)b, 256, 0, st>>>(L, B, label0, counts);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// ------------------------------------------------------------------------------------------------
// casts / transposes feeding the bf16 tensor-core kernels
// ------------------------------------------------------------------------------------------------
__global__ void cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i + 7 < n && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
    float4 a = __ldg(reinterpret_cast<const float4*>(src + i));
    float4 b = __ldg(reinterpret_cast<const float4*>(src + i) + 1);
    __nv_bfloat162 o[4] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w),
                           __floats2bfloat162_rn(b.x, b.y), __floats2bfloat162_rn(b.z, b.w)};
    *reinterpret_cast<uint4*>(dst + i) = *reinterpret_cast<uint4*>(o);
  } else {
    for (int j = 0; j < 8 && i + j < n; ++j) dst[i + j] = __float2bfloat16_rn(src[i + j]);
  }
}
int32_t launch_cast_bf16(const float* src, void* dst, int64_t n, cudaStream_t st) {
  if (n == 0) return ILFDT_OK;
  cast_bf16_kernel<<<(unsigned)cdiv(n, 256 * 8), 256, 0, st>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

__global__ void cast_bf16_rows_kernel(const float* __restrict__ src, const float* __restrict__ rowmask, int64_t rows,
                                      int cols, __nv_bfloat16* __restrict__ dst) {
  // one warp per row-segment of 256 elements (8 per lane); cols % 8 == 0
  const int64_t total = rows * (int64_t)(cols / 8);
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int64_t r = i / (cols / 8);
  const int64_t e = i * 8;
  const bool keep = __ldg(rowmask + r) == 0.f;
  uint4 o = make_uint4(0, 0, 0, 0);
  if (keep) {
    float4 a = __ldg(reinterpret_cast<const float4*>(src + e));
    float4 b = __ldg(reinterpret_cast<const float4*>(src + e) + 1);
    __nv_bfloat162 t[4] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w),
                           __floats2bfloat162_rn(b.x, b.y), __floats2bfloat162_rn(b.z, b.w)};
    o = *reinterpret_cast<uint4*>(t);
  }
  *reinterpret_cast<uint4*>(dst + e) = o;
}
int32_t launch_cast_bf16_rows(const float* src, const float* rowmask, int64_t rows, int64_t cols, void* dst,
                              cudaStream_t st) {
  if (rows * cols == 0) return ILFDT_OK;
  if (rowmask == nullptr) return launch_cast_bf16(src, dst, rows * cols, st);
  ILFDT_REQUIRE(cols % 8 == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0,
                ILFDT_E_UNSUPPORTED_SHAPE, "cast_bf16_rows: cols %% 8 == 0 and 16-byte aligned pointers required");
  const int64_t total = rows * (cols / 8);
  cast_bf16_rows_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, rowmask, rows, (int)cols,
                                                                   reinterpret_cast<__nv_bfloat16*>(dst));
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// fp32 -> three bf16 terms hi + mid + lo (8 + 8 + 8 mantissa bits: the sum reproduces the fp32 value to 2^-24), laid out
// as SIX concatenated copies of the row so that ONE bf16 tensor-core contraction over 6*cols accumulates the six
// partial products a0*b0 + a0*b1 + a1*b0 + a0*b2 + a1*b1 + a2*b0 in fp32 (everything down to 2^-16 of the leading term;
// the dropped a1*b2, a2*b1, a2*b2 are below 2^-24): side 0 (A operand) = [2 1 0 1 0 0], side 1 (B operand) = [0 1 2 0 1 0].
// This is the fp32-grade score contraction on tcgen05. Masked rows (rowmask != 0) are zeroed.
__global__ void split3_bf16_kernel(const float* __restrict__ src, const float* __restrict__ rowmask, int64_t rows, int cols,
                                   int side, __nv_bfloat16* __restrict__ dst) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 4 source elements
  const int c4 = cols >> 2;
  if (i >= rows * c4) return;
  const int64_t r = i / c4;
  const int c = (int)(i - r * c4) * 4;
  float4 v = __ldg(reinterpret_cast<const float4*>(src + r * cols + c));
  if (rowmask != nullptr && __ldg(rowmask + r) != 0.f) v = make_float4(0.f, 0.f, 0.f, 0.f);
  const float x[4] = {v.x, v.y, v.z, v.w};
  __nv_bfloat16 t[3][4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat16 h = __float2bfloat16_rn(x[e]);
    const float r1 = x[e] - __bfloat162float(h);
    const __nv_bfloat16 m = __float2bfloat16_rn(r1);
    const float r2 = r1 - __bfloat162float(m);
    t[0][e] = h; t[1][e] = m; t[2][e] = __float2bfloat16_rn(r2);
  }
  // smallest products first: the tensor core adds each 16-term partial into the fp32 accumulator with truncation, so the
  // error of an add scales with the accumulator's magnitude -- keep it small until the leading term comes last
  const int pat[2][6] = {{2, 1, 0, 1, 0, 0}, {0, 1, 2, 0, 1, 0}};
  __nv_bfloat16* o = dst + r * (int64_t)(6 * cols) + c;
#pragma unroll
  for (int s6 = 0; s6 < 6; ++s6) {
    const int w = pat[side][s6];
    *reinterpret_cast<uint2*>(o + (int64_t)s6 * cols) = *reinterpret_cast<const uint2*>(t[w]);
  }
}
int32_t launch_split3_bf16(const float* src, const float* rowmask, int64_t rows, int64_t cols, int side, void* dst,
                           cudaStream_t st) {
  if (rows * cols == 0) return ILFDT_OK;
  ILFDT_REQUIRE(cols % 4 == 0 && ((reinterpret_cast<uintptr_t>(src) & 15) | (reinterpret_cast<uintptr_t>(dst) & 7)) == 0,
                ILFDT_E_UNSUPPORTED_SHAPE, "split3_bf16: cols %% 4 == 0 and aligned pointers required");
  const int64_t total = rows * (cols / 4);
  split3_bf16_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, rowmask, rows, (int)cols, side,
                                                                reinterpret_cast<__nv_bfloat16*>(dst));
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// bf16 -> bf16 copy with the masked (pad) rows zeroed: the q operand of the tensor-core score kernel when
// the caller already holds q in bf16 (same contract as cast_bf16_rows for fp32 q).
__global__ void mask_rows_bf16_kernel(const uint4* __restrict__ src, const float* __restrict__ rowmask, int64_t rows,
                                      int cols8, uint4* __restrict__ dst) {
  const int64_t total = rows * (int64_t)cols8;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const bool keep = __ldg(rowmask + i / cols8) == 0.f;
  dst[i] = keep ? __ldg(src + i) : make_uint4(0u, 0u, 0u, 0u);
}
int32_t launch_mask_rows_bf16(const void* src, const float* rowmask, int64_t rows, int64_t cols, void* dst,
                              cudaStream_t st) {
  if (rows * cols == 0) return ILFDT_OK;
  ILFDT_REQUIRE(cols % 8 == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0,
                ILFDT_E_UNSUPPORTED_SHAPE, "mask_rows_bf16: cols %% 8 == 0 and 16-byte aligned pointers required");
  const int64_t total = rows * (cols / 8);
  mask_rows_bf16_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(reinterpret_cast<const uint4*>(src), rowmask, rows,
                                                                   (int)(cols / 8), reinterpret_cast<uint4*>(dst));
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

template <typename T>
__global__ void transpose_bf16_kernel(const T* __restrict__ src, int64_t rows, int64_t cols,
                                      __nv_bfloat16* __restrict__ dst) {
  __shared__ float tile[32][33];
  int64_t c0 = (int64_t)blockIdx.x * 32, r0 = (int64_t)blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    int64_t r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < rows && c < cols) ? (float)src[r * cols + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    int64_t c = c0 + i, r = r0 + threadIdx.x;
    if (c < cols && r < rows) dst[c * rows + r] = __float2bfloat16_rn(tile[threadIdx.x][i]);
  }
}
int32_t launch_transpose_bf16(const void* src, int src_dtype, int64_t rows, int64_t cols, void* dst,
                              cudaStream_t st) {
  if (rows == 0 || cols == 0) return ILFDT_OK;
  ILFDT_REQUIRE(cdiv(rows, 32) <= 65535, ILFDT_E_UNSUPPORTED_SHAPE, "transpose: too many rows");
  dim3 grid((unsigned)cdiv(cols, 32), (unsigned)cdiv(rows, 32)), block(32, 8);
  if (src_dtype == ILFDT_F32)
    transpose_bf16_kernel<float><<<grid, block, 0, st>>>(reinterpret_cast<const float*>(src), rows, cols,
                                                         reinterpret_cast<__nv_bfloat16*>(dst));
  else
    transpose_bf16_kernel<__nv_bfloat16><<<grid, block, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(src), rows, cols, reinterpret_cast<__nv_bfloat16*>(dst));
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// ------------------------------------------------------------------------------------------------
// a12: Philox4x32-10 + Box-Muller N(0,1) fill. Integer stream is bit-exact with
// oracle/fdt_oracle.py:philox_words; one thread produces 4 outputs from one counter block.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}
__global__ void randn_fill_kernel(float* __restrict__ dst, int64_t n, uint32_t k0, uint32_t k1, uint32_t gen) {
  int64_t blk = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (blk * 4 >= n) return;
  uint32_t c[4] = {(uint32_t)(blk & 0xffffffffll), (uint32_t)(blk >> 32), gen, 0u};
  philox4x32_10(c, k0, k1);
  const float two_pow_m32 = 2.3283064365386963e-10f;
  float u0 = ((float)c[0] + 0.5f) * two_pow_m32, u1 = ((float)c[1] + 0.5f) * two_pow_m32;
  float u2 = ((float)c[2] + 0.5f) * two_pow_m32, u3 = ((float)c[3] + 0.5f) * two_pow_m32;
  // (float)c + 0.5 can round up to 2^32 -> u == 1: log(1) = 0 is fine; u is never 0.
  u0 = fminf(u0, 1.f); u2 = fminf(u2, 1.f);
  float r0 = sqrtf(-2.f * logf(u0)), r1 = sqrtf(-2.f * logf(u2));
  float s0, c0, s1, c1;
  sincospif(2.f * u1, &s0, &c0);
  sincospif(2.f * u3, &s1, &c1);
  float o[4] = {r0 * c0, r0 * s0, r1 * c1, r1 * s1};
  for (int j = 0; j < 4 && blk * 4 + j < n; ++j) dst[blk * 4 + j] = o[j];
}
int32_t launch_randn_fill(float* dst, int64_t n, uint64_t seed, uint32_t gen, cudaStream_t st) {
  if (n == 0) return ILFDT_OK;
  int64_t nblk = cdiv(n, 4);
  randn_fill_kernel<<<(unsigned)cdiv(nblk, 256), 256, 0, st>>>(dst, n, (uint32_t)(seed & 0xffffffffull),
                                                             (uint32_t)(seed >> 32), gen);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

}  // namespace ilfdt
