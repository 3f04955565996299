// Row kernels of the q_map producer (clip_fdt.py:86-92: LayerNorm -> Linear -> GELU -> LayerNorm ->
// Linear). The two Linear layers run on the tcgen05 contraction kernels (gemm_umma.cu); everything between them is
// here: LayerNorm forward with an optional bias + exact-erf GELU in front of it, its backward (through the GELU when
// there is one) with deterministic two-stage column sums for the affine / bias gradients, the bias add + bf16 cast of
// the last layer, and a column sum. All HBM-bound: one warp per row, the row held in registers (C/32 values per lane,
// C in {256, 512, 768, 1024}), 16-byte accesses, no atomics.
#include "common.cuh"
#include "launch.h"
#include <algorithm>
#include <mutex>

namespace ilfdt {

constexpr int QM_THREADS = 256;            // 8 warps = 8 rows in flight per CTA
constexpr int QM_WARPS = QM_THREADS / 32;

// Exact-erf GELU (nn.GELU(), clip_fdt.py:89) with erf from Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, below fp32
// rounding of the products it enters): Phi(v) = 0.5 (1 + erf(v / sqrt 2)), erf(x) = sign(x) (1 - P(t) e^{-x^2}),
// t = 1 / (1 + p |x|). With x = v / sqrt 2 the exponential is e^{-v^2 / 2}: the same one the density phi(v) needs, so the
// backward pays one MUFU.EX2 and one MUFU.RCP per element instead of erff's branchy polynomial plus an expf
// (ncu r01: the GELU forms of ln_fwd / ln_bwd were issue-bound at 58 instructions per element).
__device__ __forceinline__ void gelu_parts(float v, float& cdf, float& e) {
  const float ax = fabsf(v) * 0.70710678118654752440f;
  const float t = __fdividef(1.f, fmaf(0.3275911f, ax, 1.f));
  e = exp2f(v * v * -0.72134752044448170368f);                 // e^{-v^2/2}
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float half_tail = 0.5f * poly * t * e;                 // 0.5 (1 - erf|x|)
  cdf = v >= 0.f ? 1.f - half_tail : half_tail;
}
__device__ __forceinline__ float gelu_erf(float v) {
  float cdf, e;
  gelu_parts(v, cdf, e);
  return v * cdf;
}
// a = v Phi(v) and d a / d v = Phi(v) + v phi(v)
__device__ __forceinline__ void gelu_erf_and_grad(float v, float& a, float& g) {
  float cdf, e;
  gelu_parts(v, cdf, e);
  a = v * cdf;
  g = fmaf(v * 0.39894228040143267794f, e, cdf);
}

// 4 consecutive elements <-> fp32 registers, fp32 or bf16 storage
__device__ __forceinline__ void ld4(const float* p, float (&v)[4]) {
  float4 a = *reinterpret_cast<const float4*>(p);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
__device__ __forceinline__ void ld4(const __nv_bfloat16* p, float (&v)[4]) {
  uint2 r = *reinterpret_cast<const uint2*>(p);
  v[0] = __uint_as_float(r.x << 16); v[1] = __uint_as_float(r.x & 0xffff0000u);
  v[2] = __uint_as_float(r.y << 16); v[3] = __uint_as_float(r.y & 0xffff0000u);
}
__device__ __forceinline__ void st4(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4(__nv_bfloat16* p, const float (&v)[4]) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
  __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&a);
  r.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = r;
}

// y = LN(act(src + bias)) * gamma + beta, act = GELU (erf) or identity; mean / rstd saved per row.
// GELU: the pre-activation u = src + bias is stored as bf16 (`pre`), which is what the backward differentiates.
// NV = C / 128 float4 groups per lane; column of element e of group j: j*128 + lane*4 + e.
template <int NV, bool GELU, typename TIn>
__global__ void __launch_bounds__(QM_THREADS)
ln_fwd_kernel(const TIn* __restrict__ src, const float* __restrict__ bias, const float* __restrict__ gamma,
              const float* __restrict__ beta, int64_t rows, float eps, __nv_bfloat16* __restrict__ y,
              __nv_bfloat16* __restrict__ pre, float* __restrict__ mean, float* __restrict__ rstd) {
  constexpr int C = NV * 128;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t)blockIdx.x * QM_WARPS + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * QM_WARPS;
  float gm[NV][4], bt[NV][4], bs[NV][4];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    ld4(gamma + j * 128 + lane * 4, gm[j]);
    ld4(beta + j * 128 + lane * 4, bt[j]);
    if (GELU && bias != nullptr) ld4(bias + j * 128 + lane * 4, bs[j]);
  }
  for (int64_t r = warp0; r < rows; r += nwarps) {
    float v[NV][4];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      ld4(src + r * C + j * 128 + lane * 4, v[j]);
      if (GELU) {
        if (bias != nullptr) {
#pragma unroll
          for (int e = 0; e < 4; ++e) v[j][e] += bs[j][e];
        }
        if (pre != nullptr) st4(pre + r * C + j * 128 + lane * 4, v[j]);
#pragma unroll
        for (int e = 0; e < 4; ++e) v[j][e] = gelu_erf(v[j][e]);
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) s += v[j][e];
    }
    const float mu = warp_sum(s) * (1.f / C);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float d = v[j][e] - mu;
        q = fmaf(d, d, q);
      }
    const float rs = rsqrtf(warp_sum(q) * (1.f / C) + eps);   // biased variance, as nn.LayerNorm
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      float o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) o[e] = fmaf((v[j][e] - mu) * rs, gm[j][e], bt[j][e]);
      st4(y + r * C + j * 128 + lane * 4, o);
    }
    if (lane == 0) {
      mean[r] = mu;
      rstd[r] = rs;
    }
  }
}

// Backward of y = LN(act(src)) * gamma + beta for a given dy [rows, C] (fp32):
//   xhat = (act(src) - mean) * rstd;  dxhat = dy * gamma
//   d act = rstd * (dxhat - mean_c(dxhat) - xhat * mean_c(dxhat * xhat));  dsrc = d act * act'(src)
// and the per-CTA column partials part[cta][0] = sum_r dy*xhat (dgamma), [1] = sum_r dy (dbeta), [2] = sum_r dsrc (the
// gradient of the bias in front of the GELU). dsrc is written as TOut (bf16 when it feeds the next contraction).
// 128-thread CTAs: at 120-210 registers per thread this keeps 3 (F = 768) to 4 CTAs resident per SM instead of one
// 256-thread CTA, i.e. more rows (loads) in flight per SM
constexpr int LB_THREADS = 128, LB_WARPS = LB_THREADS / 32;

template <int NV, bool GELU, typename TIn, typename TOut, typename TDy>
__global__ void __launch_bounds__(LB_THREADS)
ln_bwd_kernel(const TDy* __restrict__ dy, const TIn* __restrict__ src, const float* __restrict__ gamma,
              const float* __restrict__ mean, const float* __restrict__ rstd, int64_t rows, TOut* __restrict__ dsrc,
              float* __restrict__ part) {
  constexpr int C = NV * 128;
  __shared__ float red[LB_WARPS][128];
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  const int64_t warp0 = (int64_t)blockIdx.x * LB_WARPS + wid;
  const int64_t nwarps = (int64_t)gridDim.x * LB_WARPS;
  float gm[NV][4], ag[NV][4], ab[NV][4], as[NV][4];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    ld4(gamma + j * 128 + lane * 4, gm[j]);
#pragma unroll
    for (int e = 0; e < 4; ++e) ag[j][e] = ab[j][e] = as[j][e] = 0.f;
  }
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const float mu = mean[r], rs = rstd[r];
    float xh[NV][4], dx[NV][4], gp[NV][4];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      float u[4], d[4];
      ld4(src + r * C + j * 128 + lane * 4, u);
      ld4(dy + r * C + j * 128 + lane * 4, d);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float a = u[e];
        gp[j][e] = 1.f;
        if (GELU) gelu_erf_and_grad(u[e], a, gp[j][e]);
        xh[j][e] = (a - mu) * rs;
        ag[j][e] = fmaf(d[e], xh[j][e], ag[j][e]);
        ab[j][e] += d[e];
        dx[j][e] = d[e] * gm[j][e];
        s1 += dx[j][e];
        s2 = fmaf(dx[j][e], xh[j][e], s2);
      }
    }
    s1 = warp_sum(s1) * (1.f / C);
    s2 = warp_sum(s2) * (1.f / C);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      float o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        o[e] = rs * (dx[j][e] - s1 - xh[j][e] * s2) * gp[j][e];
        as[j][e] += o[e];
      }
      st4(dsrc + r * C + j * 128 + lane * 4, o);
    }
  }
  // CTA partials: fixed order over the 4 warps, one 128-column group at a time
  float* out = part + (int64_t)blockIdx.x * 3 * C;
#pragma unroll
  for (int which = 0; which < (GELU ? 3 : 2); ++which) {
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const float(&a)[4] = which == 0 ? ag[j] : which == 1 ? ab[j] : as[j];
      __syncthreads();
#pragma unroll
      for (int e = 0; e < 4; ++e) red[wid][lane * 4 + e] = a[e];
      __syncthreads();
      {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < LB_WARPS; ++w) t += red[w][threadIdx.x];
        out[which * C + j * 128 + threadIdx.x] = t;
      }
    }
  }
}

// out[c] = sum_p part[p * stride + c], c < cols. 32 columns x 8 partial lanes per CTA; every lane sums its partials
// p = lane, lane + 8, ... and the 8 lanes are folded in a fixed order: deterministic, and 8 loads in flight per column
// instead of one dependent chain over several hundred partials.
__global__ void __launch_bounds__(256)
partial_reduce_kernel(const float* __restrict__ part, int n_part, int64_t stride, int cols, float* __restrict__ out) {
  __shared__ float red[8][32];
  const int cx = threadIdx.x & 31, ly = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float t = 0.f;
  if (c < cols)
    for (int p = ly; p < n_part; p += 8) t += part[p * stride + c];
  red[ly][cx] = t;
  __syncthreads();
  if (ly == 0 && c < cols) {
    float a = 0.f;
#pragma unroll
    for (int l = 0; l < 8; ++l) a += red[l][cx];
    out[c] = a;
  }
}

// dst[r, :] = src[r, :] + bias   (bf16 or fp32 out)
template <typename TOut>
__global__ void bias_cast_kernel(const float* __restrict__ src, const float* __restrict__ bias, int64_t n4, int C,
                                 TOut* __restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float v[4], b[4];
    ld4(src + i * 4, v);
    ld4(bias + (int)((i * 4) % C), b);
#pragma unroll
    for (int e = 0; e < 4; ++e) v[e] += b[e];
    st4(dst + i * 4, v);
  }
}

// part[cta][c] = sum over the CTA's rows of src[r, c]. The 256 threads form RL row lanes x (C/4) column groups of four
// columns; each lane keeps 4 rows in flight, the row lanes are folded through shared memory in a fixed order.
template <typename TIn>
__global__ void __launch_bounds__(QM_THREADS)
colsum_partial_kernel(const TIn* __restrict__ src, int64_t rows, int C, float* __restrict__ part) {
  __shared__ float red[QM_THREADS * 4];
  const int groups = C >> 2;                               // column groups of 4
  const int rl_n = max(1, QM_THREADS / groups);            // row lanes per CTA (C <= 1024)
  const int cg = threadIdx.x % groups, rl = threadIdx.x / groups;
  const bool active = rl < rl_n && groups <= QM_THREADS;
  float a[4] = {0.f, 0.f, 0.f, 0.f};
  if (active) {
    const int64_t stride = (int64_t)gridDim.x * rl_n;
    int64_t r = (int64_t)blockIdx.x * rl_n + rl;
    for (; r + 3 * stride < rows; r += 4 * stride) {
      float v0[4], v1[4], v2[4], v3[4];
      ld4(src + r * C + cg * 4, v0);
      ld4(src + (r + stride) * C + cg * 4, v1);
      ld4(src + (r + 2 * stride) * C + cg * 4, v2);
      ld4(src + (r + 3 * stride) * C + cg * 4, v3);
#pragma unroll
      for (int e = 0; e < 4; ++e) a[e] += (v0[e] + v1[e]) + (v2[e] + v3[e]);
    }
    for (; r < rows; r += stride) {
      float v[4];
      ld4(src + r * C + cg * 4, v);
#pragma unroll
      for (int e = 0; e < 4; ++e) a[e] += v[e];
    }
  }
#pragma unroll
  for (int e = 0; e < 4; ++e) red[threadIdx.x * 4 + e] = active ? a[e] : 0.f;
  __syncthreads();
  if (threadIdx.x < groups && groups <= QM_THREADS) {
    float t[4] = {0.f, 0.f, 0.f, 0.f};
    for (int l = 0; l < rl_n; ++l)
#pragma unroll
      for (int e = 0; e < 4; ++e) t[e] += red[(l * groups + threadIdx.x) * 4 + e];
    st4(part + (int64_t)blockIdx.x * C + threadIdx.x * 4, t);
  }
}

static int qm_grid(int64_t rows) {
  const int64_t want = cdiv(rows, QM_WARPS);
  return (int)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)sm_count() * 4));
}

static int lb_grid(int64_t rows) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(cdiv(rows, LB_WARPS), (int64_t)sm_count() * 6));
}
int64_t ln_bwd_ws_floats(int64_t rows, int64_t C) { return (int64_t)lb_grid(rows) * 3 * C; }
int64_t colsum_ws_floats(int64_t rows, int64_t C) { return (int64_t)qm_grid(rows) * C; }

template <int NV>
static int32_t ln_fwd_nv(const void* src, int src_dtype, const float* bias, const float* gamma, const float* beta,
                         int64_t rows, float eps, int gelu, void* y, void* pre, float* mean, float* rstd,
                         cudaStream_t st) {
  const int grid = qm_grid(rows);
  auto* yb = (__nv_bfloat16*)y;
  auto* pb = (__nv_bfloat16*)pre;
  if (gelu && src_dtype == ILFDT_F32) {
    ln_fwd_kernel<NV, true, float><<<grid, QM_THREADS, 0, st>>>((const float*)src, bias, gamma, beta, rows, eps, yb, pb,
                                                                 mean, rstd);
  } else if (gelu) {
    ln_fwd_kernel<NV, true, __nv_bfloat16><<<grid, QM_THREADS, 0, st>>>((const __nv_bfloat16*)src, bias, gamma, beta, rows,
                                                                         eps, yb, pb, mean, rstd);
  } else if (src_dtype == ILFDT_F32) {
    ln_fwd_kernel<NV, false, float><<<grid, QM_THREADS, 0, st>>>((const float*)src, nullptr, gamma, beta, rows, eps, yb,
                                                                  nullptr, mean, rstd);
  } else {
    ln_fwd_kernel<NV, false, __nv_bfloat16><<<grid, QM_THREADS, 0, st>>>((const __nv_bfloat16*)src, nullptr, gamma, beta,
                                                                          rows, eps, yb, nullptr, mean, rstd);
  }
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_ln_fwd(const void* src, int src_dtype, const float* bias, const float* gamma, const float* beta,
                      int64_t rows, int64_t C, float eps, int gelu, void* y, void* pre, float* mean, float* rstd,
                      cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  switch (C) {
    case 256: return ln_fwd_nv<2>(src, src_dtype, bias, gamma, beta, rows, eps, gelu, y, pre, mean, rstd, st);
    case 512: return ln_fwd_nv<4>(src, src_dtype, bias, gamma, beta, rows, eps, gelu, y, pre, mean, rstd, st);
    case 768: return ln_fwd_nv<6>(src, src_dtype, bias, gamma, beta, rows, eps, gelu, y, pre, mean, rstd, st);
    case 1024: return ln_fwd_nv<8>(src, src_dtype, bias, gamma, beta, rows, eps, gelu, y, pre, mean, rstd, st);
  }
  return fail(ILFDT_E_UNSUPPORTED_SHAPE, "ln_fwd: width %lld not in {256,512,768,1024}", (long long)C);
}

template <int NV, typename TDy>
static int32_t ln_bwd_nv(const TDy* dy, const void* src, int src_dtype, const float* gamma, const float* mean,
                         const float* rstd, int64_t rows, int gelu, void* dsrc, int dsrc_dtype, float* dgamma,
                         float* dbeta, float* dbias, float* ws, cudaStream_t st) {
  constexpr int C = NV * 128;
  const int grid = lb_grid(rows);
  if (gelu) {
    ILFDT_REQUIRE(src_dtype == ILFDT_BF16 && dsrc_dtype == ILFDT_BF16, ILFDT_E_INVALID_ARG,
                  "ln_bwd: the GELU form reads the bf16 pre-activation and writes a bf16 gradient");
    ln_bwd_kernel<NV, true, __nv_bfloat16, __nv_bfloat16, TDy><<<grid, LB_THREADS, 0, st>>>(
        dy, (const __nv_bfloat16*)src, gamma, mean, rstd, rows, (__nv_bfloat16*)dsrc, ws);
  } else {
    ILFDT_REQUIRE(dsrc_dtype == ILFDT_F32, ILFDT_E_INVALID_ARG, "ln_bwd: the plain form writes an fp32 gradient");
    if (src_dtype == ILFDT_F32)
      ln_bwd_kernel<NV, false, float, float, TDy><<<grid, LB_THREADS, 0, st>>>(dy, (const float*)src, gamma, mean, rstd,
                                                                              rows, (float*)dsrc, ws);
    else
      ln_bwd_kernel<NV, false, __nv_bfloat16, float, TDy><<<grid, LB_THREADS, 0, st>>>(
          dy, (const __nv_bfloat16*)src, gamma, mean, rstd, rows, (float*)dsrc, ws);
  }
  ILFDT_LAUNCH_OK();
  const int tb = 256, gb = (int)cdiv(C, 32);
  partial_reduce_kernel<<<gb, tb, 0, st>>>(ws, grid, 3 * C, C, dgamma);
  partial_reduce_kernel<<<gb, tb, 0, st>>>(ws + C, grid, 3 * C, C, dbeta);
  if (gelu && dbias) partial_reduce_kernel<<<gb, tb, 0, st>>>(ws + 2 * C, grid, 3 * C, C, dbias);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

template <typename TDy>
static int32_t ln_bwd_c(const TDy* dy, const void* src, int src_dtype, const float* gamma, const float* mean,
                        const float* rstd, int64_t rows, int64_t C, int gelu, void* dsrc, int dsrc_dtype, float* dgamma,
                        float* dbeta, float* dbias, float* ws, cudaStream_t st) {
  switch (C) {
    case 256: return ln_bwd_nv<2>(dy, src, src_dtype, gamma, mean, rstd, rows, gelu, dsrc, dsrc_dtype, dgamma, dbeta, dbias, ws, st);
    case 512: return ln_bwd_nv<4>(dy, src, src_dtype, gamma, mean, rstd, rows, gelu, dsrc, dsrc_dtype, dgamma, dbeta, dbias, ws, st);
    case 768: return ln_bwd_nv<6>(dy, src, src_dtype, gamma, mean, rstd, rows, gelu, dsrc, dsrc_dtype, dgamma, dbeta, dbias, ws, st);
    case 1024: return ln_bwd_nv<8>(dy, src, src_dtype, gamma, mean, rstd, rows, gelu, dsrc, dsrc_dtype, dgamma, dbeta, dbias, ws, st);
  }
  return fail(ILFDT_E_UNSUPPORTED_SHAPE, "ln_bwd: width %lld not in {256,512,768,1024}", (long long)C);
}

int32_t launch_ln_bwd(const void* dy, int dy_dtype, const void* src, int src_dtype, const float* gamma, const float* mean,
                      const float* rstd, int64_t rows, int64_t C, int gelu, void* dsrc, int dsrc_dtype, float* dgamma,
                      float* dbeta, float* dbias, float* ws, cudaStream_t st) {
  ILFDT_REQUIRE(rows > 0, ILFDT_E_INVALID_ARG, "ln_bwd: no rows");
  if (dy_dtype == ILFDT_BF16)
    return ln_bwd_c((const __nv_bfloat16*)dy, src, src_dtype, gamma, mean, rstd, rows, C, gelu, dsrc, dsrc_dtype, dgamma,
                    dbeta, dbias, ws, st);
  return ln_bwd_c((const float*)dy, src, src_dtype, gamma, mean, rstd, rows, C, gelu, dsrc, dsrc_dtype, dgamma, dbeta,
                  dbias, ws, st);
}

int32_t launch_bias_cast(const float* src, const float* bias, int64_t rows, int64_t C, void* dst, int dst_dtype,
                         cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  ILFDT_REQUIRE(C % 4 == 0, ILFDT_E_UNSUPPORTED_SHAPE, "bias_cast: width must be a multiple of 4");
  const int64_t n4 = rows * C / 4;
  const int grid = (int)std::min<int64_t>(cdiv(n4, 256), (int64_t)sm_count() * 16);
  if (dst_dtype == ILFDT_BF16)
    bias_cast_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(src, bias, n4, (int)C, (__nv_bfloat16*)dst);
  else
    bias_cast_kernel<float><<<grid, 256, 0, st>>>(src, bias, n4, (int)C, (float*)dst);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_colsum(const void* src, int src_dtype, int64_t rows, int64_t C, float* out, float* ws, cudaStream_t st) {
  ILFDT_REQUIRE(rows > 0 && C % 4 == 0 && C <= 1024, ILFDT_E_UNSUPPORTED_SHAPE,
                "colsum: rows > 0 and width a multiple of 4, <= 1024");
  const int grid = qm_grid(rows);
  if (src_dtype == ILFDT_BF16)
    colsum_partial_kernel<__nv_bfloat16><<<grid, QM_THREADS, 0, st>>>((const __nv_bfloat16*)src, rows, (int)C, ws);
  else
    colsum_partial_kernel<float><<<grid, QM_THREADS, 0, st>>>((const float*)src, rows, (int)C, ws);
  ILFDT_LAUNCH_OK();
  partial_reduce_kernel<<<(int)cdiv(C, 32), 256, 0, st>>>(ws, grid, C, (int)C, out);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// ---- colsum without any allocation (ilfdt_colsum_static; no counterpart in the reference) ---------------------------
// The per-CTA partial sums of launch_colsum live in a file-scope device array instead of a caller workspace: every GPU
// that loads libilfdt holds these COLSUM_STATIC_FLOATS * 4 bytes until the process exits. `g_colsum_static_scratch` is,
// in HOST code, only a stand-in for the symbol: the device address is fetched once per device with
// cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
// Nothing in it has to start at zero: colsum_partial_kernel writes every part[cta][c] the reduce kernel reads.
static __device__ __align__(256) float g_colsum_static_scratch[ILFDT_COLSUM_STATIC_SCRATCH_BYTES / 4];

static int32_t colsum_static_scratch(float** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static float* ptr[MAX_DEV] = {};
  int dev = -1;
  ILFDT_CUDA_OK(cudaGetDevice(&dev));
  ILFDT_REQUIRE(dev >= 0 && dev < MAX_DEV, ILFDT_E_INVALID_ARG, "colsum_static: device ordinal %d outside 0..%d", dev, MAX_DEV - 1);
  std::lock_guard<std::mutex> lock(mu);
  if (ptr[dev] == nullptr) {
    void* p = nullptr;
    ILFDT_CUDA_OK(cudaGetSymbolAddress(&p, g_colsum_static_scratch));
    ILFDT_REQUIRE(p != nullptr, ILFDT_E_CUDA, "colsum_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    ptr[dev] = reinterpret_cast<float*>(p);
  }
  *out = ptr[dev];
  return ILFDT_OK;
}

int32_t launch_colsum_static(const void* src, int src_dtype, int64_t rows, int64_t C, float* out, cudaStream_t st) {
  ILFDT_REQUIRE(rows > 0 && C % 4 == 0 && C <= 1024, ILFDT_E_UNSUPPORTED_SHAPE,
                "colsum_static: rows > 0 and width a multiple of 4, <= 1024");
  const int grid = qm_grid(rows);                        // the same grid as launch_colsum: bit-identical sums
  const int64_t need = (int64_t)grid * C * 4;
  ILFDT_REQUIRE(need <= (int64_t)ILFDT_COLSUM_STATIC_SCRATCH_BYTES, ILFDT_E_INVALID_ARG,
                "colsum_static: %lld rows x %lld columns need %lld bytes of partial sums (%d CTAs), the static scratch "
                "holds %lld: use ilfdt_colsum with a caller workspace",
                (long long)rows, (long long)C, (long long)need, grid, (long long)ILFDT_COLSUM_STATIC_SCRATCH_BYTES);
  float* ws = nullptr;
  if (int32_t rc = colsum_static_scratch(&ws)) return rc;
  if (src_dtype == ILFDT_BF16)
    colsum_partial_kernel<__nv_bfloat16><<<grid, QM_THREADS, 0, st>>>((const __nv_bfloat16*)src, rows, (int)C, ws);
  else
    colsum_partial_kernel<float><<<grid, QM_THREADS, 0, st>>>((const float*)src, rows, (int)C, ws);
  ILFDT_LAUNCH_OK();
  partial_reduce_kernel<<<(int)cdiv(C, 32), 256, 0, st>>>(ws, grid, C, (int)C, out);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

}  // namespace ilfdt
