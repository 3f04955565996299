// a2: Sparsemax forward / backward (prototype/model/sparsemax.py:22-71).
//
// The reference sorts each row (O(K log K) radix sort + cumsum + ~8 elementwise passes). The
// simplex projection only needs the threshold tau with sum(max(0, z - tau)) = 1, so here one CTA
// per row stages the row in shared memory (one HBM read), subtracts the row max (sparsemax.py:40),
// and runs Michelot's fixed-point iteration tau <- (sum_{z>tau} z - 1) / |{z>tau}| with fp64
// block reductions (finite, monotone, converges to the same tau the sorted-prefix rule of
// sparsemax.py:50-59 defines), then writes max(0, z - tau) (one HBM write).
// Algorithmic bytes: 8*K per row (read scores + write weights) -> HBM-bound.
#include "common.cuh"
#include "launch.h"

namespace ilfdt {

constexpr int SPX_THREADS = 512;

struct SpxRed {
  double s[SPX_THREADS / 32];
  int c[SPX_THREADS / 32];
  float f[SPX_THREADS / 32];
};

__device__ __forceinline__ float block_max(float v, SpxRed& r) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) r.f[threadIdx.x >> 5] = v;
  __syncthreads();
  float m = r.f[0];
#pragma unroll
  for (int w = 1; w < SPX_THREADS / 32; ++w) m = fmaxf(m, r.f[w]);
  __syncthreads();
  return m;
}

__device__ __forceinline__ void block_sum_cnt(double& s, int& c, SpxRed& r) {
  s = warp_sum(s);
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0) { r.s[threadIdx.x >> 5] = s; r.c[threadIdx.x >> 5] = c; }
  __syncthreads();
  double ts = 0.0;
  int tc = 0;
#pragma unroll
  for (int w = 0; w < SPX_THREADS / 32; ++w) { ts += r.s[w]; tc += r.c[w]; }
  __syncthreads();
  s = ts;
  c = tc;
}

__global__ void __launch_bounds__(SPX_THREADS)
sparsemax_fwd_kernel(const float* __restrict__ z, int K, float* __restrict__ w,
                     __nv_bfloat16* __restrict__ wb, float* __restrict__ tau_out, int* __restrict__ nnz_out) {
  extern __shared__ float row[];
  __shared__ SpxRed red;
  const int64_t r = blockIdx.x;
  const float* zr = z + r * K;
  const bool vec = (K % 4 == 0) && ((reinterpret_cast<uintptr_t>(zr) & 15) == 0);

  float mx = -INFINITY;
  if (vec) {
    for (int i = threadIdx.x * 4; i < K; i += SPX_THREADS * 4) {
      float4 v = __ldg(reinterpret_cast<const float4*>(zr + i));
      *reinterpret_cast<float4*>(row + i) = v;
      mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    }
  } else {
    for (int i = threadIdx.x; i < K; i += SPX_THREADS) {
      float v = __ldg(zr + i);
      row[i] = v;
      mx = fmaxf(mx, v);
    }
  }
  mx = block_max(mx, red);  // also orders the smem writes (__syncthreads inside)

  // z <- z - max (sparsemax.py:40), in fp32 like the reference; thread-private smem slots.
  for (int i = threadIdx.x; i < K; i += SPX_THREADS) row[i] = row[i] - mx;
  // (each thread only ever touches its own stride of `row` from here on -> no barrier needed)

  // Michelot: start from the whole row active. tau is strictly increasing until the active set
  // is stable; the count is the convergence test.
  double tau = -INFINITY;
  int cnt_prev = -1;
  for (int it = 0; it < 256; ++it) {
    double s = 0.0;
    int c = 0;
    const float tf = (float)tau;  // compare in the precision the output uses
    for (int i = threadIdx.x; i < K; i += SPX_THREADS) {
      float v = row[i];
      if (it == 0 || v > tf) { s += (double)v; ++c; }
    }
    block_sum_cnt(s, c, red);
    if (c == cnt_prev) break;
    cnt_prev = c;
    tau = fmax((s - 1.0) / (double)c, -1.0);   // tau >= max - 1: every weight is <= 1
  }
  const float tf = (float)tau;
  int nz = 0;
  float* wr = w ? w + r * K : nullptr;   // optional: the tuned training path keeps only the bf16 copy
  __nv_bfloat16* wbr = wb ? wb + r * K : nullptr;
  // the subtraction loop wrote stride-1 slots; the vector pass below reads float4 groups.
  __syncthreads();
  if (vec) {
    for (int i = threadIdx.x * 4; i < K; i += SPX_THREADS * 4) {
      float4 v = *reinterpret_cast<const float4*>(row + i);
      float4 o = make_float4(fmaxf(v.x - tf, 0.f), fmaxf(v.y - tf, 0.f), fmaxf(v.z - tf, 0.f),
                             fmaxf(v.w - tf, 0.f));
      nz += (o.x > 0.f) + (o.y > 0.f) + (o.z > 0.f) + (o.w > 0.f);
      if (wr) *reinterpret_cast<float4*>(wr + i) = o;
      if (wbr) {
        __nv_bfloat162 p[2] = {__floats2bfloat162_rn(o.x, o.y), __floats2bfloat162_rn(o.z, o.w)};
        *reinterpret_cast<uint2*>(wbr + i) = *reinterpret_cast<uint2*>(p);
      }
    }
  } else {
    for (int i = threadIdx.x; i < K; i += SPX_THREADS) {
      float o = fmaxf(row[i] - tf, 0.f);
      nz += (o > 0.f);
      if (wr) wr[i] = o;
      if (wbr) wbr[i] = __float2bfloat16_rn(o);
    }
  }
  double dummy = 0.0;
  block_sum_cnt(dummy, nz, red);
  if (threadIdx.x == 0) {
    if (tau_out) tau_out[r] = tf;
    if (nnz_out) nnz_out[r] = nz;
  }
}

// Register-resident variant for K <= 16384 (K % 4 == 0): the row lives in 8 float4 per thread, so the threshold
// iterations cost no shared-memory traffic (ncu r01: the shared-memory kernel ran at 0.22 of the HBM copy rate, bound
// by its ~8 passes over the staged row) and one barrier each (double-buffered partials).
constexpr int SPX_MAXV = 8;

// shared-memory accesses by 32-bit shared address (pointer arithmetic on the integer keeps the .shared state space)
__device__ __forceinline__ uint32_t spx_saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void spx_sts(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ void spx_sts(uint32_t a, int v) { asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ double spx_lds_f64(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ int spx_lds_s32(uint32_t a) { int v; asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }

// (s0, s1) += (m0, m1) * (x0, x1);  (c0, c1) += (m0, m1)   -- one packed fp32x2 FMA and one packed add
__device__ __forceinline__ void masked_acc2(float& s0, float& s1, float& c0, float& c1, float m0, float m1, float x0, float x1) {
  asm("{\n .reg .b64 rm, rx, rs, rc;\n mov.b64 rm, {%4, %5};\n mov.b64 rx, {%6, %7};\n mov.b64 rs, {%0, %1};\n"
      " mov.b64 rc, {%2, %3};\n fma.rn.f32x2 rs, rm, rx, rs;\n add.rn.f32x2 rc, rc, rm;\n mov.b64 {%0, %1}, rs;\n"
      " mov.b64 {%2, %3}, rc;\n}\n"
      : "+f"(s0), "+f"(s1), "+f"(c0), "+f"(c1)
      : "f"(m0), "f"(m1), "f"(x0), "f"(x1));
}

__global__ void __launch_bounds__(SPX_THREADS, 2)
sparsemax_fwd_reg_kernel(const float* __restrict__ z, int K, float* __restrict__ w, __nv_bfloat16* __restrict__ wb,
                         float* __restrict__ tau_out, int* __restrict__ nnz_out) {
  __shared__ __align__(256) double red_s[2][SPX_THREADS / 32];
  __shared__ __align__(128) int red_c[2][SPX_THREADS / 32];
  __shared__ float red_f[SPX_THREADS / 32];
  __shared__ double red_q[SPX_THREADS / 32];
  __shared__ uint32_t red_m[2][SPX_THREADS / 32];
  const int64_t r = blockIdx.x;
  const float* zr = z + r * K;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4 v[SPX_MAXV];
  float mx = -INFINITY;
  // one pass, one barrier: row maximum (exact: it is the shift of sparsemax.py:40) and the raw first two moments of
  // the row. The moments only place the START of the threshold iteration (fp32 per thread, two chains; fp64 across
  // threads): their rounding cannot change the result.
  float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
  for (int j = 0; j < SPX_MAXV; ++j) {
    const int i = (j * SPX_THREADS + threadIdx.x) * 4;
    if (i < K) {
      v[j] = __ldg(reinterpret_cast<const float4*>(zr + i));
      mx = fmaxf(fmaxf(mx, fmaxf(v[j].x, v[j].y)), fmaxf(v[j].z, v[j].w));
      s0 += v[j].x; s1 += v[j].y; s0 += v[j].z; s1 += v[j].w;
      q0 = fmaf(v[j].x, v[j].x, q0); q1 = fmaf(v[j].y, v[j].y, q1);
      q0 = fmaf(v[j].z, v[j].z, q0); q1 = fmaf(v[j].w, v[j].w, q1);
    } else {
      v[j] = make_float4(-1e30f, -1e30f, -1e30f, -1e30f);  // never above any threshold (tau >= -1), never counted
    }
  }
  double ts, tq;
  {
    mx = warp_max(mx);
    const double s = warp_sum((double)s0 + (double)s1), q = warp_sum((double)q0 + (double)q1);
    if (lane == 0) { red_f[warp] = mx; red_s[0][warp] = s; red_q[warp] = q; }
    __syncthreads();
    mx = red_f[lane & 15];
    ts = red_s[0][lane & 15];
    tq = red_q[lane & 15];
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      ts += __shfl_xor_sync(0xffffffffu, ts, o);
      tq += __shfl_xor_sync(0xffffffffu, tq, o);
    }
  }
  // z <- z - max (sparsemax.py:40), in fp32 like the reference
#pragma unroll
  for (int j = 0; j < SPX_MAXV; ++j) { v[j].x -= mx; v[j].y -= mx; v[j].z -= mx; v[j].w -= mx; }
  // Michelot's iteration tau <- (sum_{z > tau} z - 1) / |{z > tau}| is Newton's method on the convex, decreasing
  // f(tau) = sum max(z - tau, 0) - 1: from ANY tau below the row maximum one step lands at or left of the root and the
  // iterates then increase monotonically to it, the support only shrinking. Started from the whole row (tau = mean - 1/K)
  // it needs 6-10 passes at K = 16384; here the start is the root of the same equation for a NORMAL row with this row's
  // mean and variance -- K sigma g(t) = 1, g(t) = phi(t) - t Q(t), ln g(t) ~ -0.919 - 1.207 t - 0.359 t^2 on [-1.7, 3] --
  // and 3-4 passes remain. The final tau is the fixed point of the same update on the same (unique) support: the
  // result does not depend on the start.
  const double kd = (double)K, mu_raw = ts / kd, mu = mu_raw - (double)mx;   // mean of the shifted row
  double tau = fmax(mu - 1.0 / kd, -1.0);            // the whole-row step: always a valid start
  {
    const double var = tq / kd - mu_raw * mu_raw;    // (the variance does not see the shift)
    if (var > 0.0) {
      const float sigma = sqrtf((float)var);
      const float L = -__logf(1.f / ((float)K * sigma)) - 0.919f;
      const float disc = 1.457f + 1.436f * L;
      const float t = disc > 0.f ? (-1.207f + sqrtf(disc)) * (1.f / 0.718f) : -1.68f;
      const double guess = mu + (double)t * (double)sigma;
      // strictly below the row maximum (0 after the shift): the count can never be 0
      tau = fmin(fmax(tau, guess), -1e-30);
    }
  }
  // (the first pass at the model start must always be followed by an update: if the start is right of the root and
  // still below every value, its count is K although tau is not the fixed point. From the second pass on, equal counts
  // mean equal sets: a start right of the root sees a subset A of the support, the next one a superset B of it)
  int cnt_prev = -1;
  // this pass writes buffer 1 first (pass 0 used red_s[0]); the buffers are 128-byte / 64-byte aligned blocks, so
  // "the other buffer" is the same address with one bit flipped
  uint32_t wr_s = spx_saddr(&red_s[1][warp]), rd_s = spx_saddr(&red_s[1][lane & 15]);
  uint32_t wr_c = spx_saddr(&red_c[1][warp]), rd_c = spx_saddr(&red_c[1][lane & 15]);
  for (int it = 1; it < 256; ++it) {
    const float tf = (float)tau;  // compare in the precision the output uses
    // per-thread partial over its <= 32 values in fp32 (two independent chains; |error| ~ 1e-8 of a partial that is
    // itself 1/512 of the row), everything across threads in fp64. Two instructions per value: the mask (1.0 / 0.0,
    // FSET) of a pair of values feeds ONE packed FMA for the two sums and ONE packed add for the two counts
    // (the predicated add + predicated count of the first version were three; the kernel is issue-bound).
    float a0 = 0.f, a1 = 0.f, c0 = 0.f, c1 = 0.f;
#pragma unroll
    for (int j = 0; j < SPX_MAXV; ++j) {
      // padding slots hold a large negative FINITE value: never above a threshold >= -1, and 0 * it stays 0
      masked_acc2(a0, a1, c0, c1, v[j].x > tf ? 1.f : 0.f, v[j].y > tf ? 1.f : 0.f, v[j].x, v[j].y);
      masked_acc2(a0, a1, c0, c1, v[j].z > tf ? 1.f : 0.f, v[j].w > tf ? 1.f : 0.f, v[j].z, v[j].w);
    }
    // From the third pass on (the iterates are left of the root and rising) the pass also finds the SMALLEST value of its
    // support, exactly: for v <= 0 and tf < 0, v > tf  <=>  bits(v) < bits(tf) as unsigned integers, and the smallest such
    // v has the largest bits -- the minimum over all values of (bits(tf) - 1 - bits(v)) as an unsigned integer (values
    // outside the support wrap to huge numbers), one subtraction per value and one three-input minimum per pair.
    // If that value is still above the NEXT threshold the support cannot shrink again: the next pass would only
    // confirm the count (94 % of the rows at random-init statistics stop after 3 passes instead of 4).
    const bool track = it >= 3;
    uint32_t ymin = 0xffffffffu;
    const uint32_t tm1 = __float_as_uint(tf) - 1u;
    if (track) {
#pragma unroll
      for (int j = 0; j < SPX_MAXV; ++j) {
        ymin = __vimin3_u32(ymin, tm1 - __float_as_uint(v[j].x), tm1 - __float_as_uint(v[j].y));
        ymin = __vimin3_u32(ymin, tm1 - __float_as_uint(v[j].z), tm1 - __float_as_uint(v[j].w));
      }
      ymin = __reduce_min_sync(0xffffffffu, ymin);
    }
    const double s = warp_sum((double)a0 + (double)a1);
    const int c = __reduce_add_sync(0xffffffffu, (int)(c0 + c1));
    if (lane == 0) { spx_sts(wr_s, s); spx_sts(wr_c, c); if (track) red_m[it & 1][warp] = ymin; }
    __syncthreads();
    // second stage: the 16 warp partials, one per lane (both half-warps hold the same 16), folded with shuffles --
    // 4 steps instead of a serial sum over 16 shared-memory loads in every thread
    static_assert(SPX_THREADS / 32 == 16, "second-stage reduction folds 16 partials");
    double ws = spx_lds_f64(rd_s);
    int tc = spx_lds_s32(rd_c);
    // (the two partial buffers alternate: pointers swapped by a constant XOR of their shared-memory byte offsets)
    wr_s ^= (uint32_t)sizeof(double) * (SPX_THREADS / 32); rd_s ^= (uint32_t)sizeof(double) * (SPX_THREADS / 32);
    wr_c ^= (uint32_t)sizeof(int) * (SPX_THREADS / 32); rd_c ^= (uint32_t)sizeof(int) * (SPX_THREADS / 32);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) ws += __shfl_xor_sync(0xffffffffu, ws, o);
    tc = __reduce_add_sync(0xffffffffu, tc) >> 1;
    if (tc == cnt_prev) break;
    cnt_prev = tc;
    // every weight is <= 1, so tau >= max - 1 = -1 after the shift. Quotient by a refined reciprocal (the IEEE fp64
    // division is a ~40-instruction subroutine per thread and pass): relative error ~1e-15, tau is then rounded to fp32.
    const double num = ws - 1.0, den = (double)tc;
    float rcp_f;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcp_f) : "f"((float)tc));
    const double rcp = (double)rcp_f;
    double qt = num * rcp;
    qt = fma(fma(-qt, den, num), rcp, qt);
    qt = fma(fma(-qt, den, num), rcp, qt);
    tau = fmax(qt, -1.0);
    if (track) {
      const uint32_t ym = __reduce_min_sync(0xffffffffu, red_m[it & 1][lane & 15]);
      const float vstar = __uint_as_float(tm1 - ym);      // smallest value above tf
      const float tf_next = (float)tau;
      if (tf_next >= tf && vstar > tf_next) break;         // (cnt_prev is already this pass's count)
    }
  }
  const float tf = (float)tau;
  float* wr = w ? w + r * K : nullptr;   // optional: the tuned training path keeps only the bf16 copy
  __nv_bfloat16* wbr = wb ? wb + r * K : nullptr;
#pragma unroll
  for (int j = 0; j < SPX_MAXV; ++j) {
    const int i = (j * SPX_THREADS + threadIdx.x) * 4;
    if (i < K) {
      const float4 o = make_float4(fmaxf(v[j].x - tf, 0.f), fmaxf(v[j].y - tf, 0.f), fmaxf(v[j].z - tf, 0.f),
                                   fmaxf(v[j].w - tf, 0.f));
      if (wr) *reinterpret_cast<float4*>(wr + i) = o;
      if (wbr) {
        __nv_bfloat162 p2[2] = {__floats2bfloat162_rn(o.x, o.y), __floats2bfloat162_rn(o.z, o.w)};
        *reinterpret_cast<uint2*>(wbr + i) = *reinterpret_cast<uint2*>(p2);
      }
    }
  }
  if (threadIdx.x == 0) {
    if (tau_out) tau_out[r] = tf;
    // the support size is the count the iteration stopped on: z - tf > 0 exactly when z > tf
    if (nnz_out) nnz_out[r] = cnt_prev;
  }
}

// dz = nz * (g - mean_nz(g)). One CTA per row; g row cached in registers is not possible for
// arbitrary K, so two passes (second one hits L2/L1).
__global__ void __launch_bounds__(SPX_THREADS)
sparsemax_bwd_kernel(const float* __restrict__ w, const float* __restrict__ g, int K, float* __restrict__ dz) {
  __shared__ SpxRed red;
  const int64_t r = blockIdx.x;
  const float* wr = w + r * K;
  const float* gr = g + r * K;
  double s = 0.0;
  int c = 0;
  for (int i = threadIdx.x; i < K; i += SPX_THREADS)
    if (wr[i] > 0.f) { s += (double)gr[i]; ++c; }
  block_sum_cnt(s, c, red);
  const float mean = c > 0 ? (float)(s / (double)c) : 0.f;
  for (int i = threadIdx.x; i < K; i += SPX_THREADS) dz[r * K + i] = wr[i] > 0.f ? gr[i] - mean : 0.f;
}

int32_t launch_sparsemax_fwd(const float* z, int64_t rows, int64_t K, float* w, void* wb, float* tau, int32_t* nnz,
                             cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  ILFDT_REQUIRE(K >= 1 && K * 4 <= 200 * 1024, ILFDT_E_UNSUPPORTED_SHAPE,
                "sparsemax: K=%lld does not fit one CTA's shared memory (max 51200)", (long long)K);
  ILFDT_REQUIRE(rows < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "sparsemax: too many rows");
  if (K % 4 == 0 && K <= SPX_MAXV * SPX_THREADS * 4 && ((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(w)) & 15) == 0 &&
      (wb == nullptr || (reinterpret_cast<uintptr_t>(wb) & 7) == 0)) {
    sparsemax_fwd_reg_kernel<<<(unsigned)rows, SPX_THREADS, 0, st>>>(z, (int)K, w, reinterpret_cast<__nv_bfloat16*>(wb), tau,
                                                                     nnz);
    ILFDT_LAUNCH_OK();
    return ILFDT_OK;
  }
  size_t smem = (size_t)K * 4;
  static thread_local int configured_dev = -1;
  int dev = 0;
  ILFDT_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    ILFDT_CUDA_OK(cudaFuncSetAttribute(sparsemax_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       200 * 1024));
    configured_dev = dev;
  }
  sparsemax_fwd_kernel<<<(unsigned)rows, SPX_THREADS, smem, st>>>(z, (int)K, w, reinterpret_cast<__nv_bfloat16*>(wb),
                                                                  tau, nnz);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_sparsemax_bwd(const float* w, const float* g, int64_t rows, int64_t K, float* dz, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  sparsemax_bwd_kernel<<<(unsigned)rows, SPX_THREADS, 0, st>>>(w, g, (int)K, dz);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

}  // namespace ilfdt
