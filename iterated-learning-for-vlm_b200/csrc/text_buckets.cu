// a1 for padded text (clip_fdt.py:113-145 with a pad mask): the token rows that cannot change the result are dropped
// BEFORE the score GEMM.
//
// The reference multiplies the scores of pad tokens by 0 (clip_fdt.py:125-130): every pad contributes exactly 0 to the
// max over tokens, whichever pad it is. With the pad rows of the bf16 operand zeroed it is therefore enough to keep, for
// sample s, the rows [0, len_s] -- its valid tokens plus ONE pad (the lowest index among ties wins in the max-pool
// epilogue, and the first pad is the lowest pad, so pooled AND argmax are bit-identical to the full-width kernel).
// Rounded up to the 16-row UMMA granularity that is R = min(16*ceil((len+1)/16), n_tok) rows. CLIP captions are short
// against the context length 77: the benchmark's U{5..77} lengths keep 62 % of the rows.
//
// To feed the statically unrolled tcgen05 kernel (gemm_umma.cu, one TMA box per stage), samples are bucketed by R:
//   text_len_kernel    groups[s] = ceil(R/16)
//   text_plan_kernel   one CTA: stable counting sort of the samples by group count -> for every sample its slot in
//                      bucket order and the first row of its slab in the compacted operand; per bucket {samples,
//                      first row, first slot}; slot -> sample map for the epilogue's stores
//   text_compact_*     q (f32 or bf16) -> compacted bf16 operand, pad rows zeroed (the cast kernel's job, same traffic)
// and umma_kernel<EPI_POOLMAX, NTOK = R> runs once per bucket, reading its bucket's extent from device memory: no host
// synchronisation anywhere. The backward's q-row gathers go through row_base[s] (ground_bwd.cu).
#include "common.cuh"
#include "launch.h"

namespace ilfdt {

constexpr int TB_MAX_BUCKETS = 16;      // n_tok <= 256
constexpr int TB_PLAN_THREADS = 512;

__global__ void text_len_kernel(const float* __restrict__ mask, int b, int n_tok, uint8_t* __restrict__ groups) {
  const int s = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (s >= b) return;
  int last = 0;  // index of the last valid token + 1
  for (int t = lane; t < n_tok; t += 32)
    if (__ldg(mask + (int64_t)s * n_tok + t) == 0.f) last = t + 1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  const int rows = min(n_tok, last + 1);
  if (lane == 0) groups[s] = (uint8_t)((rows + 15) >> 4);
}

// plan layout (ints): [0, 4*TB_MAX_BUCKETS): per bucket g = 1..16 at 4*(g-1): {samples, first row, first slot, rows per sample}
//                     then slot_of[b], row_base[b], sample_of[b]
__global__ void __launch_bounds__(TB_PLAN_THREADS)
text_plan_kernel(const uint8_t* __restrict__ groups, int b, int n_tok, int* __restrict__ plan) {
  __shared__ int cnt[TB_MAX_BUCKETS][TB_PLAN_THREADS + 1];   // thread-private columns, then exclusive prefixes
  __shared__ int bucket_n[TB_MAX_BUCKETS], bucket_slot[TB_MAX_BUCKETS], bucket_row[TB_MAX_BUCKETS];
  int* slot_of = plan + 4 * TB_MAX_BUCKETS;
  int* row_base = slot_of + b;
  int* sample_of = row_base + b;
  const int tid = threadIdx.x;
  const int per = (b + TB_PLAN_THREADS - 1) / TB_PLAN_THREADS;
  const int lo = min(b, tid * per), hi = min(b, lo + per);
  for (int g = 0; g < TB_MAX_BUCKETS; ++g) cnt[g][tid] = 0;
  for (int s = lo; s < hi; ++s) cnt[groups[s] - 1][tid] += 1;
  __syncthreads();
  // per bucket: exclusive scan over the thread columns (warp g handles bucket g: 32 values per lane)
  const int warp = tid >> 5, lane = tid & 31;
  if (warp < TB_MAX_BUCKETS) {
    int loc = 0;
    for (int j = 0; j < TB_PLAN_THREADS / 32; ++j) loc += cnt[warp][lane * (TB_PLAN_THREADS / 32) + j];
    int incl = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - loc;
    for (int j = 0; j < TB_PLAN_THREADS / 32; ++j) {
      const int i = lane * (TB_PLAN_THREADS / 32) + j;
      const int v = cnt[warp][i];
      cnt[warp][i] = run;
      run += v;
    }
    if (lane == 31) bucket_n[warp] = incl;
  }
  __syncthreads();
  if (tid == 0) {
    int slot = 0, row = 0;
    for (int g = 0; g < TB_MAX_BUCKETS; ++g) {
      const int rps = min(16 * (g + 1), n_tok);
      bucket_slot[g] = slot;
      bucket_row[g] = row;
      plan[4 * g + 0] = bucket_n[g];
      plan[4 * g + 1] = row;
      plan[4 * g + 2] = slot;
      plan[4 * g + 3] = rps;
      slot += bucket_n[g];
      row += bucket_n[g] * rps;
    }
  }
  __syncthreads();
  for (int s = lo; s < hi; ++s) {
    const int g = groups[s] - 1;
    const int r = cnt[g][tid]++;
    const int slot = bucket_slot[g] + r;
    slot_of[s] = slot;
    row_base[s] = bucket_row[g] + r * min(16 * (g + 1), n_tok);
    sample_of[slot] = s;
  }
}

// One warp per (sample, token row) of the compacted operand: valid rows are converted, pad rows zeroed.
template <typename T>
__global__ void text_compact_kernel(const T* __restrict__ q, const float* __restrict__ mask,
                                    const uint8_t* __restrict__ groups, const int* __restrict__ row_base, int b, int n_tok,
                                    int D, __nv_bfloat16* __restrict__ out) {
  const int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int s = (int)(wid / n_tok), t = (int)(wid % n_tok);
  if (s >= b) return;
  const int rows = min(16 * (int)groups[s], n_tok);
  if (t >= rows) return;
  const bool keep = __ldg(mask + (int64_t)s * n_tok + t) == 0.f;
  const T* src = q + ((int64_t)s * n_tok + t) * D;
  __nv_bfloat16* dst = out + ((int64_t)row_base[s] + t) * D;
  for (int c = lane * 8; c < D; c += 256) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (keep) {
      if constexpr (sizeof(T) == 2) {
        o = __ldg(reinterpret_cast<const uint4*>(src + c));
      } else {
        const float4 a = __ldg(reinterpret_cast<const float4*>(src + c));
        const float4 bb = __ldg(reinterpret_cast<const float4*>(src + c) + 1);
        __nv_bfloat162 p[4] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w),
                               __floats2bfloat162_rn(bb.x, bb.y), __floats2bfloat162_rn(bb.z, bb.w)};
        o = *reinterpret_cast<uint4*>(p);
      }
    }
    *reinterpret_cast<uint4*>(dst + c) = o;
  }
}

// The same compaction for the fp32-grade score contraction (3-term bf16 split, kernels_simt.cu split3_bf16_kernel, side 1):
// row (row_base[s] + t) of the output holds the six concatenated bf16 copies [x0 x1 x2 x0 x1 x0] of the fp32 source row
// (zeros for a pad), 6 * D wide. One warp per (sample, token row).
__global__ void text_compact_x3_kernel(const float* __restrict__ q, const float* __restrict__ mask,
                                       const uint8_t* __restrict__ groups, const int* __restrict__ row_base, int b,
                                       int n_tok, int D, __nv_bfloat16* __restrict__ out) {
  const int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int s = (int)(wid / n_tok), t = (int)(wid % n_tok);
  if (s >= b) return;
  const int rows = min(16 * (int)groups[s], n_tok);
  if (t >= rows) return;
  const bool keep = __ldg(mask + (int64_t)s * n_tok + t) == 0.f;
  const float* src = q + ((int64_t)s * n_tok + t) * D;
  __nv_bfloat16* dst = out + ((int64_t)row_base[s] + t) * (int64_t)(6 * D);
  for (int c = lane * 4; c < D; c += 128) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (keep) v = __ldg(reinterpret_cast<const float4*>(src + c));
    const float x[4] = {v.x, v.y, v.z, v.w};
    __nv_bfloat16 tm[3][4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat16 h = __float2bfloat16_rn(x[e]);
      const float r1 = x[e] - __bfloat162float(h);
      const __nv_bfloat16 m = __float2bfloat16_rn(r1);
      const float r2 = r1 - __bfloat162float(m);
      tm[0][e] = h; tm[1][e] = m; tm[2][e] = __float2bfloat16_rn(r2);
    }
    const int pat[6] = {0, 1, 2, 0, 1, 0};   // side 1 of split3_bf16_kernel
#pragma unroll
    for (int s6 = 0; s6 < 6; ++s6)
      *reinterpret_cast<uint2*>(dst + (int64_t)s6 * D + c) = *reinterpret_cast<const uint2*>(tm[pat[s6]]);
  }
}

int64_t text_plan_ints(int64_t b) { return 4 * TB_MAX_BUCKETS + 3 * b; }

// groups_ws: b bytes of scratch. plan: text_plan_ints(b) ints. q_out: [b * min(16*ceil(n_tok/16), n_tok) rows... upper bound b*n_tok, D] bf16.
int32_t launch_text_compact(const void* q, int q_dtype, const float* mask, int64_t b, int64_t n_tok, int64_t D,
                            uint8_t* groups_ws, int* plan, void* q_out, cudaStream_t st, int x3) {
  if (b == 0) return ILFDT_OK;
  ILFDT_REQUIRE(n_tok >= 1 && n_tok <= 256 && D % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE,
                "text_compact: n_tok in [1,256], D %% 8 == 0 required");
  ILFDT_REQUIRE(b * n_tok < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "text_compact: b*n_tok too large");
  ILFDT_REQUIRE(((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(q_out)) & 15) == 0, ILFDT_E_INVALID_ARG,
                "text_compact: 16-byte aligned pointers required");
  text_len_kernel<<<(unsigned)cdiv(b * 32, 256), 256, 0, st>>>(mask, (int)b, (int)n_tok, groups_ws);
  ILFDT_LAUNCH_OK();
  text_plan_kernel<<<1, TB_PLAN_THREADS, 0, st>>>(groups_ws, (int)b, (int)n_tok, plan);
  ILFDT_LAUNCH_OK();
  const int64_t warps = b * n_tok;
  const int* row_base = plan + 4 * TB_MAX_BUCKETS + b;
  if (x3) {
    ILFDT_REQUIRE(q_dtype == ILFDT_F32 && D % 4 == 0, ILFDT_E_UNSUPPORTED_SHAPE,
                  "text_compact_x3: fp32 rows with D %% 4 == 0 required");
    text_compact_x3_kernel<<<(unsigned)cdiv(warps * 32, 256), 256, 0, st>>>(
        reinterpret_cast<const float*>(q), mask, groups_ws, row_base, (int)b, (int)n_tok, (int)D,
        reinterpret_cast<__nv_bfloat16*>(q_out));
  } else if (q_dtype == ILFDT_F32)
    text_compact_kernel<float><<<(unsigned)cdiv(warps * 32, 256), 256, 0, st>>>(
        reinterpret_cast<const float*>(q), mask, groups_ws, row_base, (int)b, (int)n_tok, (int)D,
        reinterpret_cast<__nv_bfloat16*>(q_out));
  else
    text_compact_kernel<__nv_bfloat16><<<(unsigned)cdiv(warps * 32, 256), 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(q), mask, groups_ws, row_base, (int)b, (int)n_tok, (int)D,
        reinterpret_cast<__nv_bfloat16*>(q_out));
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

}  // namespace ilfdt
