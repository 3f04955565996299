// a13: backward of the score block (clip_fdt.py:113-149) through token max-pool and sparsemax.
//
// The reference back-propagates a dense [b, n_tok, K] gradient through two full-size matmuls. With
// max-pooling only one token per (sample, code) receives gradient, and only codes in the sparsemax
// support have any, so the gradient is a sparse set of (sample b, code k, token t=arg[b,k]) entries:
//     coef[b,k] = nz[b,k] * (g[b,k] - mean_nz(g[b,:])) * keep[b,t] / (sqrt_d * temperature)
//     dq[b,t,:]  += coef[b,k] * sd[k,:]          (kernel 1: one CTA per sample, gathers sd rows from L2)
//     dsd[k,:]   += coef[b,k] * q[b,t,:]         (kernel 2: one warp per code, gathers q rows)
// Both kernels are deterministic (fixed visiting order, no floating-point atomics unless the sample
// range is split). Work and traffic scale with nnz * D, not n_tok * K * D.
#include "common.cuh"
#include "launch.h"
#include <algorithm>
#include <cstdlib>

namespace ilfdt {

constexpr int GB_THREADS = 512;
constexpr int GB_WARPS = GB_THREADS / 32;
constexpr int GB_CHUNK = 8192;  // codes staged per pass: 8 B * 8192 = 64 KB of shared memory

struct __align__(8) GbEntry {
  uint16_t k;   // code index inside the chunk
  int16_t tok;  // token
  float coef;
};

// acc += c * row, row = 8-element vectors owned by this lane. NV > 0: D == NV*256 (vector path).
template <typename T, int NV>
__device__ __forceinline__ void gb_axpy(const T* __restrict__ row, float c, int lane, float (&acc)[NV][8]) {
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float v[8];
    Load8<T>::ld(row + (i * 32 + lane) * 8, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(c, v[j], acc[i][j]);
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 1: per-sample. Pass 0: mean of g over the support. Per chunk of codes: coef -> written back
// over g, support entries compacted (code order) into shared memory; then each warp owns tokens
// t = warp, warp+16, ... and accumulates its entries' sd rows in registers.
// ---------------------------------------------------------------------------------------------
template <typename T, int NV>
__global__ void __launch_bounds__(GB_THREADS)
ground_bwd_dq_kernel(float* __restrict__ g, const float* __restrict__ w, const int16_t* __restrict__ arg,
                     const T* __restrict__ sd, const float* __restrict__ mask, int n_tok, int K, int D,
                     float inv_scale, float* __restrict__ dq) {
  extern __shared__ __align__(16) unsigned char gb_smem[];
  GbEntry* ent = reinterpret_cast<GbEntry*>(gb_smem);
  __shared__ double red_s[GB_WARPS];
  __shared__ int red_c[GB_WARPS];
  __shared__ int warp_off[GB_WARPS + 1];
  const int64_t b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* gr = g + b * K;
  // w == nullptr: g already holds dz (softmax / sigmoid attention) -- no support, no mean to subtract
  const float* wr = w ? w + b * K : nullptr;
  const int16_t* ar = arg + b * K;
  const float* mr = mask ? mask + b * n_tok : nullptr;

  // pass 0: mean of g over the support
  double s = 0.0;
  int c = 0;
  for (int i = threadIdx.x; i < K; i += GB_THREADS)
    if (wr != nullptr && wr[i] > 0.f) { s += (double)gr[i]; ++c; }
  s = warp_sum(s);
  c = warp_sum(c);
  if (lane == 0) { red_s[warp] = s; red_c[warp] = c; }
  __syncthreads();
  s = 0.0; c = 0;
#pragma unroll
  for (int i = 0; i < GB_WARPS; ++i) { s += red_s[i]; c += red_c[i]; }
  const float mean = c > 0 ? (float)(s / (double)c) : 0.f;

  for (int k0 = 0; k0 < K; k0 += GB_CHUNK) {
    const int kn = min(GB_CHUNK, K - k0);
    const int per_warp = (kn + GB_WARPS - 1) / GB_WARPS;
    const int wb = warp * per_warp, we = min(kn, wb + per_warp);
    // count this warp's surviving entries, computing coef on the way
    int cnt = 0;
    for (int i = wb + lane; i < ((we - wb + 31) / 32) * 32 + wb; i += 32) {
      bool live = false;
      if (i < we) {
        float cf = 0.f;
        if (wr == nullptr || wr[k0 + i] > 0.f) {
          int t = ar[k0 + i];
          float keep = (mr == nullptr || mr[t] == 0.f) ? 1.f : 0.f;
          cf = (gr[k0 + i] - mean) * keep * inv_scale;
        }
        gr[k0 + i] = cf;  // g now holds coef for kernel 2
        live = (cf != 0.f);
      }
      cnt += __popc(__ballot_sync(0xffffffffu, live));
    }
    __syncthreads();  // previous chunk's token loop is done with `ent`
    if (lane == 0) warp_off[warp + 1] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
      warp_off[0] = 0;
      for (int i = 0; i < GB_WARPS; ++i) warp_off[i + 1] += warp_off[i];
    }
    __syncthreads();
    int pos = warp_off[warp];
    const int total = warp_off[GB_WARPS];
    for (int i = wb + lane; i < ((we - wb + 31) / 32) * 32 + wb; i += 32) {
      float cf = (i < we) ? gr[k0 + i] : 0.f;
      bool live = (cf != 0.f);
      unsigned m = __ballot_sync(0xffffffffu, live);
      if (live) {
        GbEntry e;
        e.k = (uint16_t)i;
        e.tok = ar[k0 + i];
        e.coef = cf;
        ent[pos + __popc(m & ((1u << lane) - 1u))] = e;
      }
      pos += __popc(m);
    }
    __syncthreads();

    for (int t = warp; t < n_tok; t += GB_WARPS) {
      float* out = dq + (b * n_tok + t) * (int64_t)D;
      if constexpr (NV > 0) {
        float acc[NV > 0 ? NV : 1][8];
#pragma unroll
        for (int i = 0; i < (NV > 0 ? NV : 1); ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
        for (int base = 0; base < total; base += 32) {
          GbEntry e;
          e.k = 0; e.tok = -1; e.coef = 0.f;
          if (base + lane < total) e = ent[base + lane];
          unsigned m = __ballot_sync(0xffffffffu, e.tok == t);
          while (m) {
            int j0 = __ffs(m) - 1;
            m &= m - 1;
            int kk0 = __shfl_sync(0xffffffffu, (int)e.k, j0);
            float c0 = __shfl_sync(0xffffffffu, e.coef, j0);
            if (m) {  // two independent row gathers in flight
              int j1 = __ffs(m) - 1;
              m &= m - 1;
              int kk1 = __shfl_sync(0xffffffffu, (int)e.k, j1);
              float c1 = __shfl_sync(0xffffffffu, e.coef, j1);
              float v0[NV > 0 ? NV : 1][8], v1[NV > 0 ? NV : 1][8];
#pragma unroll
              for (int i = 0; i < (NV > 0 ? NV : 1); ++i) {
                Load8<T>::ld(sd + (int64_t)(k0 + kk0) * D + (i * 32 + lane) * 8, v0[i]);
                Load8<T>::ld(sd + (int64_t)(k0 + kk1) * D + (i * 32 + lane) * 8, v1[i]);
              }
#pragma unroll
              for (int i = 0; i < (NV > 0 ? NV : 1); ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(c1, v1[i][j], fmaf(c0, v0[i][j], acc[i][j]));
            } else {
              gb_axpy<T, (NV > 0 ? NV : 1)>(sd + (int64_t)(k0 + kk0) * D, c0, lane, acc);
            }
          }
        }
#pragma unroll
        for (int i = 0; i < (NV > 0 ? NV : 1); ++i) {
          float* o = out + (i * 32 + lane) * 8;
          float4 a = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
          float4 bb = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
          if (k0 > 0) {
            float4 pa = *reinterpret_cast<float4*>(o), pb = *reinterpret_cast<float4*>(o + 4);
            a.x += pa.x; a.y += pa.y; a.z += pa.z; a.w += pa.w;
            bb.x += pb.x; bb.y += pb.y; bb.z += pb.z; bb.w += pb.w;
          }
          *reinterpret_cast<float4*>(o) = a;
          *reinterpret_cast<float4*>(o + 4) = bb;
        }
      } else {
        // generic D (tests, odd widths): lane-strided scalar path, D <= 1024
        float acc[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = 0.f;
        for (int base = 0; base < total; base += 32) {
          GbEntry e;
          e.k = 0; e.tok = -1; e.coef = 0.f;
          if (base + lane < total) e = ent[base + lane];
          unsigned m = __ballot_sync(0xffffffffu, e.tok == t);
          while (m) {
            int j0 = __ffs(m) - 1;
            m &= m - 1;
            int kk0 = __shfl_sync(0xffffffffu, (int)e.k, j0);
            float c0 = __shfl_sync(0xffffffffu, e.coef, j0);
            const T* row = sd + (int64_t)(k0 + kk0) * D;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j * 32 + lane < D) acc[j] = fmaf(c0, (float)row[j * 32 + lane], acc[j]);
          }
        }
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j * 32 + lane < D) out[j * 32 + lane] = (k0 > 0 ? out[j * 32 + lane] : 0.f) + acc[j];
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 2: per-code. One warp owns code k and sweeps the samples of its range 32 at a time
// (lane = sample): coef (left in g by kernel 1) and arg are read at stride K -- the 8 warps of a
// CTA own 8 consecutive codes, i.e. one 32-byte sector per sample row -- then the q rows of the
// non-zero entries are gathered and accumulated in registers. dsd[k,:] += acc.
// ---------------------------------------------------------------------------------------------
constexpr int GS_WARPS = 8;

template <typename T, int NV>
__global__ void __launch_bounds__(GS_WARPS * 32)
ground_bwd_dsd_kernel(const float* __restrict__ coef, const int16_t* __restrict__ arg, const T* __restrict__ q,
                      int64_t b_total, int n_tok, int K, int D, int64_t b_per_split, int use_atomic,
                      float* __restrict__ dsd) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * GS_WARPS + warp;
  if (k >= K) return;
  const int64_t b_begin = (int64_t)blockIdx.y * b_per_split;
  const int64_t b_end = min(b_total, b_begin + b_per_split);
  float* out = dsd + (int64_t)k * D;

  if constexpr (NV > 0) {
    float acc[NV > 0 ? NV : 1][8];
#pragma unroll
    for (int i = 0; i < (NV > 0 ? NV : 1); ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int64_t b0 = b_begin; b0 < b_end; b0 += 32) {
      const int64_t bb = b0 + lane;
      float cf = 0.f;
      int t = 0;
      if (bb < b_end) {
        cf = __ldg(coef + bb * K + k);
        t = __ldg(arg + bb * K + k);
      }
      unsigned m = __ballot_sync(0xffffffffu, cf != 0.f);
      while (m) {
        int j0 = __ffs(m) - 1;
        m &= m - 1;
        int t0 = __shfl_sync(0xffffffffu, t, j0);
        float c0 = __shfl_sync(0xffffffffu, cf, j0);
        const T* r0 = q + ((b0 + j0) * n_tok + t0) * (int64_t)D;
        if (m) {
          int j1 = __ffs(m) - 1;
          m &= m - 1;
          int t1 = __shfl_sync(0xffffffffu, t, j1);
          float c1 = __shfl_sync(0xffffffffu, cf, j1);
          const T* r1 = q + ((b0 + j1) * n_tok + t1) * (int64_t)D;
          float v0[NV > 0 ? NV : 1][8], v1[NV > 0 ? NV : 1][8];
#pragma unroll
          for (int i = 0; i < (NV > 0 ? NV : 1); ++i) {
            Load8<T>::ld(r0 + (i * 32 + lane) * 8, v0[i]);
            Load8<T>::ld(r1 + (i * 32 + lane) * 8, v1[i]);
          }
#pragma unroll
          for (int i = 0; i < (NV > 0 ? NV : 1); ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(c1, v1[i][j], fmaf(c0, v0[i][j], acc[i][j]));
        } else {
          gb_axpy<T, (NV > 0 ? NV : 1)>(r0, c0, lane, acc);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < (NV > 0 ? NV : 1); ++i) {
      float* o = out + (i * 32 + lane) * 8;
      if (use_atomic) {
#pragma unroll
        for (int j = 0; j < 8; ++j) atomicAdd(o + j, acc[i][j]);
      } else {
        float4 pa = *reinterpret_cast<float4*>(o), pb = *reinterpret_cast<float4*>(o + 4);
        pa.x += acc[i][0]; pa.y += acc[i][1]; pa.z += acc[i][2]; pa.w += acc[i][3];
        pb.x += acc[i][4]; pb.y += acc[i][5]; pb.z += acc[i][6]; pb.w += acc[i][7];
        *reinterpret_cast<float4*>(o) = pa;
        *reinterpret_cast<float4*>(o + 4) = pb;
      }
    }
  } else {
    float acc[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) acc[j] = 0.f;
    for (int64_t b0 = b_begin; b0 < b_end; b0 += 32) {
      const int64_t bb = b0 + lane;
      float cf = 0.f;
      int t = 0;
      if (bb < b_end) {
        cf = __ldg(coef + bb * K + k);
        t = __ldg(arg + bb * K + k);
      }
      unsigned m = __ballot_sync(0xffffffffu, cf != 0.f);
      while (m) {
        int j0 = __ffs(m) - 1;
        m &= m - 1;
        int t0 = __shfl_sync(0xffffffffu, t, j0);
        float c0 = __shfl_sync(0xffffffffu, cf, j0);
        const T* r0 = q + ((b0 + j0) * n_tok + t0) * (int64_t)D;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j * 32 + lane < D) acc[j] = fmaf(c0, (float)r0[j * 32 + lane], acc[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j * 32 + lane < D) {
        if (use_atomic) atomicAdd(out + j * 32 + lane, acc[j]);
        else out[j * 32 + lane] += acc[j];
      }
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 1, v3: per-sample CTA of 256 threads; the preprocessing that dominated the sorted kernel above
// (ncu r01: ~60 % of its stall samples sat in the scalar, dependent-load coef / histogram / placement
// passes) is rebuilt around coalesced 16-byte loads and thread-private counters:
//   pass 0  mean of g over the support (float4 loads, fp64 accumulation)
//   pass A  each thread owns 64 codes of the 16384-code chunk as 16 coalesced float4 groups: coef is
//           computed, written back over g (for kernel 2) and the thread's per-token counts go to its
//           private column of hist[token][thread] (uint16, no atomics, no match_any)
//   scan    per token over the 256 thread columns, then over tokens -> bucket bases
//   pass B  the thread re-walks its (register-resident) live mask / tokens and drops each code index at
//           base[token] + its running column offset: a counting sort whose order (thread, group, element)
//           is fixed, so the summation order and the result are deterministic
//   gather  warps take tokens dynamically; a bucket is read 32 entries at a time (code index from shared
//           memory, coef from the L2-resident row just written), 4 codebook rows in flight per warp.
// ---------------------------------------------------------------------------------------------
constexpr int G3_THREADS = 256, G3_WARPS = G3_THREADS / 32, G3_PER = 64, G3_CHUNK = G3_THREADS * G3_PER;  // 16384
constexpr int G3_GROUPS = G3_PER / 4;
constexpr int G3_MAX_TOK = 96;

// L2 eviction policy for the gathered rows: the rows that are reused by every CTA (the codebook in kernel 1, the current
// slice of q in kernel 2) are loaded evict-last so that the streamed per-sample arrays (g / w / arg, coef) do not push
// them out of L2.
__device__ __forceinline__ uint64_t l2_policy(bool keep) {
  uint64_t pol;
  if (keep) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint4 ldg_hint(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
  return r;
}

template <typename TO>
struct StoreRow8;
template <>
struct StoreRow8<float> {
  __device__ __forceinline__ static void st(float* o, const float (&a)[8], bool add) {
    float4 x = make_float4(a[0], a[1], a[2], a[3]), y = make_float4(a[4], a[5], a[6], a[7]);
    if (add) {
      const float4 px = *reinterpret_cast<float4*>(o), py = *reinterpret_cast<float4*>(o + 4);
      x.x += px.x; x.y += px.y; x.z += px.z; x.w += px.w;
      y.x += py.x; y.y += py.y; y.z += py.z; y.w += py.w;
    }
    *reinterpret_cast<float4*>(o) = x;
    *reinterpret_cast<float4*>(o + 4) = y;
  }
};
template <>
struct StoreRow8<__nv_bfloat16> {
  __device__ __forceinline__ static void st(__nv_bfloat16* o, const float (&a)[8], bool add) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = a[j];
    if (add) {
      float p[8];
      Load8<__nv_bfloat16>::ld(o, p);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += p[j];
    }
    uint4 r;
    __nv_bfloat162 b0 = __floats2bfloat162_rn(v[0], v[1]), b1 = __floats2bfloat162_rn(v[2], v[3]);
    __nv_bfloat162 b2 = __floats2bfloat162_rn(v[4], v[5]), b3 = __floats2bfloat162_rn(v[6], v[7]);
    r.x = *reinterpret_cast<uint32_t*>(&b0); r.y = *reinterpret_cast<uint32_t*>(&b1);
    r.z = *reinterpret_cast<uint32_t*>(&b2); r.w = *reinterpret_cast<uint32_t*>(&b3);
    *reinterpret_cast<uint4*>(o) = r;
  }
};

// support bits of four consecutive sparsemax weights (bit i: w[i] > 0). TW = float: the fp32 att_weight;
// TW = bf16: the tensor-core operand copy the forward keeps anyway (weights are >= 0 and bf16 has fp32's
// exponent range, so w > 0 survives the rounding) -- half the bytes, and the fp32 copy need not exist.
template <typename TW> struct Support4;
template <> struct Support4<float> {
  static __device__ __forceinline__ uint32_t ld(const float* p) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    return (v.x > 0.f ? 1u : 0u) | (v.y > 0.f ? 2u : 0u) | (v.z > 0.f ? 4u : 0u) | (v.w > 0.f ? 8u : 0u);
  }
};
template <> struct Support4<__nv_bfloat16> {
  static __device__ __forceinline__ uint32_t pos(uint32_t h) { return (h & 0x8000u) == 0u && (h & 0x7fffu) != 0u; }
  static __device__ __forceinline__ uint32_t ld(const __nv_bfloat16* p) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    return pos(v.x & 0xffffu) | (pos(v.x >> 16) << 1) | (pos(v.y & 0xffffu) << 2) | (pos(v.y >> 16) << 3);
  }
};

__device__ __forceinline__ bool weight_pos(float v) { return v > 0.f; }
__device__ __forceinline__ bool weight_pos(__nv_bfloat16 v) { return __bfloat162float(v) > 0.f; }

// ---------------------------------------------------------------------------------------------
// kernel 1, token-sorted: same per-sample structure, but the support entries of a chunk are bucketed by
// token with a stable counting sort (per-warp histograms over contiguous code ranges, __match_any_sync
// for the rank inside a 32-entry group), so a token's warp walks only its own entries instead of
// scanning the whole list (n_tok x nnz/32 ballots, a third of the old kernel's instructions, ncu r01).
// Order inside a bucket is code order: the summation order, and therefore the result, is deterministic.
// ---------------------------------------------------------------------------------------------
struct __align__(8) GbEnt2 {
  int k;
  float coef;
};

// TW / TO as in the v3 kernel: the support may be read from the bf16 copy of the weights and dq written in bf16
// (token counts above G3_MAX_TOK -- ViT-B/16's 196 patches, BASELINE.json config 5 -- take this kernel).
template <typename T, int NV, typename TO, typename TW>
__global__ void __launch_bounds__(GB_THREADS)
ground_bwd_dq_sorted_kernel(float* __restrict__ g, const TW* __restrict__ w, const int16_t* __restrict__ arg,
                            const T* __restrict__ sd, const float* __restrict__ mask, int n_tok, int K, int D,
                            float inv_scale, TO* __restrict__ dq) {
  extern __shared__ __align__(16) unsigned char gb_smem[];
  GbEnt2* ent = reinterpret_cast<GbEnt2*>(gb_smem);
  int* hist = reinterpret_cast<int*>(gb_smem + GB_CHUNK * sizeof(GbEnt2));  // [GB_WARPS][n_tok]
  int* base = hist + GB_WARPS * n_tok;                                      // [n_tok + 1]
  __shared__ double red_s[GB_WARPS];
  __shared__ int red_c[GB_WARPS];
  __shared__ int next_tok;
  const int64_t b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* gr = g + b * K;
  const TW* wr = w ? w + b * K : nullptr;
  const int16_t* ar = arg + b * K;
  const float* mr = mask ? mask + b * n_tok : nullptr;
  int* myhist = hist + warp * n_tok;

  double s = 0.0;
  int c = 0;
  for (int i = threadIdx.x; i < K; i += GB_THREADS)
    if (wr != nullptr && weight_pos(wr[i])) { s += (double)gr[i]; ++c; }
  s = warp_sum(s);
  c = warp_sum(c);
  if (lane == 0) { red_s[warp] = s; red_c[warp] = c; }
  __syncthreads();
  s = 0.0; c = 0;
#pragma unroll
  for (int i = 0; i < GB_WARPS; ++i) { s += red_s[i]; c += red_c[i]; }
  const float mean = c > 0 ? (float)(s / (double)c) : 0.f;

  for (int k0 = 0; k0 < K; k0 += GB_CHUNK) {
    const int kn = min(GB_CHUNK, K - k0);
    const int per_warp = (kn + GB_WARPS - 1) / GB_WARPS;
    const int wb = warp * per_warp, we = min(kn, wb + per_warp);
    const int wend = ((max(we - wb, 0) + 31) / 32) * 32 + wb;
    for (int t = lane; t < n_tok; t += 32) myhist[t] = 0;
    __syncwarp();
    // pass 1: coef (left in g for kernel 2) + per-warp token histogram
    for (int i = wb + lane; i < wend; i += 32) {
      float cf = 0.f;
      int t = 0;
      if (i < we) {
        if (wr == nullptr || weight_pos(wr[k0 + i])) {
          t = ar[k0 + i];
          const float keep = (mr == nullptr || mr[t] == 0.f) ? 1.f : 0.f;
          cf = (gr[k0 + i] - mean) * keep * inv_scale;
        }
        gr[k0 + i] = cf;
      }
      const bool live = cf != 0.f;
      const unsigned m = __match_any_sync(0xffffffffu, live ? t : (0x10000 + lane));
      if (live && lane == __ffs(m) - 1) myhist[t] += __popc(m);
      __syncwarp();
    }
    __syncthreads();  // histograms complete; every warp is done with the previous chunk's entries
    // per token: exclusive prefix over warps (in place), total -> base[t + 1]
    for (int t = threadIdx.x; t < n_tok; t += GB_THREADS) {
      int run = 0;
      for (int wq = 0; wq < GB_WARPS; ++wq) {
        const int v = hist[wq * n_tok + t];
        hist[wq * n_tok + t] = run;
        run += v;
      }
      base[t + 1] = run;
    }
    __syncthreads();
    if (warp == 0) {  // inclusive scan of base[1..n_tok] (n_tok <= 256: 8 per lane)
      const int per = (n_tok + 31) / 32;
      int loc = 0;
      for (int j = 0; j < per; ++j) {
        const int t = lane * per + j;
        if (t < n_tok) loc += base[t + 1];
      }
      int incl = loc;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
      }
      int run = incl - loc;
      for (int j = 0; j < per; ++j) {
        const int t = lane * per + j;
        if (t < n_tok) { run += base[t + 1]; base[t + 1] = run; }
      }
      if (lane == 0) { base[0] = 0; next_tok = 0; }
    }
    __syncthreads();
    // pass 2: stable placement
    for (int i = wb + lane; i < wend; i += 32) {
      const float cf = (i < we) ? gr[k0 + i] : 0.f;
      const bool live = cf != 0.f;
      const int t = live ? (int)ar[k0 + i] : 0;
      const unsigned m = __match_any_sync(0xffffffffu, live ? t : (0x10000 + lane));
      if (live) {
        GbEnt2 e;
        e.k = k0 + i;
        e.coef = cf;
        ent[base[t] + myhist[t] + __popc(m & ((1u << lane) - 1u))] = e;
      }
      __syncwarp();
      if (live && lane == __ffs(m) - 1) myhist[t] += __popc(m);
      __syncwarp();
    }
    __syncthreads();

    for (;;) {  // tokens are handed out dynamically: entry counts per token are uneven
      int t = 0;
      if (lane == 0) t = atomicAdd(&next_tok, 1);
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t >= n_tok) break;
      const int lo = base[t], hi = base[t + 1];
      float acc[NV][8];
#pragma unroll
      for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
      int e = lo;
      for (; e + 1 < hi; e += 2) {
        const GbEnt2 e0 = ent[e], e1 = ent[e + 1];
        float v0[NV][8], v1[NV][8];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          Load8<T>::ld(sd + (int64_t)e0.k * D + (i * 32 + lane) * 8, v0[i]);
          Load8<T>::ld(sd + (int64_t)e1.k * D + (i * 32 + lane) * 8, v1[i]);
        }
#pragma unroll
        for (int i = 0; i < NV; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(e1.coef, v1[i][j], fmaf(e0.coef, v0[i][j], acc[i][j]));
      }
      if (e < hi) {
        const GbEnt2 e0 = ent[e];
        gb_axpy<T, NV>(sd + (int64_t)e0.k * D, e0.coef, lane, acc);
      }
      TO* out = dq + (b * n_tok + t) * (int64_t)D;
#pragma unroll
      for (int i = 0; i < NV; ++i) StoreRow8<TO>::st(out + (i * 32 + lane) * 8, acc[i], k0 > 0);
    }
  }
}


// PG > 0 (bf16 rows): the gather loop is software-pipelined like kernel 2 v4 -- groups of PG codebook rows, the loads of
// group i+1 issued before group i is consumed, the next 32-entry chunk's indices and coefficients prefetched behind the
// current chunk's gathers -- so the per-batch L2 latency is no longer exposed (see ground_bwd_dsd_v4_kernel).
// NB = codebook rows a warp keeps in flight in the plain (PG == 0) gather loop: 4 at 3 CTAs per SM; 8 (MINB = 2) for
// launches with fewer samples than resident CTA slots, where a sample's 3000 gathers are a latency chain on an SM that
// holds one or two CTAs (ncu launch list of a 256-row step: 95 us per dq launch at 8.5 TB/s).
template <typename T, int NV, typename TO, typename TW, int PG = 0, int MINB = 3, int NB = 4>
__global__ void __launch_bounds__(G3_THREADS, MINB)
ground_bwd_dq_v3_kernel(float* __restrict__ g, const TW* __restrict__ w, const int16_t* __restrict__ arg,
                        const T* __restrict__ sd, const float* __restrict__ mask, int n_tok, int K, int D,
                        float inv_scale, TO* __restrict__ dq, const int* __restrict__ sparse_flag) {
  extern __shared__ __align__(16) unsigned char gb_smem[];
  if (sparse_flag != nullptr && *sparse_flag != 0) return;   // the sparse-regime kernel handles this launch
  const uint64_t row_pol = l2_policy(false);
  uint16_t* hist = reinterpret_cast<uint16_t*>(gb_smem);                                    // [n_tok][256]
  uint16_t* sorted_k = reinterpret_cast<uint16_t*>(gb_smem + (size_t)n_tok * G3_THREADS * 2);  // [G3_CHUNK]
  int* base = reinterpret_cast<int*>(sorted_k + G3_CHUNK);                                  // [n_tok + 1]
  float* keep_s = reinterpret_cast<float*>(base + G3_MAX_TOK + 1);                          // [n_tok]
  __shared__ double red_s[G3_WARPS];
  __shared__ int red_c[G3_WARPS];
  __shared__ int next_tok;
  const int64_t b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float* gr = g + b * K;
  const TW* wr = w ? w + b * K : nullptr;
  const int16_t* ar = arg + b * K;

  for (int t = tid; t < n_tok; t += G3_THREADS) keep_s[t] = (mask == nullptr || mask[b * n_tok + t] == 0.f) ? 1.f : 0.f;

  // pass 0: mean of g over the support
  double s = 0.0;
  int c = 0;
  if (wr != nullptr) {
    for (int i = tid * 4; i < K; i += G3_THREADS * 4) {
      const uint32_t m = Support4<TW>::ld(wr + i);
      const float4 gv = *reinterpret_cast<const float4*>(gr + i);
      if (m & 1u) { s += (double)gv.x; ++c; }
      if (m & 2u) { s += (double)gv.y; ++c; }
      if (m & 4u) { s += (double)gv.z; ++c; }
      if (m & 8u) { s += (double)gv.w; ++c; }
    }
  }
  s = warp_sum(s);
  c = warp_sum(c);
  if (lane == 0) { red_s[warp] = s; red_c[warp] = c; }
  __syncthreads();
  s = 0.0; c = 0;
#pragma unroll
  for (int i = 0; i < G3_WARPS; ++i) { s += red_s[i]; c += red_c[i]; }
  const float mean = c > 0 ? (float)(s / (double)c) : 0.f;

  for (int k0 = 0; k0 < K; k0 += G3_CHUNK) {
    // zero the counters (the previous chunk's gather phase is over: barrier at the loop tail)
    {
      uint4* h4 = reinterpret_cast<uint4*>(hist);
      const int n16 = n_tok * G3_THREADS * 2 / 16;
      for (int i = tid; i < n16; i += G3_THREADS) h4[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    // pass A
    uint32_t live[G3_GROUPS / 8];   // 4 bits per group
    uint32_t tk[G3_GROUPS];         // 4 tokens per group, one byte each
#pragma unroll
    for (int i = 0; i < G3_GROUPS / 8; ++i) live[i] = 0u;
#pragma unroll
    for (int i = 0; i < G3_GROUPS; ++i) {
      const int kk = (i * G3_THREADS + tid) * 4;
      const int k = k0 + kk;
      uint32_t lv = 0u, tkp = 0u;
      if (k < K) {
        uint32_t m = 0xfu;
        if (wr != nullptr) m = Support4<TW>::ld(wr + k);
        const float4 gv = *reinterpret_cast<const float4*>(gr + k);
        const uint2 av = __ldg(reinterpret_cast<const uint2*>(ar + k));
        const int t0 = av.x & 0xffff, t1 = av.x >> 16, t2 = av.y & 0xffff, t3 = av.y >> 16;
        float4 cf;
        cf.x = (m & 1u) ? (gv.x - mean) * keep_s[t0] * inv_scale : 0.f;
        cf.y = (m & 2u) ? (gv.y - mean) * keep_s[t1] * inv_scale : 0.f;
        cf.z = (m & 4u) ? (gv.z - mean) * keep_s[t2] * inv_scale : 0.f;
        cf.w = (m & 8u) ? (gv.w - mean) * keep_s[t3] * inv_scale : 0.f;
        *reinterpret_cast<float4*>(gr + k) = cf;   // g now holds coef for kernel 2
        tkp = (uint32_t)t0 | ((uint32_t)t1 << 8) | ((uint32_t)t2 << 16) | ((uint32_t)t3 << 24);
        if (cf.x != 0.f) { lv |= 1u; hist[t0 * G3_THREADS + tid] += 1; }
        if (cf.y != 0.f) { lv |= 2u; hist[t1 * G3_THREADS + tid] += 1; }
        if (cf.z != 0.f) { lv |= 4u; hist[t2 * G3_THREADS + tid] += 1; }
        if (cf.w != 0.f) { lv |= 8u; hist[t3 * G3_THREADS + tid] += 1; }
      }
      tk[i] = tkp;
      live[i >> 3] |= lv << ((i & 7) * 4);
    }
    __syncthreads();
    // scan over the thread columns of each token: hist[t][*] -> exclusive prefix, total -> base[t + 1]
    for (int t = warp; t < n_tok; t += G3_WARPS) {
      uint4 v = *reinterpret_cast<uint4*>(hist + t * G3_THREADS + lane * 8);
      uint32_t x[8] = {v.x & 0xffff, v.x >> 16, v.y & 0xffff, v.y >> 16, v.z & 0xffff, v.z >> 16, v.w & 0xffff, v.w >> 16};
      uint32_t run = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { const uint32_t y = x[j]; x[j] = run; run += y; }
      uint32_t incl = run;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
      }
      const uint32_t off = incl - run;
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] += off;
      v.x = x[0] | (x[1] << 16); v.y = x[2] | (x[3] << 16); v.z = x[4] | (x[5] << 16); v.w = x[6] | (x[7] << 16);
      *reinterpret_cast<uint4*>(hist + t * G3_THREADS + lane * 8) = v;
      if (lane == 31) base[t + 1] = (int)incl;
    }
    __syncthreads();
    if (warp == 0) {  // inclusive scan of base[1..n_tok] (n_tok <= 96: 3 per lane)
      int loc[3], sum = 0;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int t = lane * 3 + j;
        loc[j] = t < n_tok ? base[t + 1] : 0;
        sum += loc[j];
      }
      int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
      }
      int run = incl - sum;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int t = lane * 3 + j;
        run += loc[j];
        if (t < n_tok) base[t + 1] = run;
      }
      if (lane == 0) { base[0] = 0; next_tok = 0; }
    }
    __syncthreads();
    // pass B: placement
#pragma unroll
    for (int i = 0; i < G3_GROUPS; ++i) {
      const uint32_t lv = (live[i >> 3] >> ((i & 7) * 4)) & 15u;
      const int kk = (i * G3_THREADS + tid) * 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (lv & (1u << j)) {
          const int t = (tk[i] >> (8 * j)) & 255;
          const int h = hist[t * G3_THREADS + tid];
          hist[t * G3_THREADS + tid] = (uint16_t)(h + 1);
          sorted_k[base[t] + h] = (uint16_t)(kk + j);
        }
      }
    }
    __syncthreads();
    // gather
    const float* coef_row = gr + k0;
    for (;;) {
      int t = 0;
      if (lane == 0) t = atomicAdd(&next_tok, 1);
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t >= n_tok) break;
      const int lo = base[t], hi = base[t + 1];
      float acc[NV][8];
#pragma unroll
      for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
      if constexpr (PG > 0 && sizeof(T) == 2) {
        uint4 bufA[PG][NV], bufB[PG][NV];
        float cfA[PG], cfB[PG];
        int nA = 0, nB = 0, phase = 0;
        bool pend = false;
        int myk = 0;
        float myc = 0.f;
        if (lo + lane < hi) {
          myk = sorted_k[lo + lane];
          myc = coef_row[myk];
        }
        auto issue = [&](uint4 (&buf)[PG][NV], float (&cf)[PG], int& n, int j, int cnt) {
          n = cnt;
#pragma unroll
          for (int u = 0; u < PG; ++u) {
            const int kq = __shfl_sync(0xffffffffu, myk, (j + u) & 31);
            cf[u] = __shfl_sync(0xffffffffu, myc, (j + u) & 31);
            if (u < cnt) {
              const T* row = sd + (int64_t)(k0 + kq) * D;
#pragma unroll
              for (int i = 0; i < NV; ++i) buf[u][i] = ldg_hint(row + (i * 32 + lane) * 8, row_pol);
            }
          }
        };
        auto consume = [&](const uint4 (&buf)[PG][NV], const float (&cf)[PG], int n) {
#pragma unroll
          for (int u = 0; u < PG; ++u) {
            if (u < n) {
#pragma unroll
              for (int i = 0; i < NV; ++i) {
                const uint32_t ww[4] = {buf[u][i].x, buf[u][i].y, buf[u][i].z, buf[u][i].w};
#pragma unroll
                for (int q2 = 0; q2 < 4; ++q2) {
                  acc[i][2 * q2] = fmaf(cf[u], __uint_as_float(ww[q2] << 16), acc[i][2 * q2]);
                  acc[i][2 * q2 + 1] = fmaf(cf[u], __uint_as_float(ww[q2] & 0xffff0000u), acc[i][2 * q2 + 1]);
                }
              }
            }
          }
        };
        for (int e0 = lo; e0 < hi; e0 += 32) {
          int nk = 0;
          float nc = 0.f;
          if (e0 + 32 + lane < hi) {     // next chunk's indices + coefficients: in flight behind this chunk's gathers
            nk = sorted_k[e0 + 32 + lane];
            nc = coef_row[nk];
          }
          const int n = min(32, hi - e0);
          for (int j = 0; j < n; j += PG) {
            const int cnt = min(PG, n - j);
            if (phase == 0) {
              issue(bufA, cfA, nA, j, cnt);
              if (pend) consume(bufB, cfB, nB);
              phase = 1;
            } else {
              issue(bufB, cfB, nB, j, cnt);
              if (pend) consume(bufA, cfA, nA);
              phase = 0;
            }
            pend = true;
          }
          myk = nk;
          myc = nc;
        }
        if (pend) {
          if (phase == 1) consume(bufA, cfA, nA);
          else consume(bufB, cfB, nB);
        }
      } else {
      int nxk = 0;
      float nxc = 0.f;
      if (lo + lane < hi) {
        nxk = sorted_k[lo + lane];
        nxc = coef_row[nxk];
      }
      for (int e0 = lo; e0 < hi; e0 += 32) {
          const int myk = nxk;
          const float myc = nxc;
          nxk = 0;
          nxc = 0.f;
          if (e0 + 32 + lane < hi) {   // the next chunk's indices and coefficients: in flight behind this chunk's gathers
            nxk = sorted_k[e0 + 32 + lane];
            nxc = coef_row[nxk];
          }
          const int n = min(32, hi - e0);
          int j = 0;
          if constexpr (sizeof(T) == 2)  // bf16 rows: keep NB of them in flight as raw 16-byte vectors
          for (; j + NB <= n; j += NB) {
            uint4 raw[NB][NV];
            float cc[NB];
#pragma unroll
            for (int u = 0; u < NB; ++u) {
              const int kq = __shfl_sync(0xffffffffu, myk, j + u);
              cc[u] = __shfl_sync(0xffffffffu, myc, j + u);
              const T* row = sd + (int64_t)(k0 + kq) * D;
#pragma unroll
              for (int i = 0; i < NV; ++i) raw[u][i] = ldg_hint(row + (i * 32 + lane) * 8, row_pol);
            }
#pragma unroll
            for (int u = 0; u < NB; ++u)
#pragma unroll
              for (int i = 0; i < NV; ++i) {
                const uint32_t ww[4] = {raw[u][i].x, raw[u][i].y, raw[u][i].z, raw[u][i].w};
#pragma unroll
                for (int q2 = 0; q2 < 4; ++q2)   // packed fp32x2 FMA: same bits, half the issue slots
                  ffma2(acc[i][2 * q2], acc[i][2 * q2 + 1], cc[u], __uint_as_float(ww[q2] << 16),
                        __uint_as_float(ww[q2] & 0xffff0000u));
              }
          }
          if constexpr (sizeof(T) == 2) {
            // the last 1..NB-1 rows of the chunk as ONE partial batch: one at a time they cost a full L2 latency each --
            // with ~39 entries per token (text, 77 tokens) that was 3 of the 12 round trips of a token
            const int rem = n - j;   // warp-uniform, < NB
            if (rem > 0) {
              uint4 raw[NB - 1][NV];
              float cc[NB - 1];
#pragma unroll
              for (int u = 0; u < NB - 1; ++u) {
                const int kq = __shfl_sync(0xffffffffu, myk, (j + u) & 31);
                cc[u] = __shfl_sync(0xffffffffu, myc, (j + u) & 31);
                if (u < rem) {
                  const T* row = sd + (int64_t)(k0 + kq) * D;
#pragma unroll
                  for (int i = 0; i < NV; ++i) raw[u][i] = ldg_hint(row + (i * 32 + lane) * 8, row_pol);
                }
              }
#pragma unroll
              for (int u = 0; u < NB - 1; ++u)
                if (u < rem) {
#pragma unroll
                  for (int i = 0; i < NV; ++i) {
                    const uint32_t ww[4] = {raw[u][i].x, raw[u][i].y, raw[u][i].z, raw[u][i].w};
#pragma unroll
                    for (int q2 = 0; q2 < 4; ++q2)
                      ffma2(acc[i][2 * q2], acc[i][2 * q2 + 1], cc[u], __uint_as_float(ww[q2] << 16),
                            __uint_as_float(ww[q2] & 0xffff0000u));
                  }
                }
              j = n;
            }
          }
          for (; j < n; ++j) {
            const int kq = __shfl_sync(0xffffffffu, myk, j);
            const float c0 = __shfl_sync(0xffffffffu, myc, j);
            gb_axpy<T, NV>(sd + (int64_t)(k0 + kq) * D, c0, lane, acc);
          }
        }
      }
      TO* out = dq + (b * n_tok + t) * (int64_t)D;
#pragma unroll
      for (int i = 0; i < NV; ++i) StoreRow8<TO>::st(out + (i * 32 + lane) * 8, acc[i], k0 > 0);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 2, v3: the tiled kernel above with its shape opened up for tuning: WARPS x CPW codes per CTA and U
// q rows in flight per warp (ncu r01: the 4x4/U=2 shape ran 3.5 warps per scheduler, stalled on the row
// gathers with neither L2 nor the issue slots saturated -- the kernel is bound by loads in flight).
// ---------------------------------------------------------------------------------------------
template <typename T, int NV, int WARPS, int CPW, int U, int TS = 32>
__global__ void __launch_bounds__(WARPS * 32)
ground_bwd_dsd_v3_kernel(const float* __restrict__ coef, const int16_t* __restrict__ arg, const T* __restrict__ q,
                         const int* __restrict__ row_base, int64_t b_total, int n_tok, int K, int D,
                         int64_t b_per_split, int use_atomic, float* __restrict__ dsd,
                         const int* __restrict__ sparse_flag) {
  if (sparse_flag != nullptr && *sparse_flag != 0) return;   // the sparse-regime kernel handles this launch
  // TS samples per cooperative tile (one CTA barrier per tile). A/B on one box: TS = 32 is ~2 ms per step faster at the
  // dense random-init statistics (the headline), TS = 128 is 3 ms faster when the support is a few tens of codes and
  // the kernel is a pure scan of coef / arg (ILFDT_DSD_CFG=128)
  constexpr int SUB = TS / 32;
  const uint64_t row_pol = l2_policy((use_atomic & 4) != 0);
  constexpr int CODES = WARPS * CPW, THREADS = WARPS * 32, QUADS = CODES / 4, SLOTS = TS * QUADS;
  constexpr int LPT = (SLOTS + THREADS - 1) / THREADS;   // (sample, 4-code quad) slots per thread per tile
  __shared__ float cf_s[2][CODES][TS + 1];
  __shared__ int16_t tk_s[2][CODES][TS + 2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k0 = blockIdx.x * CODES;
  const int64_t b_begin = (int64_t)blockIdx.y * b_per_split;
  const int64_t b_end = min(b_total, b_begin + b_per_split);
  const int nblk = (int)((b_end - b_begin + TS - 1) / TS);

  float acc[CPW][NV][8];
#pragma unroll
  for (int c = 0; c < CPW; ++c)
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[c][i][j] = 0.f;

  float4 rc[LPT];
  uint2 rt[LPT];
  auto load_tile = [&](int blk) {
#pragma unroll
    for (int l = 0; l < LPT; ++l) {
      const int slot = threadIdx.x + l * THREADS;
      const int ld_s = slot / QUADS, ld_q = (slot % QUADS) * 4;
      const int64_t bb = b_begin + (int64_t)blk * TS + ld_s;
      rc[l] = make_float4(0.f, 0.f, 0.f, 0.f);
      rt[l] = make_uint2(0u, 0u);
      if (slot < SLOTS && bb < b_end) {
        if (use_atomic & 2) {   // streaming (evict-first) loads: coef / arg are read once, the q rows are the reused data
          rc[l] = __ldcs(reinterpret_cast<const float4*>(coef + bb * K + k0 + ld_q));
          rt[l] = __ldcs(reinterpret_cast<const uint2*>(arg + bb * K + k0 + ld_q));
        } else {
          rc[l] = __ldg(reinterpret_cast<const float4*>(coef + bb * K + k0 + ld_q));
          rt[l] = __ldg(reinterpret_cast<const uint2*>(arg + bb * K + k0 + ld_q));
        }
      }
    }
  };
  auto store_tile = [&](int buf) {
#pragma unroll
    for (int l = 0; l < LPT; ++l) {
      const int slot = threadIdx.x + l * THREADS;
      if (slot < SLOTS) {
        const int ld_s = slot / QUADS, ld_q = (slot % QUADS) * 4;
        cf_s[buf][ld_q + 0][ld_s] = rc[l].x; cf_s[buf][ld_q + 1][ld_s] = rc[l].y;
        cf_s[buf][ld_q + 2][ld_s] = rc[l].z; cf_s[buf][ld_q + 3][ld_s] = rc[l].w;
        tk_s[buf][ld_q + 0][ld_s] = (int16_t)(rt[l].x & 0xffff); tk_s[buf][ld_q + 1][ld_s] = (int16_t)(rt[l].x >> 16);
        tk_s[buf][ld_q + 2][ld_s] = (int16_t)(rt[l].y & 0xffff); tk_s[buf][ld_q + 3][ld_s] = (int16_t)(rt[l].y >> 16);
      }
    }
  };

  if (nblk > 0) {
    load_tile(0);
    store_tile(0);
  }
  __syncthreads();
  for (int blk = 0; blk < nblk; ++blk) {
    const int buf = blk & 1;
    if (blk + 1 < nblk) load_tile(blk + 1);
    for (int sub = 0; sub < SUB; ++sub) {
    const int64_t b0 = b_begin + (int64_t)blk * TS + sub * 32;
    if (b0 >= b_end) break;
    // first q row of this lane's sample: sample * n_tok, or its slab in the compacted text operand (text_buckets.cu)
    const int64_t mybase = row_base ? (b0 + lane < b_end ? (int64_t)__ldg(row_base + b0 + lane) : 0) : (b0 + lane) * n_tok;
#pragma unroll
    for (int c = 0; c < CPW; ++c) {
      const float cf = cf_s[buf][warp * CPW + c][sub * 32 + lane];
      const int t = tk_s[buf][warp * CPW + c][sub * 32 + lane];
      const int64_t myrow = mybase + t;
      unsigned m = __ballot_sync(0xffffffffu, cf != 0.f);
      while (m) {
        const int n = min(__popc(m), U);
        float cc[U];
        uint4 raw[U][NV];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (u < n) {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            cc[u] = __shfl_sync(0xffffffffu, cf, j);
            const int64_t rr = __shfl_sync(0xffffffffu, myrow, j);
            const T* row = q + rr * D;
            if constexpr (sizeof(T) == 2) {
#pragma unroll
              for (int i = 0; i < NV; ++i) raw[u][i] = ldg_hint(row + (i * 32 + lane) * 8, row_pol);
            } else {
              gb_axpy<T, NV>(row, cc[u], lane, acc[c]);
            }
          }
        }
        if constexpr (sizeof(T) == 2) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (u < n) {
#pragma unroll
              for (int i = 0; i < NV; ++i) {
                const uint32_t ww[4] = {raw[u][i].x, raw[u][i].y, raw[u][i].z, raw[u][i].w};
#pragma unroll
                for (int q2 = 0; q2 < 4; ++q2) {
                  acc[c][i][2 * q2] = fmaf(cc[u], __uint_as_float(ww[q2] << 16), acc[c][i][2 * q2]);
                  acc[c][i][2 * q2 + 1] = fmaf(cc[u], __uint_as_float(ww[q2] & 0xffff0000u), acc[c][i][2 * q2 + 1]);
                }
              }
            }
          }
        }
      }
    }
    }
    if (blk + 1 < nblk) store_tile(buf ^ 1);
    __syncthreads();
  }
#pragma unroll
  for (int c = 0; c < CPW; ++c) {
    float* out = dsd + (int64_t)(k0 + warp * CPW + c) * D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float* o = out + (i * 32 + lane) * 8;
      if (use_atomic & 1) {
        // a lane owns 8 consecutive floats = one 32-byte sector: two 16-byte vector reductions (REDG.E.ADD.F32x4)
        // instead of eight scalar atomics
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o), "f"(acc[c][i][0]), "f"(acc[c][i][1]),
                     "f"(acc[c][i][2]), "f"(acc[c][i][3]) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o + 4), "f"(acc[c][i][4]), "f"(acc[c][i][5]),
                     "f"(acc[c][i][6]), "f"(acc[c][i][7]) : "memory");
      } else {
        StoreRow8<float>::st(o, acc[c][i], true);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Sparse regime (trained, peaked attention: a support of tens of codes per row instead of ~0.2 K). The two kernels above
// are dense SCANS of three [b,K] arrays (coef written and re-read, w, arg) that find ~0.1 % useful entries: 13 ms per
// step at the headline shape for ~23 entries per row. Here one CTA per sample scans only the bf16 / fp32 weights of its
// row (the support), and for each of its few entries does both updates itself:
//     dq[b, arg, :] += coef * sd[k, :]   a warp per token walks the sample's short entry list (fixed order: deterministic)
//     dsd[k, :]     += coef * q[b, arg, :]   with 16-byte vector reductions to global memory (fp32 sum order free)
// No coef array, no second sweep. Which regime a launch is in is decided ON THE DEVICE (gb_regime_kernel: mean support
// from the forward's nnz[b]) -- no host synchronisation, CUDA-graph safe: both code paths are enqueued and the one that
// does not apply returns at once. fp32 sums are taken in arrival order (shared / global reductions).
// ---------------------------------------------------------------------------------------------
constexpr int GSP_THREADS = 256, GSP_WARPS = GSP_THREADS / 32;

__global__ void __launch_bounds__(256) gb_regime_kernel(const int* __restrict__ nnz, int64_t b, int64_t max_total,
                                                        int* __restrict__ flag) {
  __shared__ long long sm[8];
  long long s = 0;
  for (int64_t i = threadIdx.x; i < b; i += 256) s += nnz[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t = 0;
    for (int w = 0; w < 8; ++w) t += sm[w];
    flag[0] = (t <= max_total) ? 1 : 0;
  }
}

constexpr int GSP_CAP = 2048;   // support entries of one sample staged per round

template <typename T, int NV, typename TO, typename TW>
__global__ void __launch_bounds__(GSP_THREADS)
ground_bwd_sparse_kernel(const float* __restrict__ g, const TW* __restrict__ w, const int16_t* __restrict__ arg,
                         const T* __restrict__ q, const T* __restrict__ sd, const float* __restrict__ mask,
                         const int* __restrict__ row_base, const int* __restrict__ sparse_flag, int n_tok, int K, int D,
                         float inv_scale, TO* __restrict__ dq, float* __restrict__ dsd) {
  if (*sparse_flag == 0) return;
  __shared__ int list_k[GSP_CAP];
  __shared__ int16_t list_t[GSP_CAP];
  __shared__ float list_c[GSP_CAP];
  __shared__ double red_s[GSP_WARPS];
  __shared__ int red_c[GSP_WARPS];
  __shared__ int warp_off[GSP_WARPS + 1];
  const int64_t b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* gr = g + b * K;
  const TW* wr = w + b * K;
  const int16_t* ar = arg + b * K;
  const int64_t qbase = row_base ? (int64_t)row_base[b] : b * n_tok;
  // The row is walked in blocks of 256 threads x 4 codes x 16 = 16384 codes: per block, every thread issues its 16
  // independent support loads at once and keeps the 64 membership bits in registers.
  constexpr int PER = 16;
  int done_rounds = 0;
  // Almost every row of this sample's [n_tok, D] gradient block is zero. Every row is written exactly ONCE, by the warp
  // that owns its token in the first round below -- its accumulated entries, or the zeros the accumulators start from
  // (r2 first zero-filled the block up front and wrote the rows with entries a second time: ncu showed 7.1 GB of DRAM
  // writes per text launch against 5.2 GB of dq). A sample without any entry is zero-filled at the end.
  for (int k0 = 0; k0 < K; k0 += GSP_THREADS * 4 * PER) {
    uint64_t bits = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int k = k0 + (j * GSP_THREADS + tid) * 4;
      if (k < K) bits |= (uint64_t)Support4<TW>::ld(wr + k) << (4 * j);
    }
    // (s, c) over the WHOLE row are needed before any coefficient: first block pass only counts when K <= 16384;
    // longer rows take a separate counting pass below
    int cnt = __popcll(bits);
    // deterministic placement: exclusive scan of the per-thread counts (thread order, then bit order)
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += n;
    }
    if (lane == 31) warp_off[warp + 1] = incl;
    __syncthreads();
    if (tid == 0) {
      warp_off[0] = 0;
      for (int i = 0; i < GSP_WARPS; ++i) warp_off[i + 1] += warp_off[i];
    }
    __syncthreads();
    const int total = warp_off[GSP_WARPS];
    int pos = warp_off[warp] + incl - cnt;
    // mean of g over the support of the whole row (computed once, on the first block; rows longer than one block
    // count the remaining blocks here as well)
    if (k0 == 0) {
      double s = 0.0;
      int c = 0;
      uint64_t bb = bits;
      while (bb) {
        const int bit = __ffsll((long long)bb) - 1;
        bb &= bb - 1;
        s += (double)gr[((bit >> 2) * GSP_THREADS + tid) * 4 + (bit & 3)];
        ++c;
      }
      for (int k1 = GSP_THREADS * 4 * PER; k1 < K; k1 += GSP_THREADS * 4 * PER)
#pragma unroll 4
        for (int j = 0; j < PER; ++j) {
          const int k = k1 + (j * GSP_THREADS + tid) * 4;
          if (k < K) {
            uint32_t m = Support4<TW>::ld(wr + k);
            while (m) {
              const int e = __ffs(m) - 1;
              m &= m - 1;
              s += (double)gr[k + e];
              ++c;
            }
          }
        }
      s = warp_sum(s);
      c = warp_sum(c);
      if (lane == 0) { red_s[warp] = s; red_c[warp] = c; }
      __syncthreads();
      if (tid == 0) {
        double ts = 0.0;
        int tc = 0;
        for (int i = 0; i < GSP_WARPS; ++i) { ts += red_s[i]; tc += red_c[i]; }
        red_s[0] = tc > 0 ? ts / (double)tc : 0.0;
      }
      __syncthreads();
    }
    const float mean = (float)red_s[0];
    // rounds of at most GSP_CAP entries (one round unless a single sample of a sparse launch has > 2048 support codes)
    for (int base = 0; base < total; base += GSP_CAP, ++done_rounds) {
      uint64_t bb = bits;
      int p = pos;
      while (bb) {
        const int bit = __ffsll((long long)bb) - 1;
        bb &= bb - 1;
        if (p >= base && p < base + GSP_CAP) {
          const int k = k0 + ((bit >> 2) * GSP_THREADS + tid) * 4 + (bit & 3);
          const int t = ar[k];
          const float keep = (mask == nullptr || mask[b * n_tok + t] == 0.f) ? 1.f : 0.f;
          list_k[p - base] = k;
          list_t[p - base] = (int16_t)t;
          list_c[p - base] = (gr[k] - mean) * keep * inv_scale;
        }
        ++p;
      }
      __syncthreads();
      const int n_ent = min(GSP_CAP, total - base);
      for (int t = warp; t < n_tok; t += GSP_WARPS) {
        float acc[NV][8];
#pragma unroll
        for (int u = 0; u < NV; ++u)
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[u][e] = 0.f;
        float vq[NV][8];
        bool have_q = false;   // also: this token owns at least one entry
        for (int e0 = 0; e0 < n_ent; e0 += 32) {
          const int e = e0 + lane;
          const bool mine = e < n_ent && list_t[e] == t && list_c[e] != 0.f;
          unsigned hits = __ballot_sync(0xffffffffu, mine);
          while (hits) {
            const int src = __ffs(hits) - 1;
            hits &= hits - 1;
            const int k = list_k[e0 + src];
            const float cf = list_c[e0 + src];
            if (!have_q) {
#pragma unroll
              for (int u = 0; u < NV; ++u) Load8<T>::ld(q + (qbase + t) * (int64_t)D + (u * 32 + lane) * 8, vq[u]);
              have_q = true;
            }
            const T* srow = sd + (int64_t)k * D;
            float* drow = dsd + (int64_t)k * D;
#pragma unroll
            for (int u = 0; u < NV; ++u) {
              float vs[8];
              Load8<T>::ld(srow + (u * 32 + lane) * 8, vs);
#pragma unroll
              for (int x = 0; x < 8; ++x) acc[u][x] = fmaf(cf, vs[x], acc[u][x]);
              float* o = drow + (u * 32 + lane) * 8;
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o), "f"(cf * vq[u][0]), "f"(cf * vq[u][1]),
                           "f"(cf * vq[u][2]), "f"(cf * vq[u][3]) : "memory");
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o + 4), "f"(cf * vq[u][4]), "f"(cf * vq[u][5]),
                           "f"(cf * vq[u][6]), "f"(cf * vq[u][7]) : "memory");
            }
          }
        }
        if (have_q || done_rounds == 0) {
          TO* out = dq + (b * n_tok + t) * (int64_t)D;
#pragma unroll
          for (int u = 0; u < NV; ++u) StoreRow8<TO>::st(out + (u * 32 + lane) * 8, acc[u], done_rounds > 0);
        }
      }
      __syncthreads();
    }
  }
  if (done_rounds == 0) {   // no entry at all (cannot happen behind a sparsemax, whose support is never empty)
    uint4* z16 = reinterpret_cast<uint4*>(dq + b * n_tok * (int64_t)D);
    const int n16 = (int)((size_t)n_tok * D * sizeof(TO) / 16);
    for (int i = tid; i < n16; i += GSP_THREADS) z16[i] = make_uint4(0u, 0u, 0u, 0u);
  }
}

template <typename T, int NV, typename TO, typename TW>
static int32_t launch_sparse(const float* g, const TW* w, const int16_t* arg, const T* q, const T* sd, const float* mask,
                             const int* row_base, const int* flag, int64_t b, int64_t n_tok, int64_t K, int64_t D,
                             float inv_scale, TO* dq, float* dsd, cudaStream_t st) {
  ground_bwd_sparse_kernel<T, NV, TO, TW><<<(unsigned)b, GSP_THREADS, 0, st>>>(g, w, arg, q, sd, mask, row_base, flag,
                                                                             (int)n_tok, (int)K, (int)D, inv_scale, dq, dsd);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

// ---------------------------------------------------------------------------------------------
// kernel 2, v4 (bf16 rows): the gather stream of v3 without its bubbles. A measured L2 -> SM gather of 1 KB rows runs
// at 18.2 TB/s on this part (tools/l2_probe.cu: 64 B/clk/SM, the SM's L2 ingest port; profiles/r02_l2_probe.jsonl),
// v3 reached 12.3: a warp found ~6 entries per 32-sample tile, gathered them as one batch of 4 and one of 2 and waited
// out a full L2 latency for each, with a CTA barrier per tile. Here entry DISCOVERY and row GATHERING are decoupled:
// a warp (one code) pushes the (row, coef) pairs it finds in a tile into a private ring in shared memory and the
// gather loop pops groups of G rows through a two-deep register pipeline -- the loads of group i+1 are issued before
// group i is consumed, across tile boundaries and barriers -- so G..2G rows stay in flight per warp at all times.
// Summation order per code is still sample order: results equal v3's bit for bit.
// ---------------------------------------------------------------------------------------------
template <typename T, int NV, int WARPS, int TS, int G, bool F2 = false>
__global__ void __launch_bounds__(WARPS * 32)
ground_bwd_dsd_v4_kernel(const float* __restrict__ coef, const int16_t* __restrict__ arg, const T* __restrict__ q,
                         const int* __restrict__ row_base, int64_t b_total, int n_tok, int K, int D,
                         int64_t b_per_split, int use_atomic, float* __restrict__ dsd,
                         const int* __restrict__ sparse_flag) {
  static_assert(sizeof(T) == 2, "v4 gathers raw bf16 rows");
  if (sparse_flag != nullptr && *sparse_flag != 0) return;
  constexpr int CODES = WARPS, THREADS = WARPS * 32, QUADS = CODES / 4, SLOTS = TS * QUADS;
  constexpr int LPT = (SLOTS + THREADS - 1) / THREADS, RING = 2 * TS, SUB = TS / 32;
  static_assert(CODES % 4 == 0 && TS % 32 == 0 && (RING & (RING - 1)) == 0 && RING >= TS + G, "shape");
  const uint64_t row_pol = l2_policy(false);
  __shared__ float cf_s[2][CODES][TS + 1];
  __shared__ int16_t tk_s[2][CODES][TS + 2];
  __shared__ int rb_s[2][TS];
  __shared__ int ring_row[WARPS][RING];
  __shared__ float ring_cf[WARPS][RING];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k0 = blockIdx.x * CODES;
  const int64_t b_begin = (int64_t)blockIdx.y * b_per_split;
  const int64_t b_end = min(b_total, b_begin + b_per_split);
  const int nblk = (int)((b_end - b_begin + TS - 1) / TS);

  float acc[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float4 rc[LPT];
  uint2 rt[LPT];
  int rb[LPT];
  auto load_tile = [&](int blk) {
#pragma unroll
    for (int l = 0; l < LPT; ++l) {
      const int slot = threadIdx.x + l * THREADS;
      const int ld_s = slot / QUADS, ld_q = (slot % QUADS) * 4;
      const int64_t bb = b_begin + (int64_t)blk * TS + ld_s;
      rc[l] = make_float4(0.f, 0.f, 0.f, 0.f);
      rt[l] = make_uint2(0u, 0u);
      rb[l] = 0;
      if (slot < SLOTS && bb < b_end) {
        if (use_atomic & 2) {   // streaming (evict-first) loads: coef / arg are read once, the q rows are the reused data
          rc[l] = __ldcs(reinterpret_cast<const float4*>(coef + bb * K + k0 + ld_q));
          rt[l] = __ldcs(reinterpret_cast<const uint2*>(arg + bb * K + k0 + ld_q));
        } else {
          rc[l] = __ldg(reinterpret_cast<const float4*>(coef + bb * K + k0 + ld_q));
          rt[l] = __ldg(reinterpret_cast<const uint2*>(arg + bb * K + k0 + ld_q));
        }
        // first q row of the sample: sample * n_tok, or its slab in the compacted text operand (text_buckets.cu)
        if (ld_q == 0) rb[l] = row_base ? __ldg(row_base + bb) : (int)(bb * n_tok);
      }
    }
  };
  auto store_tile = [&](int buf) {
#pragma unroll
    for (int l = 0; l < LPT; ++l) {
      const int slot = threadIdx.x + l * THREADS;
      if (slot < SLOTS) {
        const int ld_s = slot / QUADS, ld_q = (slot % QUADS) * 4;
        cf_s[buf][ld_q + 0][ld_s] = rc[l].x; cf_s[buf][ld_q + 1][ld_s] = rc[l].y;
        cf_s[buf][ld_q + 2][ld_s] = rc[l].z; cf_s[buf][ld_q + 3][ld_s] = rc[l].w;
        tk_s[buf][ld_q + 0][ld_s] = (int16_t)(rt[l].x & 0xffff); tk_s[buf][ld_q + 1][ld_s] = (int16_t)(rt[l].x >> 16);
        tk_s[buf][ld_q + 2][ld_s] = (int16_t)(rt[l].y & 0xffff); tk_s[buf][ld_q + 3][ld_s] = (int16_t)(rt[l].y >> 16);
        if (ld_q == 0) rb_s[buf][ld_s] = rb[l];
      }
    }
  };

  // two-deep register pipeline over the ring
  uint4 bufA[G][NV], bufB[G][NV];
  float cfA[G], cfB[G];
  int nA = 0, nB = 0, phase = 0, head = 0, tail = 0;
  bool pend = false;
  int* my_row = ring_row[warp];
  float* my_cf = ring_cf[warp];
  auto issue = [&](uint4 (&buf)[G][NV], float (&cf)[G], int& n) {
    n = min(G, tail - head);
#pragma unroll
    for (int u = 0; u < G; ++u) {
      if (u < n) {
        const int idx = (head + u) & (RING - 1);
        cf[u] = my_cf[idx];
        const T* row = q + (int64_t)my_row[idx] * D;
#pragma unroll
        for (int i = 0; i < NV; ++i) buf[u][i] = ldg_hint(row + (i * 32 + lane) * 8, row_pol);
      }
    }
    head += n;
  };
  auto consume = [&](const uint4 (&buf)[G][NV], const float (&cf)[G], int n) {
#pragma unroll
    for (int u = 0; u < G; ++u) {
      if (u < n) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const uint32_t ww[4] = {buf[u][i].x, buf[u][i].y, buf[u][i].z, buf[u][i].w};
#pragma unroll
          for (int q2 = 0; q2 < 4; ++q2) {
            if constexpr (F2) {
              ffma2(acc[i][2 * q2], acc[i][2 * q2 + 1], cf[u], __uint_as_float(ww[q2] << 16),
                    __uint_as_float(ww[q2] & 0xffff0000u));
            } else {
              acc[i][2 * q2] = fmaf(cf[u], __uint_as_float(ww[q2] << 16), acc[i][2 * q2]);
              acc[i][2 * q2 + 1] = fmaf(cf[u], __uint_as_float(ww[q2] & 0xffff0000u), acc[i][2 * q2 + 1]);
            }
          }
        }
      }
    }
  };
  auto step = [&]() {
    if (phase == 0) {
      issue(bufA, cfA, nA);
      if (pend) consume(bufB, cfB, nB);
      phase = 1;
    } else {
      issue(bufB, cfB, nB);
      if (pend) consume(bufA, cfA, nA);
      phase = 0;
    }
    pend = true;
  };

  if (nblk > 0) {
    load_tile(0);
    store_tile(0);
  }
  __syncthreads();
  for (int blk = 0; blk < nblk; ++blk) {
    const int buf = blk & 1;
    if (blk + 1 < nblk) load_tile(blk + 1);
    // discovery: this warp's code over the TS samples of the tile -> ring (sample order)
#pragma unroll
    for (int sub = 0; sub < SUB; ++sub) {
      const float cf = cf_s[buf][warp][sub * 32 + lane];
      const int row = rb_s[buf][sub * 32 + lane] + tk_s[buf][warp][sub * 32 + lane];
      const unsigned m = __ballot_sync(0xffffffffu, cf != 0.f);
      if (cf != 0.f) {
        const int pos = (tail + __popc(m & ((1u << lane) - 1u))) & (RING - 1);
        my_row[pos] = row;
        my_cf[pos] = cf;
      }
      tail += __popc(m);
    }
    __syncwarp();
    while (tail - head >= G) step();
    if (blk + 1 < nblk) store_tile(buf ^ 1);
    __syncthreads();
  }
  while (tail - head > 0) step();
  if (pend) {
    if (phase == 1) consume(bufA, cfA, nA);
    else consume(bufB, cfB, nB);
  }
  float* out = dsd + (int64_t)(k0 + warp) * D;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    float* o = out + (i * 32 + lane) * 8;
    if (use_atomic & 1) {
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o), "f"(acc[i][0]), "f"(acc[i][1]), "f"(acc[i][2]),
                   "f"(acc[i][3]) : "memory");
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o + 4), "f"(acc[i][4]), "f"(acc[i][5]), "f"(acc[i][6]),
                   "f"(acc[i][7]) : "memory");
    } else {
      StoreRow8<float>::st(o, acc[i], true);
    }
  }
}

template <typename T, int NV, int WARPS, int TS, int G, bool F2 = false>
static int32_t launch_dsd_v4(const float* g, const int16_t* arg, const T* q, const int* row_base, int64_t b, int64_t n_tok,
                             int64_t K, int64_t D, float* dsd, const int* sparse_flag, cudaStream_t st) {
  constexpr int CODES = WARPS;
  const int64_t groups = K / CODES;
  int64_t nsplit = 1;
  const int64_t target = 4ll * sm_count();
  if (groups < target) nsplit = std::min<int64_t>(cdiv(target, groups), std::max<int64_t>(1, b / 256));
  const int64_t split_rows = tune_get(ILFDT_TUNE_DSD_SPLIT_ROWS);   // see launch_dsd_v3
  if (split_rows > 0 && b >= 2 * split_rows) nsplit = std::max<int64_t>(nsplit, b / split_rows);
  const int stream_flag = tune_get(ILFDT_TUNE_DSD_STREAM) ? 2 : 0;
  nsplit = std::min<int64_t>(std::max<int64_t>(nsplit, 1), 65535);
  int64_t per = round_up(cdiv(b, nsplit), TS);
  nsplit = cdiv(b, per);
  dim3 grid((unsigned)groups, (unsigned)nsplit);
  ground_bwd_dsd_v4_kernel<T, NV, WARPS, TS, G, F2><<<grid, WARPS * 32, 0, st>>>(g, arg, q, row_base, b, (int)n_tok, (int)K, (int)D,
                                                                            per, (nsplit > 1 ? 1 : 0) | stream_flag, dsd,
                                                                            sparse_flag);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

template <typename T, int NV, int WARPS, int CPW, int U, int TS = 32>
static int32_t launch_dsd_v3(const float* g, const int16_t* arg, const T* q, const int* row_base, int64_t b, int64_t n_tok,
                             int64_t K, int64_t D, float* dsd, const int* sparse_flag, cudaStream_t st) {
  constexpr int CODES = WARPS * CPW;
  const int64_t groups = K / CODES;
  int64_t nsplit = 1;
  const int64_t target = 4ll * sm_count();
  if (groups < target) nsplit = std::min<int64_t>(cdiv(target, groups), std::max<int64_t>(1, b / 256));
  // Large batches: the sample sweep is cut into ranges (grid.y; partial rows combined with vector reductions). One
  // CTA sweeping all of q drifts away from its neighbours, and the q rows the ~740 resident CTAs gather stop sharing
  // L2 lines (ncu r01: 70 % hit rate); short sweeps restart them together: 48.3 -> 43.8 ms per step at b = 32768
  // (A/B on one box, 8 ranges). ilfdt_tune_set(ILFDT_TUNE_DSD_SPLIT_ROWS) overrides the rows per range (0 = never split).
  const int64_t split_rows = tune_get(ILFDT_TUNE_DSD_SPLIT_ROWS);
  if (split_rows > 0 && b >= 2 * split_rows) nsplit = std::max<int64_t>(nsplit, b / split_rows);
  // coef / arg are read once per launch: streaming loads (41.2 vs 42.3 ms per step for the scatter pair, ABAB on one box)
  const int stream_flag = tune_get(ILFDT_TUNE_DSD_STREAM) ? 2 : 0;
  nsplit = std::min<int64_t>(std::max<int64_t>(nsplit, 1), 65535);
  int64_t per = round_up(cdiv(b, nsplit), TS);
  nsplit = cdiv(b, per);
  dim3 grid((unsigned)groups, (unsigned)nsplit);
  ground_bwd_dsd_v3_kernel<T, NV, WARPS, CPW, U, TS><<<grid, WARPS * 32, 0, st>>>(g, arg, q, row_base, b, (int)n_tok, (int)K,
                                                                             (int)D, per, (nsplit > 1 ? 1 : 0) | stream_flag, dsd,
                                                                             sparse_flag);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

template <typename T, int NV, typename TO, typename TW, int PG, int MINB, int NB = 4>
static int32_t launch_dq_v3_t(float* g, const TW* w, const int16_t* arg, const T* sd, const float* mask, int64_t b,
                              int64_t n_tok, int64_t K, int64_t D, float inv_scale, TO* dq, const int* sparse_flag,
                              cudaStream_t st) {
  const size_t smem = (size_t)n_tok * G3_THREADS * 2 + (size_t)G3_CHUNK * 2 + (G3_MAX_TOK + 1) * sizeof(int) +
                      G3_MAX_TOK * sizeof(float);
  ILFDT_CUDA_OK(cudaFuncSetAttribute(ground_bwd_dq_v3_kernel<T, NV, TO, TW, PG, MINB, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)((size_t)G3_MAX_TOK * G3_THREADS * 2 + (size_t)G3_CHUNK * 2 +
                                           (G3_MAX_TOK + 1) * sizeof(int) + G3_MAX_TOK * sizeof(float))));
  ground_bwd_dq_v3_kernel<T, NV, TO, TW, PG, MINB, NB><<<(unsigned)b, G3_THREADS, smem, st>>>(g, w, arg, sd, mask, (int)n_tok,
                                                                                     (int)K, (int)D, inv_scale, dq, sparse_flag);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

template <typename T, int NV, typename TO, typename TW>
static int32_t launch_dq_v3(float* g, const TW* w, const int16_t* arg, const T* sd, const float* mask, int64_t b,
                            int64_t n_tok, int64_t K, int64_t D, float inv_scale, TO* dq, const int* sparse_flag,
                            cudaStream_t st) {
  if constexpr (sizeof(T) == 2 && NV <= 2) {
    const int64_t pipe = tune_get(ILFDT_TUNE_DQ_PIPE);   // 0 = batches of 4 with exposed latency (r1), 1 / 2 = pipelined
    if (pipe == 1) return launch_dq_v3_t<T, NV, TO, TW, 2, 3>(g, w, arg, sd, mask, b, n_tok, K, D, inv_scale, dq, sparse_flag, st);
    if (pipe == 2) return launch_dq_v3_t<T, NV, TO, TW, 4, 2>(g, w, arg, sd, mask, b, n_tok, K, D, inv_scale, dq, sparse_flag, st);
    // fewer samples than two thirds of the resident CTA slots (3 per SM): 8 rows in flight per warp
    if (pipe == 0 && tune_get(ILFDT_TUNE_DQ_DEEP) != 0 && b <= 2ll * sm_count())
      return launch_dq_v3_t<T, NV, TO, TW, 0, 2, 8>(g, w, arg, sd, mask, b, n_tok, K, D, inv_scale, dq, sparse_flag, st);
  }
  return launch_dq_v3_t<T, NV, TO, TW, 0, 3>(g, w, arg, sd, mask, b, n_tok, K, D, inv_scale, dq, sparse_flag, st);
}

template <typename T, int NV>
static int32_t launch_ground_bwd_t(float* g, const void* w_v, int w_dtype, const int16_t* arg, const T* q, const T* sd,
                                   const float* mask, const int* row_base, const int* nnz, int* regime_ws, int64_t b,
                                   int64_t n_tok, int64_t K, int64_t D, float inv_scale, void* dq_v, int dq_dtype, float* dsd,
                                   cudaStream_t st) {
  bool done = false;
  const int* flag = nullptr;   // device flag: 1 = this launch is handled by the sparse-regime kernel
  float* dq = reinterpret_cast<float*>(dq_v);
  // the support comes from the fp32 weights or (scatter kernel v3 only) from their bf16 copy
  const float* w = w_dtype == ILFDT_F32 ? reinterpret_cast<const float*>(w_v) : nullptr;
  const __nv_bfloat16* w16 = w_dtype == ILFDT_BF16 ? reinterpret_cast<const __nv_bfloat16*>(w_v) : nullptr;
  if constexpr (NV > 0) {
    const bool v3_ok = n_tok <= G3_MAX_TOK && K % 4 == 0 &&
                       ((reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(w)) & 15) == 0 &&
                       ((reinterpret_cast<uintptr_t>(arg) | reinterpret_cast<uintptr_t>(w16)) & 7) == 0;
    // sparse regime (device-side decision, see ground_bwd_sparse_kernel): needs the forward's nnz[b], one int of scratch,
    // and the shapes both v3 kernels take
    const int64_t sparse_nnz = tune_get(ILFDT_TUNE_SPARSE_BWD_NNZ);
    if constexpr (NV <= 2) {
      const bool dsd_v3_ok = K % 32 == 0 && ((reinterpret_cast<uintptr_t>(g) & 15) | (reinterpret_cast<uintptr_t>(arg) & 7)) == 0;
      if (v3_ok && dsd_v3_ok && !done && nnz != nullptr && regime_ws != nullptr && sparse_nnz > 0 &&
          (w != nullptr || w16 != nullptr)) {
        gb_regime_kernel<<<1, 256, 0, st>>>(nnz, b, sparse_nnz * b, regime_ws);
        ILFDT_LAUNCH_OK();
        flag = regime_ws;
        int32_t rc;
        __nv_bfloat16* dq16 = reinterpret_cast<__nv_bfloat16*>(dq_v);
        if (w16 != nullptr)
          rc = dq_dtype == ILFDT_BF16
                   ? launch_sparse<T, NV, __nv_bfloat16, __nv_bfloat16>(g, w16, arg, q, sd, mask, row_base, flag, b, n_tok, K, D, inv_scale, dq16, dsd, st)
                   : launch_sparse<T, NV, float, __nv_bfloat16>(g, w16, arg, q, sd, mask, row_base, flag, b, n_tok, K, D, inv_scale, dq, dsd, st);
        else
          rc = dq_dtype == ILFDT_BF16
                   ? launch_sparse<T, NV, __nv_bfloat16, float>(g, w, arg, q, sd, mask, row_base, flag, b, n_tok, K, D, inv_scale, dq16, dsd, st)
                   : launch_sparse<T, NV, float, float>(g, w, arg, q, sd, mask, row_base, flag, b, n_tok, K, D, inv_scale, dq, dsd, st);
        if (rc != ILFDT_OK) return rc;
      }
    }
    if (v3_ok && !done) {
      int32_t rc;
      if (w16 != nullptr)
        rc = dq_dtype == ILFDT_BF16
                 ? launch_dq_v3<T, NV, __nv_bfloat16, __nv_bfloat16>(g, w16, arg, sd, mask, b, n_tok, K, D, inv_scale,
                                                                     reinterpret_cast<__nv_bfloat16*>(dq_v), flag, st)
                 : launch_dq_v3<T, NV, float, __nv_bfloat16>(g, w16, arg, sd, mask, b, n_tok, K, D, inv_scale, dq, flag, st);
      else
        rc = dq_dtype == ILFDT_BF16
                 ? launch_dq_v3<T, NV, __nv_bfloat16, float>(g, w, arg, sd, mask, b, n_tok, K, D, inv_scale,
                                                             reinterpret_cast<__nv_bfloat16*>(dq_v), flag, st)
                 : launch_dq_v3<T, NV, float, float>(g, w, arg, sd, mask, b, n_tok, K, D, inv_scale, dq, flag, st);
      if (rc != ILFDT_OK) return rc;
      done = true;
    }
  }
  if constexpr (NV > 0) {
    if (!done && n_tok <= 256) {
      const size_t smem2 = GB_CHUNK * sizeof(GbEnt2) + (size_t)(GB_WARPS * n_tok + n_tok + 1) * sizeof(int);
      const int smem_max = (int)(GB_CHUNK * sizeof(GbEnt2) + (GB_WARPS * 256 + 257) * sizeof(int));
#define ILFDT_DQ_SORTED(TO_, TW_, WPTR, DQPTR)                                                                         \
  do {                                                                                                                 \
    ILFDT_CUDA_OK(cudaFuncSetAttribute(ground_bwd_dq_sorted_kernel<T, NV, TO_, TW_>,                                   \
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));                        \
    ground_bwd_dq_sorted_kernel<T, NV, TO_, TW_><<<(unsigned)b, GB_THREADS, smem2, st>>>(                             \
        g, WPTR, arg, sd, mask, (int)n_tok, (int)K, (int)D, inv_scale, DQPTR);                                         \
  } while (0)
      __nv_bfloat16* dq16 = reinterpret_cast<__nv_bfloat16*>(dq_v);
      if (w16 != nullptr) {
        if (dq_dtype == ILFDT_BF16) ILFDT_DQ_SORTED(__nv_bfloat16, __nv_bfloat16, w16, dq16);
        else ILFDT_DQ_SORTED(float, __nv_bfloat16, w16, dq);
      } else {
        if (dq_dtype == ILFDT_BF16) ILFDT_DQ_SORTED(__nv_bfloat16, float, w, dq16);
        else ILFDT_DQ_SORTED(float, float, w, dq);
      }
#undef ILFDT_DQ_SORTED
      ILFDT_LAUNCH_OK();
      done = true;
    }
  }
  ILFDT_REQUIRE(done || w16 == nullptr, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_bwd: bf16 weights need D in {256,512,768,1024}, n_tok <= 256, K %% 4 == 0");
  ILFDT_REQUIRE(done || dq_dtype == ILFDT_F32, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_bwd: bf16 dq needs D in {256,512,768,1024}, n_tok <= 256, K %% 4 == 0");
  if (!done) {
    const size_t smem = (size_t)std::min<int64_t>(GB_CHUNK, K) * sizeof(GbEntry);
    ILFDT_CUDA_OK(cudaFuncSetAttribute(ground_bwd_dq_kernel<T, NV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(GB_CHUNK * sizeof(GbEntry))));
    ground_bwd_dq_kernel<T, NV><<<(unsigned)b, GB_THREADS, smem, st>>>(g, w, arg, sd, mask, (int)n_tok, (int)K,
                                                                      (int)D, inv_scale, dq);
    ILFDT_LAUNCH_OK();
  }
  if constexpr (NV > 0) {
    // shape of kernel 2 (warps x codes per warp, rows in flight), tuned on B200 at b = 8192, K = 16384, D = 512
    // (profiles/r01_tune_dsd_shapes.txt): 4x1 codes with 4 rows in flight beat the original 4x4/2 by 1.4x.
    // ILFDT_TUNE_DSD_TILE = 128 selects the 128-sample cooperative tile (better when the support is tens of codes).
    if (K % 32 == 0 && ((reinterpret_cast<uintptr_t>(g) & 15) | (reinterpret_cast<uintptr_t>(arg) & 7)) == 0) {
      if constexpr (NV <= 2 && sizeof(T) == 2) {
        // bf16 rows: pipelined gather stream (v4); ILFDT_TUNE_DSD_PIPE = 0 selects the v3 kernel, 2 = groups of 4 rows
        const int64_t pipe = tune_get(ILFDT_TUNE_DSD_PIPE);
        if (b * n_tok < (1ll << 31)) {
          // shapes measured at b = 32768 (tools/time_gbwd.py, profiles/r02_time_gbwd.txt; both kernels, image + text):
          // v3 33.6 ms, 4 codes x 64 samples with groups of 2 rows 31.5 ms (default: 80 registers, 6 CTAs / SM),
          // groups of 4 rows 33.3 ms (118 registers), 128-sample tiles 31.5 ms, 8 codes per CTA 32.8 ms
          // packed fp32x2 FMAs (FFMA2) in the accumulation: 30.73 against 31.26 ms (same bits)
          if (pipe == 1) return launch_dsd_v4<T, NV, 4, 64, 2, true>(g, arg, q, row_base, b, n_tok, K, D, dsd, flag, st);
          if (pipe == 2) return launch_dsd_v4<T, NV, 4, 64, 4, true>(g, arg, q, row_base, b, n_tok, K, D, dsd, flag, st);
          if (pipe == 3) return launch_dsd_v4<T, NV, 4, 64, 2, false>(g, arg, q, row_base, b, n_tok, K, D, dsd, flag, st);
        }
      }
      if constexpr (NV <= 2) {
        if (tune_get(ILFDT_TUNE_DSD_TILE) == 128)
          return launch_dsd_v3<T, NV, 4, 1, 4, 128>(g, arg, q, row_base, b, n_tok, K, D, dsd, flag, st);
        return launch_dsd_v3<T, NV, 4, 1, 4, 32>(g, arg, q, row_base, b, n_tok, K, D, dsd, flag, st);
      } else {
        return launch_dsd_v3<T, NV, 4, 1, 2>(g, arg, q, row_base, b, n_tok, K, D, dsd, flag, st);
      }
    }
  }
  ILFDT_REQUIRE(row_base == nullptr, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_bwd: a compacted q needs D in {256,...,1024}, K %% 32 == 0 and aligned pointers");
  // split the sample range when there are too few code groups to fill the machine
  const int64_t groups = cdiv(K, GS_WARPS);
  int64_t nsplit = 1;
  const int64_t target = 4ll * sm_count();
  if (groups < target) nsplit = std::min<int64_t>(cdiv(target, groups), std::max<int64_t>(1, b / 256));
  nsplit = std::min<int64_t>(std::max<int64_t>(nsplit, 1), 65535);
  int64_t per = round_up(cdiv(b, nsplit), 32);
  nsplit = cdiv(b, per);
  dim3 grid((unsigned)groups, (unsigned)nsplit);
  ground_bwd_dsd_kernel<T, NV><<<grid, GS_WARPS * 32, 0, st>>>(g, arg, q, b, (int)n_tok, (int)K, (int)D, per,
                                                              nsplit > 1 ? 1 : 0, dsd);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

template <typename T>
static int32_t dispatch_nv(float* g, const void* w, int w_dtype, const int16_t* arg, const T* q, const T* sd, const float* mask,
                           const int* row_base, const int* nnz, int* regime_ws, int64_t b, int64_t n_tok, int64_t K, int64_t D,
                           float inv_scale, void* dq, int dq_dtype, float* dsd, cudaStream_t st) {
  const bool aligned = ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(sd) |
                         reinterpret_cast<uintptr_t>(dq) | reinterpret_cast<uintptr_t>(dsd)) & 31) == 0;
  if (aligned && D == 512) return launch_ground_bwd_t<T, 2>(g, w, w_dtype, arg, q, sd, mask, row_base, nnz, regime_ws, b, n_tok, K, D, inv_scale, dq, dq_dtype, dsd, st);
  if (aligned && D == 256) return launch_ground_bwd_t<T, 1>(g, w, w_dtype, arg, q, sd, mask, row_base, nnz, regime_ws, b, n_tok, K, D, inv_scale, dq, dq_dtype, dsd, st);
  if (aligned && D == 768) return launch_ground_bwd_t<T, 3>(g, w, w_dtype, arg, q, sd, mask, row_base, nnz, regime_ws, b, n_tok, K, D, inv_scale, dq, dq_dtype, dsd, st);
  if (aligned && D == 1024) return launch_ground_bwd_t<T, 4>(g, w, w_dtype, arg, q, sd, mask, row_base, nnz, regime_ws, b, n_tok, K, D, inv_scale, dq, dq_dtype, dsd, st);
  ILFDT_REQUIRE(D <= 1024, ILFDT_E_UNSUPPORTED_SHAPE, "ground_bwd: D=%lld > 1024", (long long)D);
  return launch_ground_bwd_t<T, 0>(g, w, w_dtype, arg, q, sd, mask, row_base, nnz, regime_ws, b, n_tok, K, D, inv_scale, dq, dq_dtype, dsd, st);
}

int32_t launch_ground_bwd(float* g, const void* w, int w_dtype, const int16_t* arg, const void* q, const void* sd, int dtype,
                          const float* mask, const int* row_base, const int* nnz, int* regime_ws, int64_t b, int64_t n_tok,
                          int64_t K, int64_t D, float sqrt_d, float temperature, void* dq, int dq_dtype, float* dsd,
                          cudaStream_t st) {
  if (b == 0) return ILFDT_OK;
  ILFDT_REQUIRE(b < (1ll << 31) && K < (1ll << 31) && n_tok <= 32767, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_bwd: shape out of range");
  const float inv_scale = 1.f / (sqrt_d * temperature);
  if (dtype == ILFDT_F32)
    return dispatch_nv<float>(g, w, w_dtype, arg, reinterpret_cast<const float*>(q), reinterpret_cast<const float*>(sd), mask,
                              row_base, nnz, regime_ws, b, n_tok, K, D, inv_scale, dq, dq_dtype, dsd, st);
  return dispatch_nv<__nv_bfloat16>(g, w, w_dtype, arg, reinterpret_cast<const __nv_bfloat16*>(q),
                                    reinterpret_cast<const __nv_bfloat16*>(sd), mask, row_base, nnz, regime_ws, b, n_tok, K, D,
                                    inv_scale, dq, dq_dtype, dsd, st);
}

}  // namespace ilfdt
