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

template <typename T, int NV>
__global__ void __launch_bounds__(GB_THREADS)
ground_bwd_dq_sorted_kernel(float* __restrict__ g, const float* __restrict__ w, const int16_t* __restrict__ arg,
                            const T* __restrict__ sd, const float* __restrict__ mask, int n_tok, int K, int D,
                            float inv_scale, float* __restrict__ dq) {
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
  const float* wr = w ? w + b * K : nullptr;
  const int16_t* ar = arg + b * K;
  const float* mr = mask ? mask + b * n_tok : nullptr;
  int* myhist = hist + warp * n_tok;

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
    const int wend = ((max(we - wb, 0) + 31) / 32) * 32 + wb;
    for (int t = lane; t < n_tok; t += 32) myhist[t] = 0;
    __syncwarp();
    // pass 1: coef (left in g for kernel 2) + per-warp token histogram
    for (int i = wb + lane; i < wend; i += 32) {
      float cf = 0.f;
      int t = 0;
      if (i < we) {
        if (wr == nullptr || wr[k0 + i] > 0.f) {
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
      float* out = dq + (b * n_tok + t) * (int64_t)D;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        float* o = out + (i * 32 + lane) * 8;
        float4 a = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        float4 bb = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
        if (k0 > 0) {
          const float4 pa = *reinterpret_cast<float4*>(o), pb = *reinterpret_cast<float4*>(o + 4);
          a.x += pa.x; a.y += pa.y; a.z += pa.z; a.w += pa.w;
          bb.x += pb.x; bb.y += pb.y; bb.z += pb.z; bb.w += pb.w;
        }
        *reinterpret_cast<float4*>(o) = a;
        *reinterpret_cast<float4*>(o + 4) = bb;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 2, tiled: a CTA owns GT_CODES = 16 consecutive codes (4 warps x 4 codes, 64 fp32 accumulators per
// lane) and sweeps the samples 32 at a time. The [32 samples x 16 codes] tiles of coef / arg are read
// with fully coalesced 16-byte loads (the strided per-warp reads of the kernel above cost 18x their
// bytes in DRAM traffic, ncu r01), transposed through shared memory (double buffered, one barrier per
// tile, next tile's global loads in flight during the gathers), then lane = sample as before.
// ---------------------------------------------------------------------------------------------
constexpr int GT_WARPS = 4, GT_CPW = 4, GT_CODES = GT_WARPS * GT_CPW, GT_THREADS = GT_WARPS * 32;

template <typename T, int NV>
__global__ void __launch_bounds__(GT_THREADS)
ground_bwd_dsd_tiled_kernel(const float* __restrict__ coef, const int16_t* __restrict__ arg, const T* __restrict__ q,
                            int64_t b_total, int n_tok, int K, int D, int64_t b_per_split, int use_atomic,
                            float* __restrict__ dsd) {
  __shared__ float cf_s[2][GT_CODES][33];
  __shared__ int16_t tk_s[2][GT_CODES][34];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k0 = blockIdx.x * GT_CODES;
  const int64_t b_begin = (int64_t)blockIdx.y * b_per_split;
  const int64_t b_end = min(b_total, b_begin + b_per_split);
  const int nblk = (int)((b_end - b_begin + 31) / 32);
  // loader role: thread -> (sample row, 4-code quad)
  const int ld_s = threadIdx.x >> 2, ld_q = (threadIdx.x & 3) * 4;

  float acc[GT_CPW][NV][8];
#pragma unroll
  for (int c = 0; c < GT_CPW; ++c)
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[c][i][j] = 0.f;

  float4 rc = make_float4(0.f, 0.f, 0.f, 0.f);
  uint2 rt = make_uint2(0u, 0u);
  auto load_tile = [&](int blk) {
    const int64_t bb = b_begin + (int64_t)blk * 32 + ld_s;
    if (bb < b_end) {
      rc = __ldg(reinterpret_cast<const float4*>(coef + bb * K + k0 + ld_q));
      rt = __ldg(reinterpret_cast<const uint2*>(arg + bb * K + k0 + ld_q));
    } else {
      rc = make_float4(0.f, 0.f, 0.f, 0.f);
      rt = make_uint2(0u, 0u);
    }
  };
  auto store_tile = [&](int buf) {
    cf_s[buf][ld_q + 0][ld_s] = rc.x; cf_s[buf][ld_q + 1][ld_s] = rc.y;
    cf_s[buf][ld_q + 2][ld_s] = rc.z; cf_s[buf][ld_q + 3][ld_s] = rc.w;
    tk_s[buf][ld_q + 0][ld_s] = (int16_t)(rt.x & 0xffff); tk_s[buf][ld_q + 1][ld_s] = (int16_t)(rt.x >> 16);
    tk_s[buf][ld_q + 2][ld_s] = (int16_t)(rt.y & 0xffff); tk_s[buf][ld_q + 3][ld_s] = (int16_t)(rt.y >> 16);
  };

  if (nblk > 0) {
    load_tile(0);
    store_tile(0);
  }
  __syncthreads();
  for (int blk = 0; blk < nblk; ++blk) {
    const int buf = blk & 1;
    if (blk + 1 < nblk) load_tile(blk + 1);
    const int64_t b0 = b_begin + (int64_t)blk * 32;
#pragma unroll
    for (int c = 0; c < GT_CPW; ++c) {
      const float cf = cf_s[buf][warp * GT_CPW + c][lane];
      const int t = tk_s[buf][warp * GT_CPW + c][lane];
      unsigned m = __ballot_sync(0xffffffffu, cf != 0.f);
      while (m) {
        const int j0 = __ffs(m) - 1;
        m &= m - 1;
        const int t0 = __shfl_sync(0xffffffffu, t, j0);
        const float c0 = __shfl_sync(0xffffffffu, cf, j0);
        const T* r0 = q + ((b0 + j0) * n_tok + t0) * (int64_t)D;
        if (m) {
          const int j1 = __ffs(m) - 1;
          m &= m - 1;
          const int t1 = __shfl_sync(0xffffffffu, t, j1);
          const float c1 = __shfl_sync(0xffffffffu, cf, j1);
          const T* r1 = q + ((b0 + j1) * n_tok + t1) * (int64_t)D;
          float v0[NV][8], v1[NV][8];
#pragma unroll
          for (int i = 0; i < NV; ++i) {
            Load8<T>::ld(r0 + (i * 32 + lane) * 8, v0[i]);
            Load8<T>::ld(r1 + (i * 32 + lane) * 8, v1[i]);
          }
#pragma unroll
          for (int i = 0; i < NV; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[c][i][j] = fmaf(c1, v1[i][j], fmaf(c0, v0[i][j], acc[c][i][j]));
        } else {
          gb_axpy<T, NV>(r0, c0, lane, acc[c]);
        }
      }
    }
    if (blk + 1 < nblk) store_tile(buf ^ 1);
    __syncthreads();
  }
#pragma unroll
  for (int c = 0; c < GT_CPW; ++c) {
    float* out = dsd + (int64_t)(k0 + warp * GT_CPW + c) * D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float* o = out + (i * 32 + lane) * 8;
      if (use_atomic) {
#pragma unroll
        for (int j = 0; j < 8; ++j) atomicAdd(o + j, acc[c][i][j]);
      } else {
        float4 pa = *reinterpret_cast<float4*>(o), pb = *reinterpret_cast<float4*>(o + 4);
        pa.x += acc[c][i][0]; pa.y += acc[c][i][1]; pa.z += acc[c][i][2]; pa.w += acc[c][i][3];
        pb.x += acc[c][i][4]; pb.y += acc[c][i][5]; pb.z += acc[c][i][6]; pb.w += acc[c][i][7];
        *reinterpret_cast<float4*>(o) = pa;
        *reinterpret_cast<float4*>(o + 4) = pb;
      }
    }
  }
}

template <typename T, int NV>
static int32_t launch_ground_bwd_t(float* g, const float* w, const int16_t* arg, const T* q, const T* sd,
                                   const float* mask, int64_t b, int64_t n_tok, int64_t K, int64_t D,
                                   float inv_scale, float* dq, float* dsd, cudaStream_t st) {
  bool done = false;
  if constexpr (NV > 0) {
    if (n_tok <= 256) {
      const size_t smem2 = GB_CHUNK * sizeof(GbEnt2) + (size_t)(GB_WARPS * n_tok + n_tok + 1) * sizeof(int);
      ILFDT_CUDA_OK(cudaFuncSetAttribute(ground_bwd_dq_sorted_kernel<T, NV>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(GB_CHUNK * sizeof(GbEnt2) + (GB_WARPS * 256 + 257) * sizeof(int))));
      ground_bwd_dq_sorted_kernel<T, NV><<<(unsigned)b, GB_THREADS, smem2, st>>>(g, w, arg, sd, mask, (int)n_tok,
                                                                                (int)K, (int)D, inv_scale, dq);
      ILFDT_LAUNCH_OK();
      done = true;
    }
  }
  if (!done) {
    const size_t smem = (size_t)std::min<int64_t>(GB_CHUNK, K) * sizeof(GbEntry);
    ILFDT_CUDA_OK(cudaFuncSetAttribute(ground_bwd_dq_kernel<T, NV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(GB_CHUNK * sizeof(GbEntry))));
    ground_bwd_dq_kernel<T, NV><<<(unsigned)b, GB_THREADS, smem, st>>>(g, w, arg, sd, mask, (int)n_tok, (int)K,
                                                                      (int)D, inv_scale, dq);
    ILFDT_LAUNCH_OK();
  }
  if constexpr (NV > 0) {
    if (K % GT_CODES == 0 && ((reinterpret_cast<uintptr_t>(g) & 15) | (reinterpret_cast<uintptr_t>(arg) & 7)) == 0) {
      const int64_t groups = K / GT_CODES;
      int64_t nsplit = 1;
      const int64_t target = 4ll * sm_count();
      if (groups < target) nsplit = std::min<int64_t>(cdiv(target, groups), std::max<int64_t>(1, b / 256));
      nsplit = std::min<int64_t>(std::max<int64_t>(nsplit, 1), 65535);
      int64_t per = round_up(cdiv(b, nsplit), 32);
      nsplit = cdiv(b, per);
      dim3 grid((unsigned)groups, (unsigned)nsplit);
      ground_bwd_dsd_tiled_kernel<T, NV><<<grid, GT_THREADS, 0, st>>>(g, arg, q, b, (int)n_tok, (int)K, (int)D, per,
                                                                     nsplit > 1 ? 1 : 0, dsd);
      ILFDT_LAUNCH_OK();
      return ILFDT_OK;
    }
  }
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
static int32_t dispatch_nv(float* g, const float* w, const int16_t* arg, const T* q, const T* sd, const float* mask,
                           int64_t b, int64_t n_tok, int64_t K, int64_t D, float inv_scale, float* dq, float* dsd,
                           cudaStream_t st) {
  const bool aligned = ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(sd) |
                         reinterpret_cast<uintptr_t>(dq) | reinterpret_cast<uintptr_t>(dsd)) & 31) == 0;
  if (aligned && D == 512) return launch_ground_bwd_t<T, 2>(g, w, arg, q, sd, mask, b, n_tok, K, D, inv_scale, dq, dsd, st);
  if (aligned && D == 256) return launch_ground_bwd_t<T, 1>(g, w, arg, q, sd, mask, b, n_tok, K, D, inv_scale, dq, dsd, st);
  if (aligned && D == 768) return launch_ground_bwd_t<T, 3>(g, w, arg, q, sd, mask, b, n_tok, K, D, inv_scale, dq, dsd, st);
  if (aligned && D == 1024) return launch_ground_bwd_t<T, 4>(g, w, arg, q, sd, mask, b, n_tok, K, D, inv_scale, dq, dsd, st);
  ILFDT_REQUIRE(D <= 1024, ILFDT_E_UNSUPPORTED_SHAPE, "ground_bwd: D=%lld > 1024", (long long)D);
  return launch_ground_bwd_t<T, 0>(g, w, arg, q, sd, mask, b, n_tok, K, D, inv_scale, dq, dsd, st);
}

int32_t launch_ground_bwd(float* g, const float* w, const int16_t* arg, const void* q, const void* sd, int dtype,
                          const float* mask, int64_t b, int64_t n_tok, int64_t K, int64_t D, float sqrt_d,
                          float temperature, float* dq, float* dsd, cudaStream_t st) {
  if (b == 0) return ILFDT_OK;
  ILFDT_REQUIRE(b < (1ll << 31) && K < (1ll << 31) && n_tok <= 32767, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_bwd: shape out of range");
  const float inv_scale = 1.f / (sqrt_d * temperature);
  if (dtype == ILFDT_F32)
    return dispatch_nv<float>(g, w, arg, reinterpret_cast<const float*>(q), reinterpret_cast<const float*>(sd), mask,
                              b, n_tok, K, D, inv_scale, dq, dsd, st);
  return dispatch_nv<__nv_bfloat16>(g, w, arg, reinterpret_cast<const __nv_bfloat16*>(q),
                                    reinterpret_cast<const __nv_bfloat16*>(sd), mask, b, n_tok, K, D, inv_scale, dq,
                                    dsd, st);
}

}  // namespace ilfdt
