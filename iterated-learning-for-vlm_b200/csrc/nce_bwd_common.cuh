// Shared by the single-CTA and the CTA-pair InfoNCE backward kernels (nce_bwd_umma.cu, nce_bwd_umma2.cu):
// launch parameters and the softmax warps' per-tile work (stream-column constants, S -> P conversion).
#pragma once
#include "umma.cuh"
#include "launch.h"

namespace ilfdt {

struct NceBwdParams {
  int RL, RS, D;
  int nkb;              // ceil(D / 64)
  int num_lane_tiles, num_halves, nsplit, tiles_per_split, num_stream_tiles;
  int label_by_lane;    // 1: label(lane i) = stream label_offset + i;  0: label(stream j) = lane label_offset + j
  long long label_offset;
  const float* lse_lane;    // [RL]  (MODE 0, 2)
  const float* lse_stream;  // [RS]  (MODE 1, 2)
  const float* go_lane;     // 1 float: upstream dLoss of the rank owning the lane rows
  const float* go_stream;   // [ceil(RS / rows_per_go)]: upstream dLoss of the rank owning each stream row
  long long rows_per_go;
  float inv_2b;
  const float* scale_dev;
  float* out;           // [RL, D]
  int use_atomic;
  float* dscale;        // accumulated only by lse_by_lane launches, half 0
  uint32_t idesc1, idesc2;
};

struct NceBwdLaunch {
  const void* XL; const void* Y;      // lane matrix [RL, D], stream matrix [RS, D] (bf16)
  int64_t RL, RS, D;
  int mode;                            // 0 lane softmax, 1 stream softmax, 2 both
  int label_by_lane;
  int64_t label_offset;
  const float* lse_lane; const float* lse_stream;
  const float* go_lane; const float* go_stream;
  int64_t rows_per_go;
  float inv_2b;
  const float* scale;
  float* out; float* dscale;
};

// per-stream-column constants of the tile a softmax group is converting, one buffer per (group [, tile parity]).
// Columns beyond RS need no masking: TMA zero-fills their Y rows, so S = 0 there and whatever finite P they get
// multiplies zeros in MMA2.
struct NceColBuf {
  alignas(16) float lf[64];   // log2(go_j / 2b) - lse_j * log2(e): the stream weight folded into the exponent (lean path)
  uint32_t neg[4];            // [loader warp]: some weight of the tile is negative (general path needed)
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// dOut epilogue of a split item: 32 consecutive columns of one output row added into global memory. A thread owns a row,
// so a warp's scalar atomics are 32 uncoalesced requests per instruction and 128 instructions per thread: ~17 us of
// LSU time per item, more than the whole tile loop at b <= 256 (ncu, 256-row step: 39 us per launch). 16-byte vector
// reductions (REDG.E.ADD.F32x4) need a quarter of the requests.
__device__ __forceinline__ void nce_red_row32(float* o, const uint32_t (&v)[32], float scale, int c0, int dcols, bool vec_ok) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    if (vec_ok && c0 + j + 3 < dcols) {
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o + j), "f"(__uint_as_float(v[j]) * scale),
                   "f"(__uint_as_float(v[j + 1]) * scale), "f"(__uint_as_float(v[j + 2]) * scale),
                   "f"(__uint_as_float(v[j + 3]) * scale) : "memory");
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (c0 + j + e < dcols) atomicAdd(o + j + e, __uint_as_float(v[j + e]) * scale);
    }
  }
}

// Executed by the first two warps (tid128 < 64) of a softmax group: fetch (global -> registers; may be issued a
// tile ahead) and publish (registers -> shared, before the group's named barrier).
template <int MODE>
__device__ __forceinline__ void nce_fetch_cols(const NceBwdParams& p, int j0, int tid128, float& ws, float& l2) {
  ws = 0.f; l2 = 0.f;
  if (MODE != 0) {
    const uint32_t n = (uint32_t)(j0 + tid128);
    if (n < (uint32_t)p.RS) {
      ws = __ldg(p.go_stream + n / (uint32_t)p.rows_per_go) * p.inv_2b;
      l2 = __ldg(p.lse_stream + n) * 1.4426950408889634f;
    }
  }
}
template <int MODE>
__device__ __forceinline__ void nce_publish_cols(NceColBuf* c, int tid128, int lane, float wl, float ws, float l2) {
  bool neg = wl < 0.f;
  if (MODE != 0) {
    c->lf[tid128] = __log2f(ws) - l2;     // ws == 0 -> -inf -> ex2 gives 0; ws < 0 -> general path (flag below)
    neg = neg || ws < 0.f;
  }
  const bool any_neg = __any_sync(0xffffffffu, neg);
  if (lane == 0) c->neg[tid128 >> 5] = any_neg ? 1u : 0u;
}

// S tile row (64 fp32 logits of one lane row, v0 = columns 0..31, v1 = 32..63) -> 64 bf16 of P packed in pairs.
//   general == false: weights folded into the exponents, P = 2^(a + lane_off) + 2^(a + lf_j), no label term
//   general == true : P = wl (2^(a - lse_lane) - [j == jd]) + ws_j (2^(a - l2_j) - [j == jd])
// ds_acc accumulates sum_j P_lane[j] * S[j] (the lane rows' own logits: dLoss / dscale).
template <int MODE>
__device__ __forceinline__ void nce_convert(const uint32_t (&v0)[32], const uint32_t (&v1)[32], uint32_t (&packed)[32],
                                            bool general, int jd, float scale_log2, float lane_off, float lse_lane,
                                            float wl, const NceColBuf* cols, const NceBwdParams& p, int j0, float& ds_acc) {
  const float* c_lf = cols->lf;
  if (!general) {
    float ds_e = 0.f;
#pragma unroll
    for (int h2 = 0; h2 < 2; ++h2) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        float4 lf = make_float4(0.f, 0.f, 0.f, 0.f);
        if (MODE != 0) lf = *reinterpret_cast<const float4*>(c_lf + h2 * 32 + j);
        const float lfa[4] = {lf.x, lf.y, lf.z, lf.w};
        float pv[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float s_raw = __uint_as_float(h2 == 0 ? v0[j + e] : v1[j + e]);
          float pr = 0.f;
          if (MODE != 1) {
            pr = ex2_approx(fmaf(s_raw, scale_log2, lane_off));
            ds_e = fmaf(pr, s_raw, ds_e);
          }
          if (MODE != 0) pr += ex2_approx(fmaf(s_raw, scale_log2, lfa[e]));
          pv[e] = pr;
        }
        __nv_bfloat162 b01 = __floats2bfloat162_rn(pv[0], pv[1]);
        __nv_bfloat162 b23 = __floats2bfloat162_rn(pv[2], pv[3]);
        packed[h2 * 16 + (j >> 1)] = *reinterpret_cast<uint32_t*>(&b01);
        packed[h2 * 16 + (j >> 1) + 1] = *reinterpret_cast<uint32_t*>(&b23);
      }
    }
    ds_acc += ds_e;
  } else {
#pragma unroll
    for (int h2 = 0; h2 < 2; ++h2) {
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        float pv[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int jj = h2 * 32 + j + e;
          const float s_raw = __uint_as_float(h2 == 0 ? v0[j + e] : v1[j + e]);
          const float dlt = (jj == jd) ? 1.f : 0.f;
          float pr = 0.f;
          if (MODE != 1) {
            pr = (exp2f(fmaf(s_raw, scale_log2, -lse_lane)) - dlt) * wl;
            ds_acc = fmaf(pr, s_raw, ds_acc);
          }
          if (MODE != 0) {
            // the few general tiles (label columns, negative weights) read the stream constants straight from global
            float ws, l2;
            nce_fetch_cols<MODE>(p, j0, jj, ws, l2);
            pr = fmaf(exp2f(fmaf(s_raw, scale_log2, -l2)) - dlt, ws, pr);
          }
          pv[e] = pr;
        }
        __nv_bfloat162 b2 = __floats2bfloat162_rn(pv[0], pv[1]);
        packed[h2 * 16 + (j >> 1)] = *reinterpret_cast<uint32_t*>(&b2);
      }
    }
  }
}

// host: the persistent grid's work split (nsplit slices of the stream range per (lane tile, D half))
inline void nce_bwd_plan(NceBwdParams& p, int base_items, int workers) {
  int nsplit = 1;
  if (base_items % workers != 0) {
    double best = -1.0;
    const int max_split =
        (int)std::min<int64_t>(p.num_stream_tiles, std::max<int64_t>(1, cdiv((int64_t)4 * workers, base_items)));
    for (int c = 1; c <= max_split; ++c) {
      const int tps = (int)cdiv(p.num_stream_tiles, c);
      const int rc = (int)cdiv(p.num_stream_tiles, tps);
      const int64_t items = (int64_t)base_items * rc;
      double eff = (double)items / (double)(cdiv(items, (int64_t)workers) * workers);
      eff *= (double)tps / (double)(tps + 6.0);   // every item pays a lane-tile load, an epilogue and (split) atomics
      if (eff > best + 1e-9) { best = eff; nsplit = rc; }
    }
  }
  p.tiles_per_split = (int)cdiv(p.num_stream_tiles, nsplit);
  p.nsplit = (int)cdiv(p.num_stream_tiles, p.tiles_per_split);
  p.use_atomic = p.nsplit > 1 ? 1 : 0;
}

// nce_bwd_umma2.cu
bool nce_bwd_pair_enabled();
int32_t nce_bwd_pair_launch(const NceBwdLaunch& a, cudaStream_t st);

}  // namespace ilfdt
