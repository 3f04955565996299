// tcgen05 / TMEM / TMA contraction kernel with fused epilogues (bf16 inputs, fp32 accumulation).
//
//   acc[m, n] = sum_k A[m, k] * B[n, k]        A [M, Kr], B [N, Kr] bf16, reduction index contiguous
//
// Mapping: the 128 rows of an A tile are the 128 TMEM lanes, the (<= 256) rows of a B tile are TMEM
// columns, so an epilogue thread owns one A row and walks the B rows sequentially in registers:
//   * grounding (a1):   A = codebook (lane = code), B = tokens  -> per-sample max/argmax over the
//                       token columns is a private running max, no cross-lane reduction;
//   * InfoNCE fwd (a10): A = local embeddings (lane = logits row), B = gathered embeddings
//                       -> online log-sum-exp over the columns is private too;
//   * plain contraction: out[n, m] store is coalesced across lanes.
// Neither the [b, n_tok, K] scores nor the [b, B] logits are ever written to memory.
//
// Structure (one CTA per SM, persistent over work items = (A tile, range of B tiles)):
//   warp 0   TMA producer: A tile either resident for the whole item (Kr <= 512: 8 x 16 KB slots,
//            loaded once per item) or streamed with B; B tiles through a 3/4-stage mbarrier ring
//   warp 1   MMA issuer: one elected lane issues tcgen05.mma (128 x block_n x 16) into one of two
//            TMEM accumulators (2 x 256 columns = all 512), tcgen05.commit frees smem stages
//   warps 2-5 epilogue: tcgen05.ld 32 columns at a time, fused reduction, release the accumulator
// so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "umma.cuh"
#include "poolmax.cuh"
#include "launch.h"
#include <algorithm>
#include <mutex>

namespace ilfdt {

using namespace umma;

constexpr int UM_BM = 128;
constexpr int UM_BK = 64;
constexpr int UM_A_BYTES = UM_BM * UM_BK * 2;  // 16 KB
constexpr int UM_B_BYTES = 256 * UM_BK * 2;    // 32 KB (max block_n)
constexpr int UM_THREADS = 320;  // TMA warp + MMA warp + 2 groups of 4 epilogue warps
constexpr int UM_DATA_BYTES = 229376;          // 224 KB of operand staging
constexpr int UM_SMEM_BYTES = UM_DATA_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;
constexpr int UM_TMEM_COLS = 512;
constexpr int UM_ACC_COLS = 256;

enum { EPI_STORE_T = 0, EPI_POOLMAX = 1, EPI_LSE = 2, EPI_RANK = 3 };

struct UmmaParams {
  int num_m_tiles, num_n_tiles, nchunk, tiles_per_chunk;
  int nkb;      // ceil(Kr / 64)
  int block_n;  // UMMA N: rows of B per tile (multiple of 16, <= 256)
  int n_step;   // rows of B advanced per tile (<= block_n)
  int m_rows, n_rows;
  uint32_t idesc;
  // EPI_STORE_T
  int mn_major;   // 1: A is stored [Kr, M], B is stored [Kr, N] (the M / N index contiguous): both read as MN-major operands
  float* out;
  int64_t ldo;
  float alpha;
  const float* alpha_dev;
  int accumulate;
  // EPI_STORE_T, MN-major, split reduction (weight-gradient shapes: short M, N and a very long Kr): item chunk index
  // = ks * num_n_tiles + n_tile; split ks contracts k-blocks [ks*kb_per, (ks+1)*kb_per) into its own output slab
  int ksplit, kb_per;
  int64_t split_stride;   // elements between slabs
  // EPI_STORE_T, Linear-layer form (q_map, clip_fdt.py:88,91): out_bf16[n, m] = bf16(alpha * acc + bias[m]) instead of `out`
  __nv_bfloat16* out_bf16;
  const float* bias;
  // EPI_POOLMAX
  const float* mask;
  int n_tok, spt, n_samples;
  float pool_scale;
  float* pooled;
  int16_t* arg;
  // EPI_POOLMAX over one bucket of the compacted text operand (text_buckets.cu): extent read from DEVICE memory
  const int* dyn;          // {samples, first row, first slot, rows per sample} or nullptr
  const int* sample_map;   // slot -> sample (rows of pooled / arg), or nullptr
  // EPI_LSE
  const float* scale_dev;
  long long label_offset;
  float* part_m;
  float* part_s;
  float* diag;
  // EPI_RANK (N1): rank[row] += #{columns whose logit exceeds the row's label logit diag[row]}
  int* rank;
};

struct UmmaBars {
  uint64_t full[4], empty[4], tmem_full[2], tmem_empty[2], a_full, a_empty;
  uint32_t tmem_ptr;
};

template <int EPI, bool A_RES, int NTOK>
__global__ void __launch_bounds__(UM_THREADS, 1)
umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const UmmaParams p) {
  extern __shared__ uint8_t um_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(um_smem_raw) + 1023) & ~(uintptr_t)1023);
  UmmaBars* bars = reinterpret_cast<UmmaBars*>(smem + UM_DATA_BYTES);
  constexpr int STAGES = A_RES ? 3 : 4;
  constexpr int STAGE_BYTES = A_RES ? UM_B_BYTES : (UM_A_BYTES + UM_B_BYTES);
  uint8_t* ring = A_RES ? smem + 8 * UM_A_BYTES : smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int num_items = p.num_m_tiles * p.nchunk;
  // bucket launches: number of B tiles, first B row and the chunking follow the device-side plan
  int dyn_samples = p.n_samples, dyn_row0 = 0, dyn_slot0 = 0, num_n_tiles = p.num_n_tiles, tiles_per_chunk = p.tiles_per_chunk;
  if (EPI == EPI_POOLMAX && p.dyn != nullptr) {
    dyn_samples = __ldg(p.dyn);
    dyn_row0 = __ldg(p.dyn + 1);
    dyn_slot0 = __ldg(p.dyn + 2);
    num_n_tiles = (dyn_samples + p.spt - 1) / p.spt;
    tiles_per_chunk = (num_n_tiles + p.nchunk - 1) / p.nchunk;
    num_items = num_n_tiles == 0 ? 0 : p.num_m_tiles * ((num_n_tiles + tiles_per_chunk - 1) / tiles_per_chunk);
  }

  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(&bars->full[i], 1); mbar_init(&bars->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->tmem_full[i], 1); mbar_init(&bars->tmem_empty[i], 4); }
    mbar_init(&bars->a_full, 1);
    mbar_init(&bars->a_empty, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    tmem_alloc(&bars->tmem_ptr, UM_TMEM_COLS);
    tmem_relinquish();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = bars->tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      uint32_t stage = 0, phase = 0, it = 0;
      const uint32_t tx_bytes = (A_RES ? 0u : (uint32_t)UM_A_BYTES) + (uint32_t)p.block_n * 128u;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
        const int m_tile = item % p.num_m_tiles;
        int chunk = item / p.num_m_tiles, kb0 = 0, kb1 = p.nkb;
        if (EPI == EPI_STORE_T && !A_RES && p.ksplit > 1) {
          const int ks = chunk / num_n_tiles;
          chunk -= ks * num_n_tiles;
          kb0 = ks * p.kb_per;
          kb1 = min(p.nkb, kb0 + p.kb_per);
        }
        const int n_begin = chunk * tiles_per_chunk, n_end = min(num_n_tiles, n_begin + tiles_per_chunk);
        if (A_RES) {
          mbar_wait(&bars->a_empty, (it & 1) ^ 1);
          mbar_expect_tx(&bars->a_full, (uint32_t)p.nkb * UM_A_BYTES);
          for (int kb = 0; kb < p.nkb; ++kb)
            tma_load_2d(smem + kb * UM_A_BYTES, &tmA, &bars->a_full, kb * UM_BK, m_tile * UM_BM, kEvictLast);
        }
        for (int n = n_begin; n < n_end; ++n) {
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(&bars->empty[stage], phase ^ 1);
            mbar_expect_tx(&bars->full[stage], tx_bytes);
            uint8_t* st = ring + stage * STAGE_BYTES;
            if (EPI == EPI_STORE_T && !A_RES && p.mn_major) {
              // boxes of [64 reduction rows x 64 M- or N-columns]: 2 for the 128-wide A tile, block_n / 64 for B
              for (int j = 0; j < 2; ++j)
                tma_load_2d(st + j * 8192, &tmA, &bars->full[stage], m_tile * UM_BM + j * 64, kb * UM_BK, kEvictNormal);
              st += UM_A_BYTES;
              for (int j = 0; j < p.block_n / 64; ++j)
                tma_load_2d(st + j * 8192, &tmB, &bars->full[stage], n * p.n_step + j * 64, kb * UM_BK, kEvictNormal);
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
              continue;
            }
            if (!A_RES) {
              tma_load_2d(st, &tmA, &bars->full[stage], kb * UM_BK, m_tile * UM_BM, kEvictNormal);
              st += UM_A_BYTES;
            }
            tma_load_2d(st, &tmB, &bars->full[stage], kb * UM_BK, dyn_row0 + n * p.n_step, kEvictNormal);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0, it = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
        int chunk = item / p.num_m_tiles, kb0 = 0, kb1 = p.nkb;
        if (EPI == EPI_STORE_T && !A_RES && p.ksplit > 1) {
          const int ks = chunk / num_n_tiles;
          chunk -= ks * num_n_tiles;
          kb0 = ks * p.kb_per;
          kb1 = min(p.nkb, kb0 + p.kb_per);
        }
        const int n_begin = chunk * tiles_per_chunk, n_end = min(num_n_tiles, n_begin + tiles_per_chunk);
        if (A_RES) {
          mbar_wait(&bars->a_full, it & 1);
          fence_after_sync();
        }
        for (int n = n_begin; n < n_end; ++n) {
          mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
          fence_after_sync();
          const uint32_t tmem_d = tmem_base + acc * UM_ACC_COLS;
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(&bars->full[stage], phase);
            fence_after_sync();
            uint8_t* st = ring + stage * STAGE_BYTES;
            if (EPI == EPI_STORE_T && !A_RES && p.mn_major) {
              // MN-major operands: 64-column blocks 8 KB apart (leading offset), a 16-row K step = 2 KB
              const uint64_t adesc = make_smem_desc_sw128_mn(smem_u32(st), 8192);
              const uint64_t bdesc = make_smem_desc_sw128_mn(smem_u32(st + UM_A_BYTES), 8192);
#pragma unroll
              for (int k4 = 0; k4 < UM_BK / 16; ++k4)
                mma_f16(tmem_d, advance_desc_k(adesc, k4 * 2048), advance_desc_k(bdesc, k4 * 2048), p.idesc,
                        ((kb - kb0) | k4) != 0 ? 1u : 0u);
            } else {
            const uint64_t adesc = make_smem_desc_sw128(smem_u32(A_RES ? smem + kb * UM_A_BYTES : st));
            const uint64_t bdesc = make_smem_desc_sw128(smem_u32(A_RES ? st : st + UM_A_BYTES));
#pragma unroll
            for (int k4 = 0; k4 < UM_BK / 16; ++k4)
              mma_f16(tmem_d, advance_desc_k(adesc, k4 * 32), advance_desc_k(bdesc, k4 * 32), p.idesc,
                      (kb | k4) != 0 ? 1u : 0u);
            }
            mma_commit(&bars->empty[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          mma_commit(&bars->tmem_full[acc]);
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1;
        }
        if (A_RES) mma_commit(&bars->a_empty);
      }
    }
  } else {
    // ===================== epilogue warps (2..9) =====================
    // two groups of four warps; group g drains accumulator g, i.e. every other tile, so each group has
    // two MMA tile times to finish one epilogue.
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const uint32_t group = (uint32_t)(warp - 2) >> 2;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    uint32_t tile_seq = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int m_tile = item % p.num_m_tiles;
      int chunk = item / p.num_m_tiles;
      float* out_base = p.out;
      if (EPI == EPI_STORE_T && !A_RES && p.ksplit > 1) {
        const int ks = chunk / num_n_tiles;
        chunk -= ks * num_n_tiles;
        out_base += (int64_t)ks * p.split_stride;
      }
      const int n_begin = chunk * tiles_per_chunk, n_end = min(num_n_tiles, n_begin + tiles_per_chunk);
      const int row = m_tile * UM_BM + quad * 32 + lane;  // A row owned by this thread
      const bool row_ok = row < p.m_rows;
      // EPI_LSE state (base-2 domain)
      float run_m = -INFINITY, run_s = 0.f;
      const float scale = (EPI == EPI_LSE || EPI == EPI_RANK) ? __ldg(p.scale_dev) : 1.f;
      const float rank_ref = (EPI == EPI_RANK && row_ok) ? __ldg(p.diag + row) : INFINITY;
      int rank_cnt = 0;
      const float scale_log2 = scale * 1.4426950408889634f;
      const float alpha = (EPI == EPI_STORE_T) ? (p.alpha_dev ? p.alpha * __ldg(p.alpha_dev) : p.alpha) : 1.f;
      const int64_t ldo_ = p.ldo;
      for (int n = n_begin; n < n_end; ++n, ++tile_seq) {
        const uint32_t acc = tile_seq & 1, acc_phase = (tile_seq >> 1) & 1;
        if (acc != group) continue;
        mbar_wait(&bars->tmem_full[acc], acc_phase);
        fence_after_sync();
        const uint32_t taddr = tmem_base + lane_addr + acc * UM_ACC_COLS;
        const int n0 = n * p.n_step;

        if (EPI == EPI_STORE_T) {
          const int ncols = min(p.n_step, p.n_rows - n0);
          for (int c0 = 0; c0 < ncols; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_ld_wait();
            if (row_ok) {
              // one pointer stepped by ldo; the full-chunk overwrite case is branch-free (the generic form cost ~24
              // instructions per stored value and made g = df.sd^T epilogue-bound: 1.07 ms for 0.26 ms of MMA)
              float* o = out_base + (int64_t)(n0 + c0) * ldo_ + row;
              if (p.out_bf16 != nullptr) {
                // the thread owns output feature `row` of the Linear layer: its bias is one scalar
                const float bias = p.bias != nullptr ? __ldg(p.bias + row) : 0.f;
                __nv_bfloat16* ob = p.out_bf16 + (int64_t)(n0 + c0) * ldo_ + row;
                if (c0 + 32 <= ncols) {
#pragma unroll
                  for (int j = 0; j < 32; ++j, ob += ldo_) *ob = __float2bfloat16_rn(fmaf(alpha, __uint_as_float(v[j]), bias));
                } else {
                  for (int j = 0; j < ncols - c0; ++j, ob += ldo_) {
                    float r = 0.f;
#pragma unroll
                    for (int e = 0; e < 32; ++e) r = (e == j) ? __uint_as_float(v[e]) : r;
                    *ob = __float2bfloat16_rn(fmaf(alpha, r, bias));
                  }
                }
              } else if (c0 + 32 <= ncols && !p.accumulate) {
#pragma unroll
                for (int j = 0; j < 32; ++j, o += ldo_) *o = alpha * __uint_as_float(v[j]);
              } else if (c0 + 32 <= ncols) {
                // read-modify-write: all 32 loads are issued before the first store. Written as `*o += ...` per column the
                // compiler cannot prove that a store does not alias the next load (the pitch is a runtime value) and
                // serialises 256 load -> add -> store round trips to L2 / HBM per warp and tile: 124 us per launch of
                // dsd += w^T.df at ANY batch size (ncu launch list of a 256-row step), against ~10 us of traffic.
                float old[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) old[j] = __ldcg(o + j * ldo_);
#pragma unroll
                for (int j = 0; j < 32; ++j) o[j * ldo_] = old[j] + alpha * __uint_as_float(v[j]);
              } else {
                for (int j = 0; j < ncols - c0; ++j, o += ldo_) {
                  float r = 0.f;
#pragma unroll
                  for (int e = 0; e < 32; ++e) r = (e == j) ? alpha * __uint_as_float(v[e]) : r;   // v stays in registers
                  *o = p.accumulate ? (*o + r) : r;
                }
              }
            }
          }
        } else if (EPI == EPI_POOLMAX && NTOK > 0) {
          poolmax_static<(NTOK > 0 ? NTOK : 1)>(taddr, n * p.spt, dyn_samples, row, row_ok, p.m_rows, p.pool_scale,
                                                p.pooled, p.arg, p.sample_map, dyn_slot0);
        } else if (EPI == EPI_POOLMAX) {
          // generic token count / explicit mask: columns = tokens of samples [s0, s0 + nsmp), n_tok columns each
          const int s0 = n * p.spt;
          const int nsmp = min(p.spt, p.n_samples - s0);
          const int ncols = nsmp * p.n_tok;
          int tok = 0, smp = s0, best_i = 0;
          float best = 0.f;
          for (int c0 = 0; c0 < ncols; c0 += 32) {
            uint32_t keep = 0xffffffffu;
            if (p.mask != nullptr) {
              const int col = c0 + lane;
              const float mv = (col < ncols) ? __ldg(p.mask + (int64_t)n0 + col) : 0.f;
              keep = __ballot_sync(0xffffffffu, mv == 0.f);
            }
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (c0 + j < ncols) {  // warp-uniform
                // pads contribute score 0 (clip_fdt.py:125-130)
                const float x = ((keep >> j) & 1u) ? __uint_as_float(v[j]) : 0.f;
                if (tok == 0 || x > best) { best = x; best_i = tok; }
                if (++tok == p.n_tok) {
                  if (row_ok) {
                    const int64_t o = (int64_t)smp * p.m_rows + row;
                    p.pooled[o] = best * p.pool_scale;
                    p.arg[o] = (int16_t)best_i;
                  }
                  tok = 0;
                  ++smp;
                }
              }
            }
          }
        } else if (EPI == EPI_RANK) {
          // same scale multiply as the EPI_LSE label pick and the materialising logits kernel: the comparison is the
          // one accuracy() would make on stored logits (misc.py:464-477), column for column
          const int ncols = min(p.n_step, p.n_rows - n0);
          for (int c0 = 0; c0 < ncols; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j)
              rank_cnt += (c0 + j < ncols && __uint_as_float(v[j]) * scale > rank_ref) ? 1 : 0;
          }
        } else {  // EPI_LSE
          const int ncols = min(p.n_step, p.n_rows - n0);
          const long long label = p.label_offset + row;
          const long long wl0 = p.label_offset + m_tile * UM_BM + quad * 32;  // labels of this warp: [wl0, wl0+32)
          for (int c0 = 0; c0 < ncols; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_ld_wait();
            float t[32];
            float cm = -INFINITY;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              t[j] = (c0 + j < ncols) ? __uint_as_float(v[j]) * scale_log2 : -INFINITY;
              cm = fmaxf(cm, t[j]);
            }
            const float nm = fmaxf(run_m, cm);
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) s += exp2f(t[j] - nm);
            run_s = run_s * exp2f(run_m - nm) + s;
            run_m = nm;
            const long long cbeg = (long long)n0 + c0;
            if (wl0 < cbeg + 32 && wl0 + 32 > cbeg) {  // warp-uniform: some label falls in this chunk
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (row_ok && cbeg + j == label && c0 + j < ncols) p.diag[row] = __uint_as_float(v[j]) * scale;
            }
          }
        }
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
      }
      if (EPI == EPI_RANK && row_ok && rank_cnt != 0) atomicAdd(p.rank + row, rank_cnt);
      if (EPI == EPI_LSE && row_ok) {  // one partial per (chunk, epilogue group)
        p.part_m[(int64_t)(chunk * 2 + group) * p.m_rows + row] = run_m;
        p.part_s[(int64_t)(chunk * 2 + group) * p.m_rows + row] = run_s;
      }
    }
  }

  fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, UM_TMEM_COLS);
  }
}

// lse[r] = ln2 * (M + log2(sum_c s_c * 2^(m_c - M)))
__global__ void lse_combine_kernel(const float* __restrict__ part_m, const float* __restrict__ part_s, int nchunk,
                                   int64_t rows, float* __restrict__ lse) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  float M = -INFINITY;
  for (int c = 0; c < nchunk; ++c) M = fmaxf(M, part_m[(int64_t)c * rows + r]);
  float s = 0.f;
  for (int c = 0; c < nchunk; ++c) s += part_s[(int64_t)c * rows + r] * exp2f(part_m[(int64_t)c * rows + r] - M);
  lse[r] = 0.6931471805599453f * (M + log2f(s));
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(f);
  });
  return fn;
}

int32_t make_tmap_bf16(CUtensorMap* out, const void* base, int64_t rows, int64_t inner, int64_t row_stride_elems,
                       int box_rows) {
  PFN_encodeTiled enc = get_encode_fn();
  ILFDT_REQUIRE(enc != nullptr, ILFDT_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
  ILFDT_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (row_stride_elems * 2) % 16 == 0,
                ILFDT_E_INVALID_ARG, "TMA operand must be 16-byte aligned with a 16-byte multiple row pitch");
  ILFDT_REQUIRE(box_rows >= 1 && box_rows <= 256, ILFDT_E_INVALID_ARG, "TMA box rows %d out of range", box_rows);
  cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)row_stride_elems * 2};
  cuuint32_t box[2] = {(cuuint32_t)UM_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  ILFDT_REQUIRE(r == CUDA_SUCCESS, ILFDT_E_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return ILFDT_OK;
}

// Pick the number of B-tile chunks so that (A tiles x chunks) fills the persistent grid evenly while
// each item still amortises its resident A tile over several B tiles.
static void plan_items(int num_m_tiles, int num_n_tiles, int sms, int& nchunk, int& tiles_per_chunk) {
  int best_c = 1;
  double best_eff = -1.0;
  const int max_c = std::max(1, std::min(num_n_tiles, 4096));
  for (int c = 1; c <= max_c; ++c) {
    const int tpc = (int)cdiv(num_n_tiles, c);
    const int real_c = (int)cdiv(num_n_tiles, tpc);
    const int64_t items = (int64_t)num_m_tiles * real_c;
    const int64_t waves = cdiv(items, sms);
    double eff = (double)items / (double)(waves * sms);
    // every item pays one A-tile load (~8 B-tile-stages worth): prefer longer items
    eff *= (double)tpc / (double)(tpc + 0.5);
    if (eff > best_eff + 1e-9) { best_eff = eff; best_c = real_c; }
    if (tpc == 1) break;
  }
  nchunk = best_c;
  tiles_per_chunk = (int)cdiv(num_n_tiles, nchunk);
  nchunk = (int)cdiv(num_n_tiles, tiles_per_chunk);
}

template <int EPI, bool A_RES, int NTOK = 0>
static int32_t launch_umma_t(const CUtensorMap& tmA, const CUtensorMap& tmB, const UmmaParams& p, cudaStream_t st) {
  static thread_local int configured_dev = -1;
  int dev = 0;
  ILFDT_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    ILFDT_CUDA_OK(cudaFuncSetAttribute(umma_kernel<EPI, A_RES, NTOK>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       UM_SMEM_BYTES));
    configured_dev = dev;
  }
  const int items = p.num_m_tiles * p.nchunk;
  const int grid = std::min(items, sm_count());
  umma_kernel<EPI, A_RES, NTOK><<<grid, UM_THREADS, UM_SMEM_BYTES, st>>>(tmA, tmB, p);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

template <int EPI>
static int32_t launch_umma(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, UmmaParams& p,
                           cudaStream_t st) {
  ILFDT_REQUIRE(Kr % 8 == 0 && Kr >= 8, ILFDT_E_UNSUPPORTED_SHAPE, "tensor-core path needs Kr %% 8 == 0 (Kr=%lld)",
                (long long)Kr);
  ILFDT_REQUIRE(M < (1ll << 31) && N < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "dimension too large");
  ILFDT_REQUIRE(p.block_n % 16 == 0 && p.block_n >= 16 && p.block_n <= 256 && p.n_step <= p.block_n,
                ILFDT_E_INVALID_ARG, "bad block_n %d / n_step %d", p.block_n, p.n_step);
  p.m_rows = (int)M;
  p.n_rows = (int)N;
  p.nkb = (int)cdiv(Kr, UM_BK);
  p.num_m_tiles = (int)cdiv(M, UM_BM);
  p.num_n_tiles = (int)cdiv(N, p.n_step);
  p.idesc = make_idesc_bf16(UM_BM, p.block_n);
  const bool a_res = p.nkb <= 8;
  if (p.nchunk <= 0) {
    if (a_res) plan_items(p.num_m_tiles, p.num_n_tiles, sm_count(), p.nchunk, p.tiles_per_chunk);
    else { p.nchunk = p.num_n_tiles; p.tiles_per_chunk = 1; }
  }
  CUtensorMap tmA, tmB;
  int32_t rc = make_tmap_bf16(&tmA, A, M, Kr, Kr, UM_BM);
  if (rc != ILFDT_OK) return rc;
  rc = make_tmap_bf16(&tmB, B, N, Kr, Kr, p.block_n);
  if (rc != ILFDT_OK) return rc;
  if (EPI == EPI_POOLMAX && p.mask == nullptr) {
    // static token counts: unrolled max-pool epilogue. With a streamed A operand (reduction longer than 512: the 6x-wide
    // operands of the fp32-grade contraction) the same epilogues exist for the bucket launches of the compacted text
    // operand, which the generic epilogue does not serve (it has no slot map).
#define ILFDT_POOLMAX_CASE(NT)                                                        \
  case NT:                                                                            \
    return a_res ? launch_umma_t<EPI_POOLMAX, true, NT>(tmA, tmB, p, st)              \
                 : launch_umma_t<EPI_POOLMAX, false, NT>(tmA, tmB, p, st);
    switch (p.n_tok) {
      ILFDT_POOLMAX_CASE(16)
      ILFDT_POOLMAX_CASE(32)
      ILFDT_POOLMAX_CASE(48)
      ILFDT_POOLMAX_CASE(64)
      ILFDT_POOLMAX_CASE(49)
      ILFDT_POOLMAX_CASE(77)
      ILFDT_POOLMAX_CASE(196)
      ILFDT_POOLMAX_CASE(256)
      default: break;
    }
#undef ILFDT_POOLMAX_CASE
  }
  ILFDT_REQUIRE(!(EPI == EPI_POOLMAX && p.dyn != nullptr), ILFDT_E_UNSUPPORTED_SHAPE,
                "bucket launch with %d tokens per sample has no static epilogue", p.n_tok);
  return a_res ? launch_umma_t<EPI, true>(tmA, tmB, p, st) : launch_umma_t<EPI, false>(tmA, tmB, p, st);
}

int32_t launch_gemm_nt_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha,
                            const float* alpha_dev, int accumulate, float* out, int64_t ldo, cudaStream_t st) {
  if (M == 0 || N == 0) return ILFDT_OK;
  if (alpha_dev == nullptr && gemm_wide2_enabled() && gemm_wide2_supported(M, N, Kr, out, ldo))
    return launch_gemm_wide2(A, B, M, N, Kr, alpha, accumulate, out, ldo, st);   // short A, long reduction: CTA pair
  UmmaParams p{};
  p.block_n = (int)std::min<int64_t>(256, round_up(N, 16));
  p.n_step = p.block_n;
  p.out = out;
  p.ldo = ldo;
  p.alpha = alpha;
  p.alpha_dev = alpha_dev;
  p.accumulate = accumulate;
  return launch_umma<EPI_STORE_T>(A, B, M, N, Kr, p, st);
}

// Linear layer: out_bf16[n, m] = bf16(sum_k A[m,k] * B[n,k] + bias[m]), A = weight [M = out features, Kr], B = activations
// [N = rows, Kr]; the bias add and the bf16 rounding sit in the TMEM read-out, so the fp32 product never reaches HBM.
int32_t launch_linear_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, const float* bias,
                           void* out_bf16, int64_t ldo, cudaStream_t st) {
  if (M == 0 || N == 0) return ILFDT_OK;
  UmmaParams p{};
  p.block_n = (int)std::min<int64_t>(256, round_up(N, 16));
  p.n_step = p.block_n;
  p.out = nullptr;
  p.out_bf16 = (__nv_bfloat16*)out_bf16;
  p.bias = bias;
  p.ldo = ldo;
  p.alpha = 1.f;
  return launch_umma<EPI_STORE_T>(A, B, M, N, Kr, p, st);
}

// out[n, m] (+)= alpha * sum_k A[k, m] * B[k, n]: A [Kr, M], B [Kr, N] bf16 with the REDUCTION index outermost (e.g.
// dsd[K,D] += w[b,K]^T . df[b,D] straight from the row-major w and df: no transposed copies). tcgen05 reads both as
// MN-major operands.
int32_t launch_gemm_tn_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha, int accumulate,
                            float* out, int64_t ldo, cudaStream_t st) {
  if (M == 0 || N == 0) return ILFDT_OK;
  ILFDT_REQUIRE(M % 8 == 0 && N % 8 == 0 && Kr >= 1, ILFDT_E_UNSUPPORTED_SHAPE,
                "gemm_tn(bf16): M, N must be multiples of 8 (TMA row pitch)");
  ILFDT_REQUIRE(M < (1ll << 31) && N < (1ll << 31) && Kr < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "dimension too large");
  UmmaParams p{};
  p.block_n = (int)std::min<int64_t>(256, round_up(N, 64));
  p.n_step = p.block_n;
  p.out = out;
  p.ldo = ldo;
  p.alpha = alpha;
  p.accumulate = accumulate;
  p.mn_major = 1;
  p.m_rows = (int)M;
  p.n_rows = (int)N;
  p.nkb = (int)cdiv(Kr, UM_BK);
  p.num_m_tiles = (int)cdiv(M, UM_BM);
  p.num_n_tiles = (int)cdiv(N, p.n_step);
  p.idesc = make_idesc_bf16(UM_BM, p.block_n, 1, 1);
  p.nchunk = p.num_n_tiles;
  p.tiles_per_chunk = 1;
  CUtensorMap tmA, tmB;
  int32_t rc = make_tmap_bf16(&tmA, A, Kr, M, M, 64);   // rows = reduction index, inner = M; box 64 x 64
  if (rc != ILFDT_OK) return rc;
  rc = make_tmap_bf16(&tmB, B, Kr, N, N, 64);
  if (rc != ILFDT_OK) return rc;
  return launch_umma_t<EPI_STORE_T, false>(tmA, tmB, p, st);
}

// The same contraction for weight-gradient shapes (dW[N,M] = B[Kr,N]^T . A[Kr,M] with M, N of a few hundred and Kr in the
// millions: q_map's Linear layers, clip_fdt.py:88,91): the plain launch has only ceil(M/128) * ceil(N/256) tiles. Here
// the reduction is split so that tiles x splits fills the SMs once; every split writes its own [N, M] slab of `ws`
// (no atomics) and a second kernel sums the slabs in a fixed order: deterministic.
static void plan_ksplit(int64_t M, int64_t N, int64_t Kr, int& ksplit, int& kb_per) {
  const int tiles = (int)(cdiv(M, UM_BM) * cdiv(N, std::min<int64_t>(256, round_up(N, 64))));
  const int nkb = (int)cdiv(Kr, UM_BK);
  int want = std::max(1, sm_count() / std::max(1, tiles));
  want = std::min(want, std::max(1, nkb / 8));   // at least 8 k-blocks per split
  kb_per = (int)cdiv(nkb, want);
  ksplit = (int)cdiv(nkb, kb_per);
}
int64_t gemm_tn_splitk_ws_floats(int64_t M, int64_t N, int64_t Kr) {
  if (M <= 0 || N <= 0 || Kr <= 0) return 0;
  int ksplit, kb_per;
  plan_ksplit(M, N, Kr, ksplit, kb_per);
  return ksplit > 1 ? (int64_t)ksplit * N * M : 0;
}

__global__ void slab_reduce_kernel(const float* __restrict__ ws, int nslab, int64_t slab, int64_t M, int64_t ldo,
                                   int accumulate, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= slab) return;
  float t = 0.f;
  for (int s = 0; s < nslab; ++s) t += ws[s * slab + i];
  float* o = out + (i / M) * ldo + (i % M);
  *o = accumulate ? *o + t : t;
}

int32_t launch_gemm_tn_splitk_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha,
                                   int accumulate, float* out, int64_t ldo, float* ws, cudaStream_t st) {
  if (M == 0 || N == 0) return ILFDT_OK;
  int ksplit, kb_per;
  plan_ksplit(M, N, Kr, ksplit, kb_per);
  if (ksplit <= 1) return launch_gemm_tn_bf16(A, B, M, N, Kr, alpha, accumulate, out, ldo, st);
  ILFDT_REQUIRE(M % 8 == 0 && N % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE,
                "gemm_tn(bf16): M, N must be multiples of 8 (TMA row pitch)");
  ILFDT_REQUIRE(M < (1ll << 31) && N < (1ll << 31) && Kr < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "dimension too large");
  ILFDT_REQUIRE(ws != nullptr, ILFDT_E_INVALID_ARG, "gemm_tn_splitk: workspace required");
  UmmaParams p{};
  p.block_n = (int)std::min<int64_t>(256, round_up(N, 64));
  p.n_step = p.block_n;
  p.out = ws;
  p.ldo = M;
  p.alpha = alpha;
  p.accumulate = 0;
  p.mn_major = 1;
  p.m_rows = (int)M;
  p.n_rows = (int)N;
  p.nkb = (int)cdiv(Kr, UM_BK);
  p.num_m_tiles = (int)cdiv(M, UM_BM);
  p.num_n_tiles = (int)cdiv(N, p.n_step);
  p.idesc = make_idesc_bf16(UM_BM, p.block_n, 1, 1);
  p.ksplit = ksplit;
  p.kb_per = kb_per;
  p.split_stride = N * M;
  p.nchunk = p.num_n_tiles * ksplit;
  p.tiles_per_chunk = 1;
  CUtensorMap tmA, tmB;
  int32_t rc = make_tmap_bf16(&tmA, A, Kr, M, M, 64);
  if (rc != ILFDT_OK) return rc;
  rc = make_tmap_bf16(&tmB, B, Kr, N, N, 64);
  if (rc != ILFDT_OK) return rc;
  rc = launch_umma_t<EPI_STORE_T, false>(tmA, tmB, p, st);
  if (rc != ILFDT_OK) return rc;
  const int64_t slab = N * M;
  slab_reduce_kernel<<<(unsigned)cdiv(slab, 256), 256, 0, st>>>(ws, ksplit, slab, M, ldo, accumulate, out);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_ground_fwd_bf16(const void* q, const void* sd, const float* mask, int64_t b, int64_t n_tok, int64_t K,
                               int64_t D, float sqrt_d, float temperature, float* pooled, int16_t* arg,
                               cudaStream_t st) {
  if (b == 0) return ILFDT_OK;
  ILFDT_REQUIRE(n_tok >= 1 && n_tok <= 256, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_fwd(bf16): n_tok=%lld not in [1,256] (use the fp32 path)", (long long)n_tok);
  // (D > 512: both operands are streamed by the single-CTA kernel; that is the route of the 3 x bf16 fp32-grade contraction)
  ILFDT_REQUIRE(D % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE, "ground_fwd(bf16): D=%lld must be a multiple of 8", (long long)D);
  ILFDT_REQUIRE(b * n_tok < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "ground_fwd(bf16): b*n_tok too large");
  if (mask == nullptr && ground2_enabled() && ground2_supported(n_tok, D) && b >= 256 / n_tok)
    return launch_ground2(q, sd, b * n_tok, b, n_tok, K, D, 1.f / (sqrt_d * temperature), pooled, arg, nullptr, nullptr, b,
                          st);
  UmmaParams p{};
  p.spt = (int)std::min<int64_t>(256 / n_tok, b);
  p.n_step = p.spt * (int)n_tok;
  p.block_n = (int)round_up(p.n_step, 16);
  p.mask = mask;
  p.n_tok = (int)n_tok;
  p.n_samples = (int)b;
  p.pool_scale = 1.f / (sqrt_d * temperature);
  p.pooled = pooled;
  p.arg = arg;
  return launch_umma<EPI_POOLMAX>(sd, q, K, b * n_tok, D, p, st);
}

// Bucketed text grounding: q is the compacted operand of text_buckets.cu, plan its device-side description. One
// statically unrolled launch per bucket size R = min(16 g, n_tok); each launch reads its bucket's extent on the device.
bool ground_fwd_text_supported(int64_t n_tok) {
  for (int64_t g = 1; 16 * (g - 1) < n_tok; ++g) {
    const int64_t R = std::min<int64_t>(16 * g, n_tok);
    if (!(R == 16 || R == 32 || R == 48 || R == 64 || R == 49 || R == 77 || R == 196 || R == 256)) return false;
  }
  return true;
}

int32_t launch_ground_fwd_text_bf16(const void* q_compact, const void* sd, const int* plan, int64_t b, int64_t n_tok,
                                    int64_t K, int64_t D, float sqrt_d, float temperature, float* pooled, int16_t* arg,
                                    cudaStream_t st) {
  if (b == 0) return ILFDT_OK;
  ILFDT_REQUIRE(ground_fwd_text_supported(n_tok), ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_fwd_text: n_tok=%lld has a bucket size without a static epilogue", (long long)n_tok);
  // (D > 512: the 6x-wide operands of the fp32-grade 3 x bf16 contraction; both operands are streamed by the single-CTA kernel)
  ILFDT_REQUIRE(D <= 6144 && D % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE, "ground_fwd_text: D=%lld must be <= 6144, %% 8",
                (long long)D);
  ILFDT_REQUIRE(b * n_tok < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "ground_fwd_text: b*n_tok too large");
  const int ngmax = (int)cdiv(n_tok, 16);
  // small batches: one launch over the tiles of all buckets (ground_umma2.cu); large ones keep a static kernel per bucket
  if (ground2_enabled() && b <= tune_get(ILFDT_TUNE_TEXT_MERGED_MAX_B) && ground2_text_merged_supported(n_tok, D))
    return launch_ground2_text_merged(q_compact, sd, b * n_tok, b, n_tok, K, D, 1.f / (sqrt_d * temperature), pooled, arg, plan,
                                      plan + 4 * 16 + 2 * b, st);
  for (int g = 1; g <= ngmax; ++g) {
    const int R = (int)std::min<int64_t>(16 * g, n_tok);
    if (ground2_enabled() && ground2_supported(R, D)) {
      int32_t rc2 = launch_ground2(q_compact, sd, b * n_tok, b, R, K, D, 1.f / (sqrt_d * temperature), pooled, arg,
                                   plan + 4 * (g - 1), plan + 4 * 16 + 2 * b, cdiv(b, ngmax), st);
      if (rc2 != ILFDT_OK) return rc2;
      continue;
    }
    UmmaParams p{};
    p.spt = 256 / R;
    p.n_step = p.spt * R;
    p.block_n = (int)round_up(p.n_step, 16);
    p.mask = nullptr;
    p.n_tok = R;
    p.n_samples = (int)b;
    p.pool_scale = 1.f / (sqrt_d * temperature);
    p.pooled = pooled;
    p.arg = arg;
    p.dyn = plan + 4 * (g - 1);
    p.sample_map = plan + 4 * 16 + 2 * b;
    // chunking from the expected bucket size (an even share); the kernel re-derives tiles per chunk on the device
    const int est_tiles = (int)std::max<int64_t>(1, cdiv(cdiv(b, ngmax), p.spt));
    plan_items((int)cdiv(K, UM_BM), est_tiles, sm_count(), p.nchunk, p.tiles_per_chunk);
    int32_t rc = launch_umma<EPI_POOLMAX>(sd, q_compact, K, b * n_tok, D, p, st);
    if (rc != ILFDT_OK) return rc;
  }
  return ILFDT_OK;
}

int64_t nce_fwd_bf16_ws_bytes(int64_t b, int64_t B) {
  // partial (m, s) per (chunk, row) for both directions; chunks <= number of B tiles
  const int64_t max_chunks = 2 * cdiv(B, 256);
  return 2 * 2 * max_chunks * b * 4 + 256;
}

// one direction: lse[r], diag[r] of the rows scale * x[r,:] . g^T (labels label_offset + r)
int32_t launch_nce_lse_bf16(const void* x, const void* g, int64_t b, int64_t B, int64_t D, const float* scale,
                            int64_t label_offset, float* lse, float* diag, void* ws, int64_t ws_bytes, cudaStream_t st) {
  ILFDT_REQUIRE(D <= 512 && D % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE, "nce_lse(bf16): D=%lld must be <= 512, %% 8",
                (long long)D);
  ILFDT_REQUIRE(ws != nullptr && ws_bytes >= nce_fwd_bf16_ws_bytes(b, B) / 2, ILFDT_E_INVALID_ARG,
                "nce_lse(bf16): workspace too small");
  const int64_t max_chunks = 2 * cdiv(B, 256);
  if (tune_get(ILFDT_TUNE_NCE_2CTA) != 0 && ground2_enabled()) {
    // CTA-pair kernel (ground_umma2.cu): 256 logits rows resident per pair, each CTA streams half of every gathered tile
    float* part_m = reinterpret_cast<float*>(ws);
    float* part_s = part_m + max_chunks * b;
    int nchunk = 1;
    int32_t rc2 = launch_nce_lse2(x, g, b, B, D, scale, label_offset, part_m, part_s, diag, &nchunk, st);
    if (rc2 != ILFDT_OK) return rc2;
    lse_combine_kernel<<<(unsigned)cdiv(b, 256), 256, 0, st>>>(part_m, part_s, 2 * nchunk, b, lse);
    ILFDT_LAUNCH_OK();
    return ILFDT_OK;
  }
  UmmaParams p{};
  p.block_n = (int)std::min<int64_t>(256, round_up(B, 16));
  p.n_step = p.block_n;
  p.scale_dev = scale;
  p.label_offset = label_offset;
  p.part_m = reinterpret_cast<float*>(ws);
  p.part_s = p.part_m + max_chunks * b;
  p.diag = diag;
  int32_t rc = launch_umma<EPI_LSE>(x, g, b, B, D, p, st);
  if (rc != ILFDT_OK) return rc;
  lse_combine_kernel<<<(unsigned)cdiv(b, 256), 256, 0, st>>>(p.part_m, p.part_s, 2 * p.nchunk, b, lse);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_nce_fwd_bf16(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt, int64_t b,
                            int64_t B, int64_t D, const float* scale, int64_t label_offset, float* lse_i, float* lse_t,
                            float* diag_i, float* diag_t, void* ws, int64_t ws_bytes, cudaStream_t st) {
  ILFDT_REQUIRE(ws != nullptr && ws_bytes >= nce_fwd_bf16_ws_bytes(b, B), ILFDT_E_INVALID_ARG,
                "nce_fwd(bf16): workspace too small");
  const int64_t half = (nce_fwd_bf16_ws_bytes(b, B) / 2) & ~(int64_t)15;
  // logits_per_image = x_img . g_txt^T ; logits_per_text = x_txt . g_img^T  (clip_fdt.py:421-422)
  int32_t rc = launch_nce_lse_bf16(x_img, g_txt, b, B, D, scale, label_offset, lse_i, diag_i, ws, ws_bytes, st);
  if (rc != ILFDT_OK) return rc;
  return launch_nce_lse_bf16(x_txt, g_img, b, B, D, scale, label_offset, lse_t, diag_t,
                             reinterpret_cast<char*>(ws) + half, ws_bytes - half, st);
}


// N1: top-1 / top-5 hits of logits = scale * x . g^T against labels label_offset + row, without storing the logits.
__global__ void rank_hits_kernel(const int* __restrict__ rank, int64_t rows, int32_t* __restrict__ counts) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int h1 = 0, h5 = 0;
  if (r < rows) {
    const int q = rank[r];
    h1 = q < 1;
    h5 = q < 5;
  }
  h1 = warp_sum(h1);
  h5 = warp_sum(h5);
  if ((threadIdx.x & 31) == 0) {
    if (h1) atomicAdd(counts + 0, h1);
    if (h5) atomicAdd(counts + 1, h5);
  }
}

int32_t launch_nce_topk_bf16(const void* x, const void* g, int64_t b, int64_t B, int64_t D, const float* scale,
                             int64_t label_offset, const float* diag, int32_t* counts, void* ws, int64_t ws_bytes,
                             cudaStream_t st) {
  ILFDT_REQUIRE(D <= 512 && D % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE, "nce_topk(bf16): D=%lld must be <= 512, %% 8",
                (long long)D);
  ILFDT_REQUIRE(ws != nullptr && ws_bytes >= b * 4, ILFDT_E_INVALID_ARG, "nce_topk(bf16): workspace too small");
  int* rank = reinterpret_cast<int*>(ws);
  ILFDT_CUDA_OK(cudaMemsetAsync(rank, 0, (size_t)b * 4, st));
  ILFDT_CUDA_OK(cudaMemsetAsync(counts, 0, 8, st));
  UmmaParams p{};
  p.block_n = (int)std::min<int64_t>(256, round_up(B, 16));
  p.n_step = p.block_n;
  p.scale_dev = scale;
  p.label_offset = label_offset;
  p.diag = const_cast<float*>(diag);
  p.rank = rank;
  int32_t rc = launch_umma<EPI_RANK>(x, g, b, B, D, p, st);
  if (rc != ILFDT_OK) return rc;
  rank_hits_kernel<<<(unsigned)cdiv(b, 256), 256, 0, st>>>(rank, b, counts);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

}  // namespace ilfdt
