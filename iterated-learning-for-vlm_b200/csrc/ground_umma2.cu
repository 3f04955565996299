// a1 on a CTA pair: the grounding score kernel (codebook tile x token rows -> per-sample max / argmax) with
// tcgen05 cta_group::2.
//
// Why: ncu (r01) shows umma_kernel<EPI_POOLMAX> moving 12 TB/s from L2 into the SMs -- the measured L2 cap -- with the
// tensor pipe at 74-77 %: with a 128-code tile resident per SM every token row is delivered to every SM once per
// 128 codes. A CTA pair computes a 256-code x N-token tile with ONE delivery of the token rows per pair: each CTA keeps
// its own 128 codes resident (its 128 TMEM lanes), loads only N/2 of the tile's token rows per stage, and the pair's
// tcgen05.mma.cta_group::2 (M = 256) reads B from both CTAs' shared memory. L2->SM bytes per FLOP halve, the stage
// shrinks to 16 KB so six stages fit where three did.
//
// Protocol (leader = CTA rank 0 of the pair):
//   producer (one thread per CTA)   waits its own empty[stage]; TMA (cta_group::2 form) into its own smem, completing
//                                   transaction bytes on the LEADER's full[stage]; the leader expects both halves
//   MMA issuer (leader only)        waits full[stage]; tcgen05.mma.cta_group::2; tcgen05.commit multicast to empty[stage]
//                                   of both CTAs; per tile a multicast commit to tmem_full[acc] of both CTAs
//   epilogue warps (both CTAs)      each CTA drains its own TMEM lanes (its 128 codes) with the same statically
//                                   unrolled max/argmax body; every epilogue warp of both CTAs arrives on the
//                                   LEADER's tmem_empty[acc] (remote arrive for the peer)
#include "umma_pair.cuh"
#include "poolmax.cuh"
#include "launch.h"
#include <algorithm>
#include <cstdlib>

namespace ilfdt {

using namespace umma;

constexpr int G2_THREADS = 320;
constexpr int G2_A_BYTES = 128 * 64 * 2;      // 16 KB per k-block of this CTA's resident codebook half
constexpr int G2_B_BYTES = 128 * 64 * 2;      // 16 KB per stage: this CTA's half (<= 128 rows) of the B tile
constexpr int G2_STAGES = 6;
constexpr int G2_DATA_BYTES = 8 * G2_A_BYTES + G2_STAGES * G2_B_BYTES;   // 224 KB
constexpr int G2_SMEM_BYTES = G2_DATA_BYTES + 1024 + 256;

struct G2Params {
  int num_m_pairs, nchunk, tiles_per_chunk, num_n_tiles, nkb;
  int block_n;      // UMMA N (multiple of 16, <= 256); each CTA holds block_n / 2 rows of the tile
  int n_step;       // rows of B advanced per tile
  int m_rows;       // K (codes)
  int spt, n_samples;
  float pool_scale;
  float* pooled;
  int16_t* arg;
  const int* dyn;          // {samples, first row, first slot, rows per sample} (text buckets) or nullptr
  const int* sample_map;
  uint32_t idesc;
  // EPI2 == G2_EPI_LSE (InfoNCE forward, a8 + a10): lanes = local logits rows, columns = gathered rows; per row an online
  // base-2 log-sum-exp over the item's columns and the label logit; one partial per (chunk, epilogue group)
  const float* scale_dev;
  long long label_offset;
  float* part_m;
  float* part_s;
  float* diag;
  int n_rows;              // gathered rows (B)
};

enum { G2_EPI_POOLMAX = 0, G2_EPI_LSE = 1 };

__device__ __forceinline__ float g2_ex2(float x) {   // 2^x for x <= 0 (or -inf): results in [0, 1], tiny ones flushed to 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct G2Bars {
  uint64_t full[G2_STAGES], empty[G2_STAGES], tmem_full[2], tmem_empty[2], a_full, a_empty;
  uint32_t tmem_ptr;
};

template <int NTOK, int EPI2 = G2_EPI_POOLMAX>
__global__ void __launch_bounds__(G2_THREADS, 1)
ground2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const G2Params p) {
  extern __shared__ uint8_t g2_smem_raw[];
  // the pair's two CTAs must use identical offsets: the dynamic window starts at the same offset in both
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(g2_smem_raw) + 1023) & ~(uintptr_t)1023);
  G2Bars* bars = reinterpret_cast<G2Bars*>(smem + G2_DATA_BYTES);
  uint8_t* ring = smem + 8 * G2_A_BYTES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  int dyn_samples = p.n_samples, dyn_row0 = 0, dyn_slot0 = 0, num_n_tiles = p.num_n_tiles, tiles_per_chunk = p.tiles_per_chunk;
  int num_items = p.num_m_pairs * p.nchunk;
  if (p.dyn != nullptr) {
    dyn_samples = __ldg(p.dyn);
    dyn_row0 = __ldg(p.dyn + 1);
    dyn_slot0 = __ldg(p.dyn + 2);
    num_n_tiles = (dyn_samples + p.spt - 1) / p.spt;
    tiles_per_chunk = (num_n_tiles + p.nchunk - 1) / p.nchunk;
    num_items = num_n_tiles == 0 ? 0 : p.num_m_pairs * ((num_n_tiles + tiles_per_chunk - 1) / tiles_per_chunk);
  }
  const int half_n = p.block_n >> 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < G2_STAGES; ++i) { mbar_init(&bars->full[i], 1); mbar_init(&bars->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->tmem_full[i], 1); mbar_init(&bars->tmem_empty[i], 8); }
    mbar_init(&bars->a_full, 1);
    mbar_init(&bars->a_empty, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    tmem_alloc_pair(&bars->tmem_ptr, 512);
    tmem_relinquish_pair();
  }
  fence_before_sync();
  cluster_sync_all();   // barrier inits and the TMEM allocation of BOTH CTAs are visible before any remote access
  fence_after_sync();
  const uint32_t tmem_base = bars->tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer (one thread in each CTA) =====================
    if (elect_one()) {
      uint32_t stage = 0, phase = 0, it = 0;
      for (int item = cluster_id; item < num_items; item += num_clusters, ++it) {
        const int m_pair = item % p.num_m_pairs, chunk = item / p.num_m_pairs;
        const int n_begin = chunk * tiles_per_chunk, n_end = min(num_n_tiles, n_begin + tiles_per_chunk);
        mbar_wait(&bars->a_empty, (it & 1) ^ 1);
        if (leader) mbar_expect_tx(&bars->a_full, 2u * (uint32_t)p.nkb * G2_A_BYTES);
        for (int kb = 0; kb < p.nkb; ++kb)
          tma_load_2d_pair(smem + kb * G2_A_BYTES, &tmA, &bars->a_full, kb * 64, (2 * m_pair + (int)rank) * 128, kEvictLast);
        for (int n = n_begin; n < n_end; ++n) {
          const int r0 = dyn_row0 + n * p.n_step + (int)rank * half_n;
          for (int kb = 0; kb < p.nkb; ++kb) {
            mbar_wait(&bars->empty[stage], phase ^ 1);
            if (leader) mbar_expect_tx(&bars->full[stage], (uint32_t)p.block_n * 128u);
            tma_load_2d_pair(ring + stage * G2_B_BYTES, &tmB, &bars->full[stage], kb * 64, r0, kEvictNormal);
            if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && elect_one()) {
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0, it = 0;
      for (int item = cluster_id; item < num_items; item += num_clusters, ++it) {
        const int chunk = item / p.num_m_pairs;
        const int n_begin = chunk * tiles_per_chunk, n_end = min(num_n_tiles, n_begin + tiles_per_chunk);
        mbar_wait(&bars->a_full, it & 1);
        fence_after_sync();
        for (int n = n_begin; n < n_end; ++n) {
          mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
          fence_after_sync();
          const uint32_t tmem_d = tmem_base + acc * 256;
          for (int kb = 0; kb < p.nkb; ++kb) {
            mbar_wait(&bars->full[stage], phase);
            fence_after_sync();
            const uint64_t adesc = make_smem_desc_sw128(smem_u32(smem + kb * G2_A_BYTES));
            const uint64_t bdesc = make_smem_desc_sw128(smem_u32(ring + stage * G2_B_BYTES));
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              mma_f16_pair(tmem_d, advance_desc_k(adesc, k4 * 32), advance_desc_k(bdesc, k4 * 32), p.idesc,
                           (kb | k4) != 0 ? 1u : 0u);
            mma_commit_pair(&bars->empty[stage]);
            if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
          }
          mma_commit_pair(&bars->tmem_full[acc]);
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1;
        }
        mma_commit_pair(&bars->a_empty);
      }
    }
  } else {
    // ===================== epilogue warps (2..9) of both CTAs: group g drains accumulator g =====================
    const int quad = warp & 3;
    const uint32_t group = (uint32_t)(warp - 2) >> 2;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    uint32_t tile_seq = 0;
    for (int item = cluster_id; item < num_items; item += num_clusters) {
      const int m_pair = item % p.num_m_pairs, chunk = item / p.num_m_pairs;
      const int n_begin = chunk * tiles_per_chunk, n_end = min(num_n_tiles, n_begin + tiles_per_chunk);
      const int row = (2 * m_pair + (int)rank) * 128 + quad * 32 + lane;  // code (LSE: logits row) owned by this thread
      const bool row_ok = row < p.m_rows;
      float run_m = -INFINITY, run_s = 0.f;   // G2_EPI_LSE: running maximum and sum of the item's columns, base-2 domain
      for (int n = n_begin; n < n_end; ++n, ++tile_seq) {
        const uint32_t acc = tile_seq & 1, acc_phase = (tile_seq >> 1) & 1;
        if (acc != group) continue;
        mbar_wait(&bars->tmem_full[acc], acc_phase);
        fence_after_sync();
        if (EPI2 == G2_EPI_POOLMAX) {
          poolmax_static<NTOK>(tmem_base + lane_addr + acc * 256, n * p.spt, dyn_samples, row, row_ok, p.m_rows,
                               p.pool_scale, p.pooled, p.arg, p.sample_map, dyn_slot0);
        } else {
          const float scale = __ldg(p.scale_dev);
          const float scale_log2 = scale * 1.4426950408889634f;
          const uint32_t taddr = tmem_base + lane_addr + acc * 256;
          const int n0 = n * p.n_step;
          const int ncols = min(p.n_step, p.n_rows - n0);
          const long long label = p.label_offset + row;
          const long long wl0 = p.label_offset + (2 * m_pair + (int)rank) * 128 + quad * 32;  // labels of this warp: [wl0, wl0 + 32)
          for (int c0 = 0; c0 < ncols; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_ld_wait();
            // lean form (the kernel is epilogue-bound otherwise: ~8 issued instructions per logit with exp2f on scaled,
            // masked copies): the chunk maximum is taken on the raw accumulators (scale > 0 keeps the order), and each
            // term is one FFMA (scale and shift folded), one MUFU.EX2, one FADD
            float cm = __uint_as_float(v[0]);
            if (c0 + 32 <= ncols) {
#pragma unroll
              for (int j = 1; j < 32; ++j) cm = fmaxf(cm, __uint_as_float(v[j]));
            } else {
#pragma unroll
              for (int j = 1; j < 32; ++j)
                if (c0 + j < ncols) cm = fmaxf(cm, __uint_as_float(v[j]));
            }
            const float nm = fmaxf(run_m, cm * scale_log2);
            const float shift = -nm;
            float sum = 0.f;
            if (c0 + 32 <= ncols) {
#pragma unroll
              for (int j = 0; j < 32; ++j) sum += g2_ex2(fmaf(__uint_as_float(v[j]), scale_log2, shift));
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (c0 + j < ncols) sum += g2_ex2(fmaf(__uint_as_float(v[j]), scale_log2, shift));
            }
            run_s = run_s * g2_ex2(run_m - nm) + sum;
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
        if (lane == 0) mbar_arrive_cluster(&bars->tmem_empty[acc], 0);
      }
      if (EPI2 == G2_EPI_LSE && row_ok) {  // one partial per (chunk, epilogue group); a group without a tile leaves (-inf, 0)
        p.part_m[(int64_t)(chunk * 2 + (int)group) * p.m_rows + row] = run_m;
        p.part_s[(int64_t)(chunk * 2 + (int)group) * p.m_rows + row] = run_s;
      }
    }
  }

  fence_before_sync();
  cluster_sync_all();   // neither CTA may free TMEM or exit while its peer can still touch its barriers / smem
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

// ---- all text buckets in ONE launch (small batches) ---------------------------------------------------
// Five launches of ground2_kernel<R> (one per bucket of the compacted text operand, text_buckets.cu) each end in their own
// partial wave, and the host cannot size their items because the bucket populations only exist on the device: at 256 rows
// the five launches take 185 us for 156 us of tensor work, at 64 rows 88 us for 40. This kernel walks the tiles of ALL
// buckets as one flat list: work unit w = code-tile pair * T + t, t = 0 .. T-1 over the buckets' token tiles in bucket
// order (T read from the plan on the device); cluster c takes the contiguous range [c W / C, (c + 1) W / C) -- equal to
// within one tile, and the resident codebook tile changes at most twice per cluster. What varies per bucket (rows per
// sample, samples per tile, UMMA N, the TMA box) is looked up per tile; the epilogue dispatches to the same statically
// unrolled bodies. Results are the same bits as the per-bucket launches (same MMAs per tile, same epilogue).
constexpr int G2T_MAX_BUCKETS = 5;

struct G2TextMaps {
  CUtensorMap b[G2T_MAX_BUCKETS];   // one per bucket: box = block_n / 2 rows x 64
};

struct G2TextParams {
  int ng;                      // buckets
  int num_m_pairs, nkb, m_rows;
  int rows[G2T_MAX_BUCKETS];   // token rows per sample of the bucket (R)
  int spt[G2T_MAX_BUCKETS], n_step[G2T_MAX_BUCKETS], block_n[G2T_MAX_BUCKETS];
  uint32_t idesc[G2T_MAX_BUCKETS];
  float pool_scale;
  float* pooled;
  int16_t* arg;
  const int* plan;             // 4 ints per bucket {samples, first row, first slot, rows per sample}
  const int* sample_map;
};

struct G2TextSched {
  int tile0[G2T_MAX_BUCKETS + 1];   // first flat tile index of each bucket; [ng] = T
  int samples[G2T_MAX_BUCKETS], row0[G2T_MAX_BUCKETS], slot0[G2T_MAX_BUCKETS];
};

__device__ __forceinline__ int g2t_bucket_of(const G2TextSched& sc, int ng, int t) {
  int g = 0;
#pragma unroll
  for (int i = 1; i < G2T_MAX_BUCKETS; ++i)
    if (i < ng && t >= sc.tile0[i]) g = i;
  return g;
}

__global__ void __launch_bounds__(G2_THREADS, 1)
ground2_text_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ G2TextMaps tmB, const G2TextParams p) {
  extern __shared__ uint8_t g2_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(g2_smem_raw) + 1023) & ~(uintptr_t)1023);
  G2Bars* bars = reinterpret_cast<G2Bars*>(smem + G2_DATA_BYTES);
  uint8_t* ring = smem + 8 * G2_A_BYTES;
  __shared__ G2TextSched sc;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < G2_STAGES; ++i) { mbar_init(&bars->full[i], 1); mbar_init(&bars->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->tmem_full[i], 1); mbar_init(&bars->tmem_empty[i], 8); }
    mbar_init(&bars->a_full, 1);
    mbar_init(&bars->a_empty, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmA);
    int t = 0;
    for (int g = 0; g < p.ng; ++g) {
      const int n = __ldg(p.plan + 4 * g);
      sc.samples[g] = n;
      sc.row0[g] = __ldg(p.plan + 4 * g + 1);
      sc.slot0[g] = __ldg(p.plan + 4 * g + 2);
      sc.tile0[g] = t;
      t += (n + p.spt[g] - 1) / p.spt[g];
    }
    for (int g = p.ng; g <= G2T_MAX_BUCKETS; ++g) sc.tile0[g] = t;
  }
  if (warp == 1) {
    tmem_alloc_pair(&bars->tmem_ptr, 512);
    tmem_relinquish_pair();
  }
  fence_before_sync();
  __syncthreads();
  cluster_sync_all();
  fence_after_sync();
  const uint32_t tmem_base = bars->tmem_ptr;
  const int T = sc.tile0[p.ng];
  const long long W = (long long)p.num_m_pairs * T;
  const int w_begin = (int)(W * cluster_id / num_clusters), w_end = (int)(W * (cluster_id + 1) / num_clusters);

  if (warp == 0) {
    // ===================== TMA producer (one thread in each CTA) =====================
    if (elect_one()) {
      uint32_t stage = 0, phase = 0, it = 0;
      int cur_m = -1;
      for (int w = w_begin; w < w_end; ++w) {
        const int m_pair = w / T, t = w - m_pair * T;
        if (m_pair != cur_m) {
          cur_m = m_pair;
          mbar_wait(&bars->a_empty, (it & 1) ^ 1);
          if (leader) mbar_expect_tx(&bars->a_full, 2u * (uint32_t)p.nkb * G2_A_BYTES);
          for (int kb = 0; kb < p.nkb; ++kb)
            tma_load_2d_pair(smem + kb * G2_A_BYTES, &tmA, &bars->a_full, kb * 64, (2 * m_pair + (int)rank) * 128, kEvictLast);
          ++it;
        }
        const int g = g2t_bucket_of(sc, p.ng, t);
        const int n = t - sc.tile0[g];
        const int r0 = sc.row0[g] + n * p.n_step[g] + (int)rank * (p.block_n[g] >> 1);
        for (int kb = 0; kb < p.nkb; ++kb) {
          mbar_wait(&bars->empty[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&bars->full[stage], (uint32_t)p.block_n[g] * 128u);
          tma_load_2d_pair(ring + stage * G2_B_BYTES, &tmB.b[g], &bars->full[stage], kb * 64, r0, kEvictNormal);
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && elect_one()) {
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0, it = 0;
      int cur_m = -1;
      for (int w = w_begin; w < w_end; ++w) {
        const int m_pair = w / T, t = w - m_pair * T;
        if (m_pair != cur_m) {
          if (cur_m >= 0) mma_commit_pair(&bars->a_empty);   // every MMA that reads the previous resident tile has been issued
          cur_m = m_pair;
          mbar_wait(&bars->a_full, it & 1);
          fence_after_sync();
          ++it;
        }
        const int g = g2t_bucket_of(sc, p.ng, t);
        const uint32_t idesc = p.idesc[g];
        mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
        fence_after_sync();
        const uint32_t tmem_d = tmem_base + acc * 256;
        for (int kb = 0; kb < p.nkb; ++kb) {
          mbar_wait(&bars->full[stage], phase);
          fence_after_sync();
          const uint64_t adesc = make_smem_desc_sw128(smem_u32(smem + kb * G2_A_BYTES));
          const uint64_t bdesc = make_smem_desc_sw128(smem_u32(ring + stage * G2_B_BYTES));
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)
            mma_f16_pair(tmem_d, advance_desc_k(adesc, k4 * 32), advance_desc_k(bdesc, k4 * 32), idesc, (kb | k4) != 0 ? 1u : 0u);
          mma_commit_pair(&bars->empty[stage]);
          if (++stage == G2_STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit_pair(&bars->tmem_full[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ===================== epilogue warps (2..9) of both CTAs: group g drains accumulator g =====================
    const int quad = warp & 3;
    const uint32_t group = (uint32_t)(warp - 2) >> 2;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    uint32_t tile_seq = 0;
    for (int w = w_begin; w < w_end; ++w, ++tile_seq) {
      const uint32_t acc = tile_seq & 1, acc_phase = (tile_seq >> 1) & 1;
      if (acc != group) continue;
      const int m_pair = w / T, t = w - m_pair * T;
      const int g = g2t_bucket_of(sc, p.ng, t);
      const int n = t - sc.tile0[g];
      const int row = (2 * m_pair + (int)rank) * 128 + quad * 32 + lane;  // code owned by this thread
      const bool row_ok = row < p.m_rows;
      const uint32_t taddr = tmem_base + lane_addr + acc * 256;
      const int s0 = n * p.spt[g], ns = sc.samples[g], slot0 = sc.slot0[g];
      mbar_wait(&bars->tmem_full[acc], acc_phase);
      fence_after_sync();
      switch (p.rows[g]) {
        case 16: poolmax_static<16>(taddr, s0, ns, row, row_ok, p.m_rows, p.pool_scale, p.pooled, p.arg, p.sample_map, slot0); break;
        case 32: poolmax_static<32>(taddr, s0, ns, row, row_ok, p.m_rows, p.pool_scale, p.pooled, p.arg, p.sample_map, slot0); break;
        case 48: poolmax_static<48>(taddr, s0, ns, row, row_ok, p.m_rows, p.pool_scale, p.pooled, p.arg, p.sample_map, slot0); break;
        case 64: poolmax_static<64>(taddr, s0, ns, row, row_ok, p.m_rows, p.pool_scale, p.pooled, p.arg, p.sample_map, slot0); break;
        default: poolmax_static<77>(taddr, s0, ns, row, row_ok, p.m_rows, p.pool_scale, p.pooled, p.arg, p.sample_map, slot0); break;
      }
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&bars->tmem_empty[acc], 0);
    }
  }

  fence_before_sync();
  cluster_sync_all();   // neither CTA may free TMEM or exit while its peer can still touch its barriers / smem
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

// ---- host side --------------------------------------------------------------------------------------
static void g2_plan(int num_m, int num_n_tiles, int workers, int& nchunk, int& tiles_per_chunk) {
  int best_c = 1;
  double best_eff = -1.0;
  const int max_c = std::max(1, std::min(num_n_tiles, 4096));
  for (int c = 1; c <= max_c; ++c) {
    const int tpc = (int)cdiv(num_n_tiles, c);
    const int real_c = (int)cdiv(num_n_tiles, tpc);
    const int64_t items = (int64_t)num_m * real_c;
    double eff = (double)items / (double)(cdiv(items, workers) * workers);
    eff *= (double)tpc / (double)(tpc + 0.5);   // every item pays one resident-tile load
    if (eff > best_eff + 1e-9) { best_eff = eff; best_c = real_c; }
    if (tpc == 1) break;
  }
  nchunk = best_c;
  tiles_per_chunk = (int)cdiv(num_n_tiles, nchunk);
  nchunk = (int)cdiv(num_n_tiles, tiles_per_chunk);
}

template <int NTOK, int EPI2 = G2_EPI_POOLMAX>
static int32_t launch_g2_t(const CUtensorMap& tmA, const CUtensorMap& tmB, G2Params& p, int est_tiles, cudaStream_t st) {
  static thread_local int configured_dev = -1;
  static thread_local int max_clusters = 0;
  int dev = 0;
  ILFDT_CUDA_OK(cudaGetDevice(&dev));
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(G2_THREADS);
  cfg.dynamicSmemBytes = G2_SMEM_BYTES;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (configured_dev != dev) {
    ILFDT_CUDA_OK(cudaFuncSetAttribute(ground2_kernel<NTOK, EPI2>, cudaFuncAttributeMaxDynamicSharedMemorySize, G2_SMEM_BYTES));
    cfg.gridDim = dim3(2 * (sm_count() / 2));
    int n = 0;
    ILFDT_CUDA_OK(cudaOccupancyMaxActiveClusters(&n, ground2_kernel<NTOK, EPI2>, &cfg));
    ILFDT_REQUIRE(n >= 1, ILFDT_E_CUDA, "ground2: no CTA pair can be resident");
    max_clusters = std::min(n, sm_count() / 2);
    configured_dev = dev;
  }
  g2_plan(p.num_m_pairs, est_tiles, max_clusters, p.nchunk, p.tiles_per_chunk);
  if (p.dyn == nullptr) p.tiles_per_chunk = (int)cdiv(p.num_n_tiles, p.nchunk);
  const int items = p.num_m_pairs * p.nchunk;
  cfg.gridDim = dim3(2 * std::min(items, max_clusters));
  ILFDT_CUDA_OK(cudaLaunchKernelEx(&cfg, ground2_kernel<NTOK, EPI2>, tmA, tmB, p));
  return ILFDT_OK;
}

// InfoNCE forward on the pair kernel: per local row r the partial (max, sum) of 2^(scale log2(e) x[r,:] . g[c,:]) over each
// chunk of gathered rows c, and the label logit diag[r] = scale x[r,:] . g[label_offset + r,:]. Partials: [2 * nchunk, b]
// (chunk-major, two epilogue groups per chunk) at part_m / part_s; nchunk is returned for the combine kernel.
int32_t launch_nce_lse2(const void* x, const void* g, int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset,
                        float* part_m, float* part_s, float* diag, int* nchunk_out, cudaStream_t st) {
  G2Params p{};
  p.block_n = (int)std::min<int64_t>(256, round_up(B, 16));
  p.n_step = p.block_n;
  p.spt = 1;
  p.nkb = (int)cdiv(D, 64);
  p.m_rows = (int)b;
  p.num_m_pairs = (int)cdiv(b, 256);
  p.n_rows = (int)B;
  p.num_n_tiles = (int)cdiv(B, p.n_step);
  p.n_samples = p.num_n_tiles;
  p.scale_dev = scale;
  p.label_offset = label_offset;
  p.part_m = part_m;
  p.part_s = part_s;
  p.diag = diag;
  p.idesc = make_idesc_bf16(256, p.block_n);
  CUtensorMap tmA, tmB;
  int32_t rc = make_tmap_bf16(&tmA, x, b, D, D, 128);
  if (rc != ILFDT_OK) return rc;
  if ((rc = make_tmap_bf16(&tmB, g, B, D, D, p.block_n / 2)) != ILFDT_OK) return rc;
  rc = launch_g2_t<16, G2_EPI_LSE>(tmA, tmB, p, p.num_n_tiles, st);
  *nchunk_out = p.nchunk;
  return rc;
}

bool ground2_enabled() {
  return tune_get(ILFDT_TUNE_GROUND_2CTA) != 0;
}
bool ground2_supported(int64_t n_tok, int64_t D) {
  return D <= 512 && D % 8 == 0 &&
         (n_tok == 16 || n_tok == 32 || n_tok == 48 || n_tok == 64 || n_tok == 49 || n_tok == 77 || n_tok == 196 ||
          n_tok == 256);
}

// q: [rows, D] bf16 (full-width [b*n_tok, D] or the compacted text operand), dyn / sample_map as in UmmaParams.
int32_t launch_ground2(const void* q, const void* sd, int64_t q_rows, int64_t n_samples, int64_t n_tok, int64_t K, int64_t D,
                       float pool_scale, float* pooled, int16_t* arg, const int* dyn, const int* sample_map,
                       int64_t est_samples, cudaStream_t st) {
  G2Params p{};
  p.spt = (int)(256 / n_tok);
  p.n_step = p.spt * (int)n_tok;
  p.block_n = (int)round_up(p.n_step, 16);
  p.nkb = (int)cdiv(D, 64);
  p.m_rows = (int)K;
  p.num_m_pairs = (int)cdiv(K, 256);
  p.n_samples = (int)n_samples;
  p.num_n_tiles = (int)cdiv(n_samples, p.spt);
  p.pool_scale = pool_scale;
  p.pooled = pooled;
  p.arg = arg;
  p.dyn = dyn;
  p.sample_map = sample_map;
  p.idesc = make_idesc_bf16(256, p.block_n);
  CUtensorMap tmA, tmB;
  int32_t rc = make_tmap_bf16(&tmA, sd, K, D, D, 128);
  if (rc != ILFDT_OK) return rc;
  rc = make_tmap_bf16(&tmB, q, q_rows, D, D, p.block_n / 2);
  if (rc != ILFDT_OK) return rc;
  const int est_tiles = (int)std::max<int64_t>(1, cdiv(est_samples, p.spt));
  switch (n_tok) {
    case 16: return launch_g2_t<16>(tmA, tmB, p, est_tiles, st);
    case 32: return launch_g2_t<32>(tmA, tmB, p, est_tiles, st);
    case 48: return launch_g2_t<48>(tmA, tmB, p, est_tiles, st);
    case 64: return launch_g2_t<64>(tmA, tmB, p, est_tiles, st);
    case 49: return launch_g2_t<49>(tmA, tmB, p, est_tiles, st);
    case 77: return launch_g2_t<77>(tmA, tmB, p, est_tiles, st);
    case 196: return launch_g2_t<196>(tmA, tmB, p, est_tiles, st);
    case 256: return launch_g2_t<256>(tmA, tmB, p, est_tiles, st);
    default: break;
  }
  return fail(ILFDT_E_UNSUPPORTED_SHAPE, "ground2: n_tok=%lld has no static epilogue", (long long)n_tok);
}

// All buckets of the compacted text operand in one launch (ground2_text_kernel). n_tok's buckets must be 16 / 32 / 48 / 64 / 77.
bool ground2_text_merged_supported(int64_t n_tok, int64_t D) {
  if (!(D <= 512 && D % 8 == 0)) return false;
  if (n_tok > 77 || cdiv(n_tok, 16) > G2T_MAX_BUCKETS) return false;
  const int64_t last = n_tok;
  return last == 16 || last == 32 || last == 48 || last == 64 || last == 77;
}

int32_t launch_ground2_text_merged(const void* q, const void* sd, int64_t q_rows, int64_t b, int64_t n_tok, int64_t K, int64_t D,
                                   float pool_scale, float* pooled, int16_t* arg, const int* plan, const int* sample_map,
                                   cudaStream_t st) {
  G2TextParams p{};
  G2TextMaps maps{};
  p.ng = (int)cdiv(n_tok, 16);
  p.nkb = (int)cdiv(D, 64);
  p.m_rows = (int)K;
  p.num_m_pairs = (int)cdiv(K, 256);
  p.pool_scale = pool_scale;
  p.pooled = pooled;
  p.arg = arg;
  p.plan = plan;
  p.sample_map = sample_map;
  CUtensorMap tmA;
  int32_t rc = make_tmap_bf16(&tmA, sd, K, D, D, 128);
  if (rc != ILFDT_OK) return rc;
  for (int g = 0; g < p.ng; ++g) {
    const int R = (int)std::min<int64_t>(16 * (g + 1), n_tok);
    p.rows[g] = R;
    p.spt[g] = 256 / R;
    p.n_step[g] = p.spt[g] * R;
    p.block_n[g] = (int)round_up(p.n_step[g], 16);
    p.idesc[g] = make_idesc_bf16(256, p.block_n[g]);
    if ((rc = make_tmap_bf16(&maps.b[g], q, q_rows, D, D, p.block_n[g] / 2)) != ILFDT_OK) return rc;
  }
  static thread_local int configured_dev = -1;
  static thread_local int max_clusters = 0;
  int dev = 0;
  ILFDT_CUDA_OK(cudaGetDevice(&dev));
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.blockDim = dim3(G2_THREADS);
  cfg.dynamicSmemBytes = G2_SMEM_BYTES;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (configured_dev != dev) {
    ILFDT_CUDA_OK(cudaFuncSetAttribute(ground2_text_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G2_SMEM_BYTES));
    cfg.gridDim = dim3(2 * (sm_count() / 2));
    int n = 0;
    ILFDT_CUDA_OK(cudaOccupancyMaxActiveClusters(&n, ground2_text_kernel, &cfg));
    ILFDT_REQUIRE(n >= 1, ILFDT_E_CUDA, "ground2_text: no CTA pair can be resident");
    max_clusters = std::min(n, sm_count() / 2);
    configured_dev = dev;
  }
  // upper bound of the flat tile count (every sample in the bucket with the fewest samples per tile) caps the grid
  const int64_t max_tiles = cdiv(b, 256 / std::min<int64_t>(n_tok, 77)) + p.ng;
  cfg.gridDim = dim3(2 * (unsigned)std::min<int64_t>(max_clusters, std::max<int64_t>(1, p.num_m_pairs * max_tiles)));
  ILFDT_CUDA_OK(cudaLaunchKernelEx(&cfg, ground2_text_kernel, tmA, maps, p));
  return ILFDT_OK;
}

}  // namespace ilfdt
