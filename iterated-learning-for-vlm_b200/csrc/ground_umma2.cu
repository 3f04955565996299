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
};

struct G2Bars {
  uint64_t full[G2_STAGES], empty[G2_STAGES], tmem_full[2], tmem_empty[2], a_full, a_empty;
  uint32_t tmem_ptr;
};

template <int NTOK>
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
      const int row = (2 * m_pair + (int)rank) * 128 + quad * 32 + lane;  // code owned by this thread
      const bool row_ok = row < p.m_rows;
      for (int n = n_begin; n < n_end; ++n, ++tile_seq) {
        const uint32_t acc = tile_seq & 1, acc_phase = (tile_seq >> 1) & 1;
        if (acc != group) continue;
        mbar_wait(&bars->tmem_full[acc], acc_phase);
        fence_after_sync();
        poolmax_static<NTOK>(tmem_base + lane_addr + acc * 256, n * p.spt, dyn_samples, row, row_ok, p.m_rows,
                             p.pool_scale, p.pooled, p.arg, p.sample_map, dyn_slot0);
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&bars->tmem_empty[acc], 0);
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

template <int NTOK>
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
    ILFDT_CUDA_OK(cudaFuncSetAttribute(ground2_kernel<NTOK>, cudaFuncAttributeMaxDynamicSharedMemorySize, G2_SMEM_BYTES));
    cfg.gridDim = dim3(2 * (sm_count() / 2));
    int n = 0;
    ILFDT_CUDA_OK(cudaOccupancyMaxActiveClusters(&n, ground2_kernel<NTOK>, &cfg));
    ILFDT_REQUIRE(n >= 1, ILFDT_E_CUDA, "ground2: no CTA pair can be resident");
    max_clusters = std::min(n, sm_count() / 2);
    configured_dev = dev;
  }
  g2_plan(p.num_m_pairs, est_tiles, max_clusters, p.nchunk, p.tiles_per_chunk);
  if (p.dyn == nullptr) p.tiles_per_chunk = (int)cdiv(p.num_n_tiles, p.nchunk);
  const int items = p.num_m_pairs * p.nchunk;
  cfg.gridDim = dim3(2 * std::min(items, max_clusters));
  ILFDT_CUDA_OK(cudaLaunchKernelEx(&cfg, ground2_kernel<NTOK>, tmA, tmB, p));
  return ILFDT_OK;
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

}  // namespace ilfdt
