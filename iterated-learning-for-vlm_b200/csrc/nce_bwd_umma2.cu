// a13 (loss part) on a CTA pair: the fused InfoNCE backward of nce_bwd_umma.cu with tcgen05 cta_group::2.
//
// Why: ncu (r01) shows the single-CTA kernel waiting on its operand feed -- per 64-row stream tile one SM takes 96 KB
// from L2, writes them to shared memory and reads 224 KB of MMA operands back out of it, all through the same
// 128 B/clk shared-memory port -- with the tensor pipe 46 % busy. A CTA pair computes a 256-lane-row tile: each CTA
// keeps its own 128 lane rows resident (its 128 TMEM lanes) and loads only HALF of every streamed operand
// (32 of the tile's 64 stream rows for S = XL.Y^T, 128 of the 256 D-columns for dOut += P.Y); the pair's
// tcgen05.mma.cta_group::2 (M = 256) reads B from both CTAs' shared memory. Per CTA and tile: 48 KB from L2.
//
// Protocol (leader = CTA rank 0 of the pair; same roles as ground_umma2.cu):
//   TMA warps (both CTAs)       wait their own *_empty; load into their own smem; transaction bytes on the LEADER's
//                               *_full (the leader expects both halves)
//   MMA issuers (leader only)   two threads: MMA1 (S tiles, as far ahead as the two S buffers allow) and MMA2 (dOut);
//                               tcgen05.commit multicast to y_empty / yt_empty / s_full / p_empty / d_full / xl_empty
//                               of BOTH CTAs
//   softmax warps (both CTAs)   each CTA converts its own 128 rows: S (own TMEM lanes) -> P (own TMEM lanes, the A
//                               operand of MMA2); arrivals on the LEADER's s_empty / p_full / d_empty
//                               (remote arrive from the peer), counts doubled
#include <algorithm>
#include <cstdlib>
#include "umma_pair.cuh"
#include "nce_bwd_common.cuh"

namespace ilfdt {

using namespace umma;

constexpr int N2_THREADS = 384;  // Y TMA warp, MMA1 warp, 2 softmax groups of 4 warps, YT TMA warp, MMA2 warp
constexpr int N2_XL_BYTES = 8 * 16384;          // this CTA's 128 lane rows, all k-blocks: 128 KB
constexpr int N2_Y_BOX = 32 * 64 * 2;           // this CTA's 32 of the tile's 64 stream rows x 64 k: 4 KB
constexpr int N2_Y_KPS = 4;                     // k-blocks (boxes) per ring stage: one barrier round trip per 16 MMAs
constexpr int N2_Y_STAGE = N2_Y_KPS * N2_Y_BOX;  // 16 KB
constexpr int N2_Y_STAGES = 3;
constexpr int N2_YT_STAGE = 2 * 64 * 64 * 2;    // this CTA's 128 D-columns (2 boxes of 64) x 64 stream rows: 16 KB
constexpr int N2_YT_STAGES = 3;
constexpr int N2_OFF_Y = N2_XL_BYTES;
constexpr int N2_OFF_YT = N2_OFF_Y + N2_Y_STAGES * N2_Y_STAGE;
constexpr int N2_OFF_BARS = N2_OFF_YT + N2_YT_STAGES * N2_YT_STAGE;   // 208 KB
constexpr int N2_SMEM_BYTES = N2_OFF_BARS + 1024 + 2048;
constexpr int N2_TMEM_S = 256;  // S buffers at columns 256 and 320
constexpr int N2_TMEM_P = 384;  // P buffers (bf16 pairs) at columns 384 and 416

struct Nce2Bars {
  uint64_t xl_full, xl_empty;
  uint64_t y_full[N2_Y_STAGES], y_empty[N2_Y_STAGES];
  uint64_t yt_full[N2_YT_STAGES], yt_empty[N2_YT_STAGES];
  uint64_t s_full[2], s_empty[2];
  uint64_t p_full[2], p_empty[2];
  uint64_t d_full, d_empty;
  uint32_t tmem_ptr;
  NceColBuf cols[4];   // [group * 2 + tile parity]: double buffered, so one named barrier per tile is enough
};
static_assert(sizeof(Nce2Bars) <= 2048, "barrier block outgrew its shared-memory slot");
static_assert(N2_SMEM_BYTES <= 232448, "over the sm_100 dynamic shared memory limit");

template <int MODE>
__global__ void __launch_bounds__(N2_THREADS, 1)
nce_bwd2_kernel(const __grid_constant__ CUtensorMap tmXL, const __grid_constant__ CUtensorMap tmY32,
                const __grid_constant__ CUtensorMap tmY64, const NceBwdParams p) {
  extern __shared__ uint8_t n2_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(n2_smem_raw) + 1023) & ~(uintptr_t)1023);
  Nce2Bars* bars = reinterpret_cast<Nce2Bars*>(smem + N2_OFF_BARS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_items = p.num_lane_tiles * p.num_halves * p.nsplit;   // num_lane_tiles counts PAIRS of 128-row tiles here

  if (threadIdx.x == 0) {
    mbar_init(&bars->xl_full, 1);
    mbar_init(&bars->xl_empty, 1);
    for (int i = 0; i < N2_Y_STAGES; ++i) { mbar_init(&bars->y_full[i], 1); mbar_init(&bars->y_empty[i], 1); }
    for (int i = 0; i < N2_YT_STAGES; ++i) { mbar_init(&bars->yt_full[i], 1); mbar_init(&bars->yt_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->s_empty[i], 8);    // 4 softmax warps of the owning group in each CTA
      mbar_init(&bars->p_full[i], 8);
      mbar_init(&bars->p_empty[i], 1);
    }
    mbar_init(&bars->d_full, 1);
    mbar_init(&bars->d_empty, 16);        // 8 softmax warps in each CTA
    fence_barrier_init();
    tma_prefetch_desc(&tmXL);
    tma_prefetch_desc(&tmY32);
    tma_prefetch_desc(&tmY64);
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
    // ===================== TMA producer: lane rows + the MMA1 stream operand (one thread in each CTA) ==========
    if (elect_one()) {
      uint32_t ys = 0, yph = 0, it = 0;
      for (int item = cluster_id; item < num_items; item += num_clusters, ++it) {
        const int split = item % p.nsplit, lh = item / p.nsplit;
        const int lp = lh / p.num_halves;
        const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
        mbar_wait(&bars->xl_empty, (it & 1) ^ 1);
        if (leader) mbar_expect_tx(&bars->xl_full, 2u * (uint32_t)p.nkb * 16384u);
        for (int kb = 0; kb < p.nkb; ++kb)
          tma_load_2d_pair(smem + kb * 16384, &tmXL, &bars->xl_full, kb * 64, (2 * lp + (int)rank) * 128, kEvictLast);
        for (int t = t_begin; t < t_end; ++t) {
          for (int kb0 = 0; kb0 < p.nkb; kb0 += N2_Y_KPS) {
            const int nb = min(N2_Y_KPS, p.nkb - kb0);
            mbar_wait(&bars->y_empty[ys], yph ^ 1);
            if (leader) mbar_expect_tx(&bars->y_full[ys], 2u * (uint32_t)nb * N2_Y_BOX);
            for (int j = 0; j < nb; ++j)
              tma_load_2d_pair(smem + N2_OFF_Y + ys * N2_Y_STAGE + j * N2_Y_BOX, &tmY32, &bars->y_full[ys],
                               (kb0 + j) * 64, t * 64 + (int)rank * 32, kEvictNormal);
            if (++ys == N2_Y_STAGES) { ys = 0; yph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 10) {
    // ===================== TMA producer of the MMA2 operand (one thread in each CTA) =====================
    if (elect_one()) {
      uint32_t ts = 0, tph = 0;
      for (int item = cluster_id; item < num_items; item += num_clusters) {
        const int split = item % p.nsplit, lh = item / p.nsplit;
        const int half = lh % p.num_halves;
        const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
        for (int t = t_begin; t < t_end; ++t) {
          mbar_wait(&bars->yt_empty[ts], tph ^ 1);
          if (leader) mbar_expect_tx(&bars->yt_full[ts], 2u * N2_YT_STAGE);
          for (int j = 0; j < 2; ++j)  // this CTA's 128 of the item's 256 D-columns (columns >= D: zero fill)
            tma_load_2d_pair(smem + N2_OFF_YT + ts * N2_YT_STAGE + j * 8192, &tmY64, &bars->yt_full[ts],
                             half * 256 + (int)rank * 128 + j * 64, t * 64, kEvictNormal);
          if (++ts == N2_YT_STAGES) { ts = 0; tph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA1 issuer (leader CTA only): S = XL . Y^T, as far ahead as the two S buffers allow ======
    // MMA1 and MMA2 are issued by DIFFERENT threads: with one issuer S(i+2) could only be issued after P(i-1) had
    // been waited for, a dependency loop through three mbarrier round trips per tile (measured: the tensor pipe idle
    // and the softmax warps waiting on S 43 % of the time).
    if (leader && elect_one()) {
      uint32_t ys = 0, yph = 0, it = 0;
      uint32_t tile_ctr = 0;  // S tiles issued (S buffer = ctr & 1)
      for (int item = cluster_id; item < num_items; item += num_clusters, ++it) {
        const int split = item % p.nsplit;
        const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
        mbar_wait(&bars->xl_full, it & 1);
        fence_after_sync();
        for (int t = t_begin; t < t_end; ++t, ++tile_ctr) {
          const uint32_t sb = tile_ctr & 1, sph = (tile_ctr >> 1) & 1;
          mbar_wait(&bars->s_empty[sb], sph ^ 1);   // both CTAs' softmax warps hold the buffer's previous tile in registers
          fence_after_sync();
          const uint32_t tmem_s = tmem_base + N2_TMEM_S + sb * 64;
          for (int kb0 = 0; kb0 < p.nkb; kb0 += N2_Y_KPS) {
            const int nb = min(N2_Y_KPS, p.nkb - kb0);
            mbar_wait(&bars->y_full[ys], yph);
            fence_after_sync();
            for (int j = 0; j < nb; ++j) {
              const uint64_t adesc = make_smem_desc_sw128(smem_u32(smem + (kb0 + j) * 16384));
              const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smem + N2_OFF_Y + ys * N2_Y_STAGE + j * N2_Y_BOX));
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                mma_f16_pair(tmem_s, advance_desc_k(adesc, k4 * 32), advance_desc_k(bdesc, k4 * 32), p.idesc1,
                             (kb0 | j | k4) != 0 ? 1u : 0u);
            }
            mma_commit_pair(&bars->y_empty[ys]);
            if (++ys == N2_Y_STAGES) { ys = 0; yph ^= 1; }
          }
          mma_commit_pair(&bars->s_full[sb]);
        }
        mma_commit_pair(&bars->xl_empty);   // every MMA that reads this item's lane rows has been issued by this thread
      }
    }
  } else if (warp == 11) {
    // ===================== MMA2 issuer (leader CTA only): dOut += P . Y =====================
    if (leader && elect_one()) {
      uint32_t ts = 0, tph = 0, it = 0;
      uint32_t p_ctr = 0;     // MMA2 issued (P buffer = ctr & 1)
      for (int item = cluster_id; item < num_items; item += num_clusters, ++it) {
        const int split = item % p.nsplit;
        const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
        mbar_wait(&bars->d_empty, (it & 1) ^ 1);  // both CTAs' epilogues of the previous item drained dOut
        fence_after_sync();
        for (int t = t_begin; t < t_end; ++t, ++p_ctr) {
          const uint32_t pb = p_ctr & 1;
          mbar_wait(&bars->p_full[pb], (p_ctr >> 1) & 1);
          mbar_wait(&bars->yt_full[ts], tph);
          fence_after_sync();
          const uint32_t a_tmem = tmem_base + N2_TMEM_P + pb * 32;
          // B = Y[64 stream rows (K), this CTA's 128 D-columns (N / 2)]: MN-major, 2 boxes of 64 columns 8 KB apart
          const uint64_t bdesc = make_smem_desc_sw128_mn(smem_u32(smem + N2_OFF_YT + ts * N2_YT_STAGE), 8192);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)
            mma_f16_ts_pair(tmem_base, a_tmem + k4 * 8, advance_desc_k(bdesc, k4 * 2048), p.idesc2,
                            (t > t_begin || k4 != 0) ? 1u : 0u);
          mma_commit_pair(&bars->p_empty[pb]);
          mma_commit_pair(&bars->yt_empty[ts]);
          if (++ts == N2_YT_STAGES) { ts = 0; tph ^= 1; }
        }
        mma_commit_pair(&bars->d_full);
      }
    }
  } else if (warp < 10) {
    // ===================== softmax + epilogue warps (2..9) of both CTAs: each CTA owns its 128 lane rows ==========
    const int quad = warp & 3;
    const uint32_t group = (uint32_t)(warp - 2) >> 2;
    const int tid128 = ((warp - 2) & 3) * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const int r_in_tile = quad * 32 + lane;
    const float wl = (MODE != 1) ? __ldg(p.go_lane) * p.inv_2b : 0.f;
    const float scale = __ldg(p.scale_dev);
    const float scale_log2 = scale * 1.4426950408889634f;
    uint32_t tile_ctr = 0, it = 0;
    for (int item = cluster_id; item < num_items; item += num_clusters, ++it) {
      const int split = item % p.nsplit, lh = item / p.nsplit;
      const int half = lh % p.num_halves, lt = 2 * (lh / p.num_halves) + (int)rank;   // this CTA's 128-row lane tile
      const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
      const int row = lt * 128 + r_in_tile;
      const long long jl = p.label_by_lane ? p.label_offset + row : (long long)row - p.label_offset;
      const bool row_ok = row < p.RL;
      const float lse_lane = (MODE != 1 && row_ok) ? __ldg(p.lse_lane + row) * 1.4426950408889634f : 0.f;
      const float lane_off = __log2f(wl) - lse_lane;
      const long long label_lo = p.label_by_lane ? p.label_offset + (long long)lt * 128 : (long long)lt * 128 - p.label_offset;
      float ds_acc = 0.f;
      float ws_n = 0.f, l2_n = 0.f;   // loader threads: next tile's column constants, fetched one tile ahead
      bool have_n = false;
      for (int t = t_begin; t < t_end; ++t, ++tile_ctr) {
        const uint32_t sb = tile_ctr & 1, sph = (tile_ctr >> 1) & 1;
        if (sb != group) continue;  // the other softmax group converts this tile
        const int j0 = t * 64;
        NceColBuf* cols = &bars->cols[group * 2 + sph];
        if (tid128 < 64) {
          if (!have_n) nce_fetch_cols<MODE>(p, j0, tid128, ws_n, l2_n);
          nce_publish_cols<MODE>(cols, tid128, lane, wl, ws_n, l2_n);
          have_n = t + 2 < t_end;
          if (have_n) nce_fetch_cols<MODE>(p, j0 + 128, tid128, ws_n, l2_n);   // this group's next tile
        }
        named_bar_sync(1 + (int)group, 128);
        const bool general = (j0 < label_lo + 128 && j0 + 64 > label_lo) || (cols->neg[0] | cols->neg[1]) != 0u;
        const int jd = (jl >= j0 && jl < (long long)j0 + 64) ? (int)(jl - j0) : -1;
        mbar_wait(&bars->s_full[sb], sph);
        fence_after_sync();
        uint32_t v0[32], v1[32];
        tmem_ld32(tmem_base + lane_addr + N2_TMEM_S + sb * 64, v0);
        tmem_ld32(tmem_base + lane_addr + N2_TMEM_S + sb * 64 + 32, v1);
        tmem_ld_wait();
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&bars->s_empty[sb], 0);
        uint32_t packed[32];
        nce_convert<MODE>(v0, v1, packed, general, jd, scale_log2, lane_off, lse_lane, wl, cols, p, j0, ds_acc);
        // the group's previous tile (two tiles back) must have been consumed by MMA2 before its P buffer is rewritten
        mbar_wait(&bars->p_empty[group], ((tile_ctr >> 1) & 1) ^ 1);
        fence_after_sync();
        tmem_st32(tmem_base + lane_addr + N2_TMEM_P + group * 32, packed);
        tmem_st_wait();
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&bars->p_full[group], 0);
      }
      // ---- epilogue: dOut (own TMEM lanes, columns 0..255) -> global ----
      mbar_wait(&bars->d_full, it & 1);
      fence_after_sync();
      const int dcols = min(256, p.D - half * 256);
      for (int c0 = (int)group * 128; c0 < min(dcols, (int)group * 128 + 128); c0 += 32) {  // each group stores 128 columns
        uint32_t v[32];
        tmem_ld32(tmem_base + lane_addr + c0, v);
        tmem_ld_wait();
        if (row_ok) {
          float* o = p.out + (int64_t)row * p.D + half * 256 + c0;
          if (p.use_atomic) {
            nce_red_row32(o, v, scale, c0, dcols, (p.D & 3) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (c0 + j + 3 < dcols) {
                *reinterpret_cast<float4*>(o + j) =
                    make_float4(__uint_as_float(v[j]) * scale, __uint_as_float(v[j + 1]) * scale,
                                __uint_as_float(v[j + 2]) * scale, __uint_as_float(v[j + 3]) * scale);
              } else {
                for (int e = 0; e < 4; ++e)
                  if (c0 + j + e < dcols) o[j + e] = __uint_as_float(v[j + e]) * scale;
              }
            }
          }
        }
      }
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&bars->d_empty, 0);
      if (MODE != 1 && p.dscale != nullptr && half == 0) {
        float s = row_ok ? ds_acc : 0.f;
        s = warp_sum(s);
        if (lane == 0) atomicAdd(p.dscale, s);
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
bool nce_bwd_pair_enabled() {
  return tune_get(ILFDT_TUNE_NCE_2CTA) != 0;
}

template <int MODE>
static int32_t launch_n2_t(const CUtensorMap& tmXL, const CUtensorMap& tmY32, const CUtensorMap& tmY64, NceBwdParams& p,
                           int base_items, float* out, int64_t out_bytes, cudaStream_t st) {
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
  cfg.blockDim = dim3(N2_THREADS);
  cfg.dynamicSmemBytes = N2_SMEM_BYTES;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (configured_dev != dev) {
    ILFDT_CUDA_OK(cudaFuncSetAttribute(nce_bwd2_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, N2_SMEM_BYTES));
    cfg.gridDim = dim3(2 * (sm_count() / 2));
    int n = 0;
    ILFDT_CUDA_OK(cudaOccupancyMaxActiveClusters(&n, nce_bwd2_kernel<MODE>, &cfg));
    ILFDT_REQUIRE(n >= 1, ILFDT_E_CUDA, "nce_bwd2: no CTA pair can be resident");
    max_clusters = std::min(n, sm_count() / 2);
    configured_dev = dev;
  }
  nce_bwd_plan(p, base_items, max_clusters);
  if (p.use_atomic) ILFDT_CUDA_OK(cudaMemsetAsync(out, 0, (size_t)out_bytes, st));
  const int items = base_items * p.nsplit;
  cfg.gridDim = dim3(2 * std::min(items, max_clusters));
  ILFDT_CUDA_OK(cudaLaunchKernelEx(&cfg, nce_bwd2_kernel<MODE>, tmXL, tmY32, tmY64, p));
  return ILFDT_OK;
}

int32_t nce_bwd_pair_launch(const NceBwdLaunch& a, cudaStream_t st) {
  NceBwdParams p{};
  const int64_t RL = a.RL, RS = a.RS, D = a.D;
  p.RL = (int)RL; p.RS = (int)RS; p.D = (int)D;
  p.nkb = (int)cdiv(D, 64);
  p.num_lane_tiles = (int)cdiv(RL, 256);          // pairs of 128-row lane tiles
  p.num_halves = (int)cdiv(D, 256);
  p.num_stream_tiles = (int)cdiv(RS, 64);
  p.label_by_lane = a.label_by_lane;
  p.label_offset = a.label_offset;
  p.lse_lane = a.lse_lane;
  p.lse_stream = a.lse_stream;
  p.go_lane = a.go_lane;
  p.go_stream = a.go_stream;
  p.rows_per_go = std::max<int64_t>(1, a.rows_per_go);
  p.inv_2b = a.inv_2b;
  p.scale_dev = a.scale;
  p.out = a.out;
  p.dscale = a.dscale;
  p.idesc1 = make_idesc_bf16(256, 64);
  p.idesc2 = make_idesc_bf16(256, 256, 0, 1);   // B = Y boxes read MN-major
  CUtensorMap tmXL, tmY32, tmY64;
  int32_t rc = make_tmap_bf16(&tmXL, a.XL, RL, D, D, 128);
  if (rc != ILFDT_OK) return rc;
  if ((rc = make_tmap_bf16(&tmY32, a.Y, RS, D, D, 32)) != ILFDT_OK) return rc;
  if ((rc = make_tmap_bf16(&tmY64, a.Y, RS, D, D, 64)) != ILFDT_OK) return rc;
  const int base_items = p.num_lane_tiles * p.num_halves;
  const int64_t out_bytes = RL * D * 4;
  if (a.mode == 0) return launch_n2_t<0>(tmXL, tmY32, tmY64, p, base_items, a.out, out_bytes, st);
  if (a.mode == 1) return launch_n2_t<1>(tmXL, tmY32, tmY64, p, base_items, a.out, out_bytes, st);
  return launch_n2_t<2>(tmXL, tmY32, tmY64, p, base_items, a.out, out_bytes, st);
}

}  // namespace ilfdt
