// a13 (loss part): fused InfoNCE backward on tcgen05 -- the [b, B] logits / probabilities are
// recomputed tile by tile in TMEM and consumed on-chip; they never reach HBM.
//
// One kernel. Per launch:
//   "lane" matrix  XL [RL, D]  (its 128-row tiles are the TMEM lanes; the gradient is w.r.t. XL)
//   "stream" matrix Y  [RS, D]  (rows streamed 64 at a time)
//   S[i, j]   = XL[i,:] . Y[j,:]                      MMA1  128 x 64 x D      -> TMEM (double buffered)
//   P[i, j]   = wl   * (exp(scale*S - lse_lane[i])   - [i ~ j])     (MODE 0, 2: "lane" softmax)
//             + ws_j * (exp(scale*S - lse_stream[j]) - [i ~ j])     (MODE 1, 2: "stream" softmax)
//                                                       softmax warps: TMEM -> regs -> bf16 -> smem (SW128)
//   dOut[i,:] += P[i, j-block] . Y[j-block, half]       MMA2  128 x 256 x 64    -> TMEM (256 columns)
//   dOut      *= scale                                  epilogue
// MMA2 reads the SAME row-major Y boxes TMA delivers for MMA1 (64 rows x 128 B, 128-byte swizzle) as an
// MN-major B operand (N = D contiguous, K = stream row), so no transposed copy of any operand exists.
// MODE 2 is the gathered form used in training: for local image rows i the total gradient is
//   dX_img[i] = scale * sum_j (P_img[i,j] + P_txt[j,i]) G_txt[j]   with P_txt[j,i] built from lse_txt[j] of ALL ranks,
// i.e. one launch per modality produces the local rows' full gradient (what dX + reduce-scatter(dG) gives in
// the reference, clip_fdt.py:182-188) and nothing has to be reduce-scattered. MODE 0 / 1 are the separate
// dX / dG launches for callers whose gathered operand is an independent tensor.
// A CTA owns (lane tile, 256-wide half of D, slice of the stream range); S is recomputed for each half
// (TMEM holds 512 columns: 256 for dOut + 2 x 64 for S).  The MMA thread software-pipelines MMA1(t+1)
// ahead of MMA2(t) so the tensor pipe stays busy while the softmax warps convert tile t.
#include <algorithm>
#include "nce_bwd_common.cuh"

namespace ilfdt {

using namespace umma;

constexpr int NB_THREADS = 352;  // Y TMA warp, MMA warp, 2 softmax groups of 4 warps (group g owns S buffer g), YT TMA warp
constexpr int NB_XL_BYTES = 8 * 16384;          // lane tile, all k-blocks: 128 KB
constexpr int NB_Y_STAGE = 64 * 64 * 2;         // 64 stream rows x 64 k      : 8 KB
constexpr int NB_Y_STAGES = 4;          // MMA1 consumes 8 of these 8 KB boxes per S tile (2 stages left it TMA-latency-bound)
constexpr int NB_YT_STAGE = 256 * 64 * 2;       // 256 D-columns x 64 stream   : 32 KB
constexpr int NB_YT_STAGES = 2;
constexpr int NB_OFF_Y = NB_XL_BYTES;
constexpr int NB_OFF_YT = NB_OFF_Y + NB_Y_STAGES * NB_Y_STAGE;
constexpr int NB_OFF_BARS = NB_OFF_YT + NB_YT_STAGES * NB_YT_STAGE;  // 229376
constexpr int NB_SMEM_BYTES = NB_OFF_BARS + 1024 + 2048;  // 232448 = the sm_100 opt-in maximum
constexpr int NB_TMEM_COLS = 512;
constexpr int NB_TMEM_S = 256;  // S buffers at columns 256 and 320
constexpr int NB_TMEM_P = 384;  // P (bf16 pairs, the A operand of MMA2) at columns 384 and 416, one buffer per softmax group

struct NceBwdBars {
  uint64_t xl_full, xl_empty;
  uint64_t y_full[NB_Y_STAGES], y_empty[NB_Y_STAGES];
  uint64_t yt_full[NB_YT_STAGES], yt_empty[NB_YT_STAGES];
  uint64_t s_full[2], s_empty[2];
  uint64_t p_full[2], p_empty[2];
  uint64_t d_full, d_empty;
  uint32_t tmem_ptr;
  NceColBuf cols[2];   // one per softmax group
};

static_assert(sizeof(NceBwdBars) <= 2048, "barrier block outgrew its shared-memory slot");
static_assert(NB_SMEM_BYTES <= 232448, "over the sm_100 dynamic shared memory limit");

template <int MODE>
__global__ void __launch_bounds__(NB_THREADS, 1)
nce_bwd_kernel(const __grid_constant__ CUtensorMap tmXL, const __grid_constant__ CUtensorMap tmY,
               const NceBwdParams p) {
  extern __shared__ uint8_t nb_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(nb_smem_raw) + 1023) & ~(uintptr_t)1023);
  NceBwdBars* bars = reinterpret_cast<NceBwdBars*>(smem + NB_OFF_BARS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_items = p.num_lane_tiles * p.num_halves * p.nsplit;

  if (threadIdx.x == 0) {
    mbar_init(&bars->xl_full, 1);
    mbar_init(&bars->xl_empty, 1);
    for (int i = 0; i < NB_Y_STAGES; ++i) { mbar_init(&bars->y_full[i], 1); mbar_init(&bars->y_empty[i], 1); }
    for (int i = 0; i < NB_YT_STAGES; ++i) { mbar_init(&bars->yt_full[i], 1); mbar_init(&bars->yt_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->s_full[i], 1); mbar_init(&bars->s_empty[i], 4); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->p_full[i], 4); mbar_init(&bars->p_empty[i], 1); }
    mbar_init(&bars->d_full, 1);
    mbar_init(&bars->d_empty, 8);
    fence_barrier_init();
    tma_prefetch_desc(&tmXL);
    tma_prefetch_desc(&tmY);
  }
  if (warp == 1) {
    tmem_alloc(&bars->tmem_ptr, NB_TMEM_COLS);
    tmem_relinquish();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = bars->tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      uint32_t ys = 0, yph = 0, it = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
        const int split = item % p.nsplit, lh = item / p.nsplit;
        const int lt = lh / p.num_halves;
        const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
        mbar_wait(&bars->xl_empty, (it & 1) ^ 1);
        mbar_expect_tx(&bars->xl_full, (uint32_t)p.nkb * 16384u);
        for (int kb = 0; kb < p.nkb; ++kb)
          tma_load_2d(smem + kb * 16384, &tmXL, &bars->xl_full, kb * 64, lt * 128, kEvictLast);
        for (int t = t_begin; t < t_end; ++t) {
          for (int kb = 0; kb < p.nkb; ++kb) {
            mbar_wait(&bars->y_empty[ys], yph ^ 1);
            mbar_expect_tx(&bars->y_full[ys], NB_Y_STAGE);
            tma_load_2d(smem + NB_OFF_Y + ys * NB_Y_STAGE, &tmY, &bars->y_full[ys], kb * 64, t * 64, kEvictNormal);
            if (++ys == NB_Y_STAGES) { ys = 0; yph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 10) {
    // ===================== TMA producer of the MMA2 operand (own warp: the Y ring must not queue behind it) =====
    if (elect_one()) {
      uint32_t ts = 0, tph = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int split = item % p.nsplit, lh = item / p.nsplit;
        const int half = lh % p.num_halves;
        const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
        for (int t = t_begin; t < t_end; ++t) {
          mbar_wait(&bars->yt_empty[ts], tph ^ 1);
          mbar_expect_tx(&bars->yt_full[ts], NB_YT_STAGE);
          for (int j = 0; j < 4; ++j)  // the CTA's 256 D-columns of the same 64 stream rows (columns >= D: zero fill)
            tma_load_2d(smem + NB_OFF_YT + ts * NB_YT_STAGE + j * 8192, &tmY, &bars->yt_full[ts], half * 256 + j * 64,
                        t * 64, kEvictNormal);
          if (++ts == NB_YT_STAGES) { ts = 0; tph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      uint32_t ys = 0, yph = 0, ts = 0, tph = 0, it = 0;
      uint32_t tile_ctr = 0;  // global count of S tiles issued by this CTA (S buffer = ctr & 1)
      uint32_t p_ctr = 0;     // global count of MMA2 issued (P buffer phases)
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
        const int split = item % p.nsplit;
        const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
        const int ntile = t_end - t_begin;
        mbar_wait(&bars->xl_full, it & 1);
        mbar_wait(&bars->d_empty, (it & 1) ^ 1);  // epilogue of the previous item drained dOut
        fence_after_sync();
        int n1 = 0;  // S tiles issued in this item
        for (int i = 0; i < ntile; ++i) {
          // ---- MMA1 runs up to two tiles ahead of MMA2 (S is double buffered and released as soon as the softmax
          //      warps hold it in registers), so the tensor pipe computes S(i+1), S(i+2) while P(i) is being formed ----
          while (n1 < ntile && n1 <= i + 2) {
            const uint32_t sb = tile_ctr & 1, sph = (tile_ctr >> 1) & 1;
            mbar_wait(&bars->s_empty[sb], sph ^ 1);
            fence_after_sync();
            const uint32_t tmem_s = tmem_base + NB_TMEM_S + sb * 64;
            for (int kb = 0; kb < p.nkb; ++kb) {
              mbar_wait(&bars->y_full[ys], yph);
              fence_after_sync();
              const uint64_t adesc = make_smem_desc_sw128(smem_u32(smem + kb * 16384));
              const uint64_t bdesc = make_smem_desc_sw128(smem_u32(smem + NB_OFF_Y + ys * NB_Y_STAGE));
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                mma_f16(tmem_s, advance_desc_k(adesc, k4 * 32), advance_desc_k(bdesc, k4 * 32), p.idesc1,
                        (kb | k4) != 0 ? 1u : 0u);
              mma_commit(&bars->y_empty[ys]);
              if (++ys == NB_Y_STAGES) { ys = 0; yph ^= 1; }
            }
            mma_commit(&bars->s_full[sb]);
            ++tile_ctr;
            ++n1;
          }
          {
            // ---- MMA2: dOut += P(tile i) . YT(tile i)^T ----
            const uint32_t pb = p_ctr & 1;   // P buffer = softmax group = S buffer of the tile
            mbar_wait(&bars->p_full[pb], (p_ctr >> 1) & 1);
            mbar_wait(&bars->yt_full[ts], tph);
            fence_after_sync();
            const uint32_t a_tmem = tmem_base + NB_TMEM_P + pb * 32;   // 16 K-elements = 8 columns per MMA
            // B = Y[64 stream rows (K), 256 D-columns (N)], N contiguous: MN-major, 4 boxes of 64 columns 8 KB
            // apart (leading offset), 8-row groups 1 KB apart (stride offset); a K step of 16 rows = 2 KB
            const uint64_t bdesc = make_smem_desc_sw128_mn(smem_u32(smem + NB_OFF_YT + ts * NB_YT_STAGE), 8192);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              mma_f16_ts(tmem_base, a_tmem + k4 * 8, advance_desc_k(bdesc, k4 * 2048), p.idesc2,
                         (i > 0 || k4 != 0) ? 1u : 0u);
            mma_commit(&bars->p_empty[pb]);
            mma_commit(&bars->yt_empty[ts]);
            if (++ts == NB_YT_STAGES) { ts = 0; tph ^= 1; }
            ++p_ctr;
          }
        }
        mma_commit(&bars->d_full);
        mma_commit(&bars->xl_empty);
      }
    }
  } else if (warp < 10) {
    // ===================== softmax + epilogue warps (2..9) =====================
    const int quad = warp & 3;
    const uint32_t group = (uint32_t)(warp - 2) >> 2;
    const int tid128 = ((warp - 2) & 3) * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const int r_in_tile = quad * 32 + lane;
    const float wl = (MODE != 1) ? __ldg(p.go_lane) * p.inv_2b : 0.f;
    const float scale = __ldg(p.scale_dev);
    const float scale_log2 = scale * 1.4426950408889634f;
    uint32_t tile_ctr = 0, it = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
      const int split = item % p.nsplit, lh = item / p.nsplit;
      const int half = lh % p.num_halves, lt = lh / p.num_halves;
      const int t_begin = split * p.tiles_per_split, t_end = min(p.num_stream_tiles, t_begin + p.tiles_per_split);
      const int row = lt * 128 + r_in_tile;
      // label column (stream index) matching this lane row, or an index that never occurs
      const long long jl = p.label_by_lane ? p.label_offset + row          // stream j == label_offset + lane row
                                           : (long long)row - p.label_offset;  // lane row == label_offset + stream j
      const bool row_ok = row < p.RL;
      const float lse_lane = (MODE != 1 && row_ok) ? __ldg(p.lse_lane + row) * 1.4426950408889634f : 0.f;
      // fast path: lane weight folded into the exponent; rows beyond RL give finite garbage that is never stored
      const float lane_off = __log2f(wl) - lse_lane;
      // stream columns that hold the labels of this lane tile: [label_lo, label_lo + 128)
      const long long label_lo = p.label_by_lane ? p.label_offset + (long long)lt * 128 : (long long)lt * 128 - p.label_offset;
      float ds_acc = 0.f;
      for (int t = t_begin; t < t_end; ++t, ++tile_ctr) {
        const uint32_t sb = tile_ctr & 1, sph = (tile_ctr >> 1) & 1;
        if (sb != group) continue;  // the other softmax group converts this tile
        const int j0 = t * 64;
        NceColBuf* cols = &bars->cols[sb];
        if (tid128 < 64) {
          float ws, l2;
          nce_fetch_cols<MODE>(p, j0, tid128, ws, l2);
          nce_publish_cols<MODE>(cols, tid128, lane, wl, ws, l2);
        }
        named_bar_sync(1 + (int)group, 128);
        // the general path is needed only by tiles that hold a label column of this lane tile or a negative weight
        const bool general = (j0 < label_lo + 128 && j0 + 64 > label_lo) || (cols->neg[0] | cols->neg[1]) != 0u;
        const int jd = (jl >= j0 && jl < (long long)j0 + 64) ? (int)(jl - j0) : -1;
        mbar_wait(&bars->s_full[sb], sph);
        fence_after_sync();
        uint32_t v0[32], v1[32];
        tmem_ld32(tmem_base + lane_addr + NB_TMEM_S + sb * 64, v0);
        tmem_ld32(tmem_base + lane_addr + NB_TMEM_S + sb * 64 + 32, v1);
        tmem_ld_wait();
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->s_empty[sb]);
        uint32_t packed[32];
        nce_convert<MODE>(v0, v1, packed, general, jd, scale_log2, lane_off, lse_lane, wl, cols, p, j0, ds_acc);
        // one P buffer per group: the group's previous tile (two tiles back) must have been consumed by MMA2.
        // (A single shared buffer would alias mbarrier parities now that S runs two tiles ahead of MMA2.)
        mbar_wait(&bars->p_empty[group], ((tile_ctr >> 1) & 1) ^ 1);
        // P stays on chip as the TMEM A operand of MMA2: thread = lane row, 32 columns of bf16 pairs (K = 64)
        tmem_st32(tmem_base + lane_addr + NB_TMEM_P + group * 32, packed);
        tmem_st_wait();
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->p_full[group]);
        named_bar_sync(1 + (int)group, 128);  // every reader of col_* is done before the next tile's values land
      }
      // ---- epilogue: dOut (TMEM columns 0..255) -> global ----
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
      if (lane == 0) mbar_arrive(&bars->d_empty);
      if (MODE != 1 && p.dscale != nullptr && half == 0) {
        float s = row_ok ? ds_acc : 0.f;
        s = warp_sum(s);
        if (lane == 0) atomicAdd(p.dscale, s);
      }
    }
  }

  fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc(tmem_base, NB_TMEM_COLS);
  }
}

int64_t nce_bwd_bf16_ws_bytes(int64_t b, int64_t B, int64_t D) {
  (void)b; (void)B; (void)D;
  return 256;
}

static int32_t launch_one(const NceBwdLaunch& a, cudaStream_t st) {
  if (nce_bwd_pair_enabled() && a.RL >= 256) return nce_bwd_pair_launch(a, st);
  NceBwdParams p{};
  const int64_t RL = a.RL, RS = a.RS, D = a.D;
  p.RL = (int)RL; p.RS = (int)RS; p.D = (int)D;
  p.nkb = (int)cdiv(D, 64);
  p.num_lane_tiles = (int)cdiv(RL, 128);
  p.num_halves = (int)cdiv(D, 256);
  p.num_stream_tiles = (int)cdiv(RS, 64);
  const int base_items = p.num_lane_tiles * p.num_halves;
  const int sms = sm_count();
  nce_bwd_plan(p, base_items, sms);
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
  p.idesc1 = make_idesc_bf16(128, 64);
  p.idesc2 = make_idesc_bf16(128, 256, 0, 1);   // B = Y box read MN-major
  if (p.use_atomic) ILFDT_CUDA_OK(cudaMemsetAsync(a.out, 0, (size_t)RL * D * 4, st));
  CUtensorMap tmXL, tmY;
  int32_t rc = make_tmap_bf16(&tmXL, a.XL, RL, D, D, 128);
  if (rc != ILFDT_OK) return rc;
  rc = make_tmap_bf16(&tmY, a.Y, RS, D, D, 64);
  if (rc != ILFDT_OK) return rc;
  static thread_local int configured_dev = -1;
  int dev = 0;
  ILFDT_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    ILFDT_CUDA_OK(cudaFuncSetAttribute(nce_bwd_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, NB_SMEM_BYTES));
    ILFDT_CUDA_OK(cudaFuncSetAttribute(nce_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, NB_SMEM_BYTES));
    ILFDT_CUDA_OK(cudaFuncSetAttribute(nce_bwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, NB_SMEM_BYTES));
    configured_dev = dev;
  }
  const int items = base_items * p.nsplit;
  const int grid = std::min(items, sms);
  if (a.mode == 0) nce_bwd_kernel<0><<<grid, NB_THREADS, NB_SMEM_BYTES, st>>>(tmXL, tmY, p);
  else if (a.mode == 1) nce_bwd_kernel<1><<<grid, NB_THREADS, NB_SMEM_BYTES, st>>>(tmXL, tmY, p);
  else nce_bwd_kernel<2><<<grid, NB_THREADS, NB_SMEM_BYTES, st>>>(tmXL, tmY, p);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

int32_t launch_nce_bwd_bf16(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt,
                            int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset, const float* lse_i,
                            const float* lse_t, const float* grad_out, float* dx_img, float* dx_txt, float* dg_img,
                            float* dg_txt, float* dscale, void* ws, int64_t ws_bytes, cudaStream_t st) {
  (void)ws; (void)ws_bytes;
  ILFDT_REQUIRE(D <= 512 && D % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE, "nce_bwd(bf16): D=%lld must be <= 512, %% 8",
                (long long)D);
  ILFDT_CUDA_OK(cudaMemsetAsync(dscale, 0, 4, st));
  NceBwdLaunch a{};
  a.D = D; a.label_offset = label_offset; a.inv_2b = 1.f / (2.f * (float)b); a.scale = scale;
  a.go_lane = grad_out; a.go_stream = grad_out; a.rows_per_go = std::max<int64_t>(B, 1);
  int32_t rc;
  // logits_per_image = x_img . g_txt^T : dX_img (lse_i by lane), dG_txt (lse_i by stream)
  a.XL = x_img; a.Y = g_txt; a.RL = b; a.RS = B; a.mode = 0; a.label_by_lane = 1;
  a.lse_lane = lse_i; a.lse_stream = nullptr; a.out = dx_img; a.dscale = dscale;
  if ((rc = launch_one(a, st)) != ILFDT_OK) return rc;
  a.XL = g_txt; a.Y = x_img; a.RL = B; a.RS = b; a.mode = 1; a.label_by_lane = 0;
  a.lse_lane = nullptr; a.lse_stream = lse_i; a.out = dg_txt; a.dscale = nullptr;
  if ((rc = launch_one(a, st)) != ILFDT_OK) return rc;
  // logits_per_text = x_txt . g_img^T : dX_txt (lse_t by lane), dG_img (lse_t by stream)
  a.XL = x_txt; a.Y = g_img; a.RL = b; a.RS = B; a.mode = 0; a.label_by_lane = 1;
  a.lse_lane = lse_t; a.lse_stream = nullptr; a.out = dx_txt; a.dscale = dscale;
  if ((rc = launch_one(a, st)) != ILFDT_OK) return rc;
  a.XL = g_img; a.Y = x_txt; a.RL = B; a.RS = b; a.mode = 1; a.label_by_lane = 0;
  a.lse_lane = nullptr; a.lse_stream = lse_t; a.out = dg_img; a.dscale = nullptr;
  return launch_one(a, st);
}

// Gathered form: g_* are the all-gather of every rank's x_* (rank-major), lse_*_all [B] the all-gather of the
// forward's per-row log-sum-exp, go_all [B / b] every rank's upstream dLoss. Two launches produce the local rows'
// TOTAL gradient; there is no gradient w.r.t. the gathered copies left to reduce-scatter.
int32_t launch_nce_bwd_gathered_bf16(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt,
                                     int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset,
                                     const float* lse_i_all, const float* lse_t_all, const float* go_all,
                                     float* dx_img, float* dx_txt, float* dscale, cudaStream_t st) {
  ILFDT_REQUIRE(D <= 512 && D % 8 == 0, ILFDT_E_UNSUPPORTED_SHAPE, "nce_bwd(bf16): D=%lld must be <= 512, %% 8",
                (long long)D);
  ILFDT_CUDA_OK(cudaMemsetAsync(dscale, 0, 4, st));
  NceBwdLaunch a{};
  a.D = D; a.label_offset = label_offset; a.inv_2b = 1.f / (2.f * (float)b); a.scale = scale;
  a.RL = b; a.RS = B; a.mode = 2; a.label_by_lane = 1;
  a.go_lane = go_all + label_offset / b; a.go_stream = go_all; a.rows_per_go = b;
  a.dscale = dscale;
  int32_t rc;
  a.XL = x_img; a.Y = g_txt; a.lse_lane = lse_i_all + label_offset; a.lse_stream = lse_t_all; a.out = dx_img;
  if ((rc = launch_one(a, st)) != ILFDT_OK) return rc;
  a.XL = x_txt; a.Y = g_img; a.lse_lane = lse_t_all + label_offset; a.lse_stream = lse_i_all; a.out = dx_txt;
  return launch_one(a, st);
}

}  // namespace ilfdt
