// a3 on a CTA pair: out[n, m] (+)= alpha * sum_k B[n,k] A[m,k] for a SHORT A (M <= 512 rows) and a long reduction --
// the pooling f[b, D] = w[b, K] . sd[K, D] (clip_fdt.py:154; A = sd^T [D, K], B = w [b, K], Kr = K = 16384).
//
// Why: nothing of this product can stay resident (the reduction is 16384 long), so the single-CTA kernel streams
// 48 KB per 128x256x64 block of MACs -- 24 TB/s over the whole chip at the tensor rate, twice what L2 delivers; it ran
// at a third of the tensor peak. Here the lanes are the rows of B (256 per CTA pair, 128 per CTA) and ALL of A's rows
// are the columns of up to two 256-column accumulators, so one 16 KB box of B rows per CTA is multiplied against the
// whole of A, and the pair's tcgen05.mma.cta_group::2 lets each CTA stage only half of A's rows:
// 48 KB per CTA per 128x512x64 block of MACs, half the bytes per FLOP.
//
// Protocol as in ground_umma2.cu (leader = CTA rank 0): producers of both CTAs load into their own smem with the
// transaction bytes on the leader's full[stage]; the leader issues the MMAs and multicasts commits; the epilogue warps
// of each CTA drain their own TMEM lanes (group g = accumulator g) and arrive on the leader's tmem_empty.
#include <algorithm>
#include <cstdlib>
#include "umma_pair.cuh"
#include "launch.h"

namespace ilfdt {

using namespace umma;

constexpr int W2_THREADS = 320;
constexpr int W2_BOX = 128 * 64 * 2;            // 16 KB: 128 rows x 64 k
constexpr int W2_STAGE = 3 * W2_BOX;            // this CTA's B rows + its half of A's rows for each accumulator: 48 KB
constexpr int W2_STAGES = 4;
constexpr int W2_DATA_BYTES = W2_STAGES * W2_STAGE;   // 192 KB
constexpr int W2_SMEM_BYTES = W2_DATA_BYTES + 1024 + 256;

struct W2Params {
  int n_rows, m_rows, nkb;      // lanes = rows of B (n), columns = rows of A (m)
  int num_pairs;                // ceil(n_rows / 256)
  int nacc;                     // accumulators in use: ceil(m_rows / 256)
  int acc_n[2];                 // UMMA N of each accumulator (multiple of 16)
  float alpha;
  int accumulate;
  float* out;
  long long ldo;
  uint32_t idesc[2];
  // split reduction (few B rows, e.g. a per-GPU batch of 256: one CTA pair would do the whole 16384-long reduction
  // alone): item = pair * ksplit + ks contracts k-blocks [ks * kb_per, (ks + 1) * kb_per) into slab ks (dense
  // [n_rows, m_rows] fp32); w2_reduce_kernel adds the slabs in a fixed order (deterministic, no atomics)
  int ksplit, kb_per;
  float* slab;
  long long slab_stride;
  // 1: A is given as [Kr, M] row-major (the reduction index outermost: the codebook sd [K, D] itself, no transposed copy);
  // each CTA stages its 128 columns of a 64-k block as two 64 x 64 boxes read as an MN-major operand
  int col_mn;
};

struct W2Bars {
  uint64_t full[W2_STAGES], empty[W2_STAGES], tmem_full, tmem_empty;
  uint32_t tmem_ptr;
};

__global__ void __launch_bounds__(W2_THREADS, 1)
gemm_wide2_kernel(const __grid_constant__ CUtensorMap tmLane, const __grid_constant__ CUtensorMap tmCol, const W2Params p) {
  extern __shared__ uint8_t w2_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(w2_smem_raw) + 1023) & ~(uintptr_t)1023);
  W2Bars* bars = reinterpret_cast<W2Bars*>(smem + W2_DATA_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < W2_STAGES; ++i) { mbar_init(&bars->full[i], 1); mbar_init(&bars->empty[i], 1); }
    mbar_init(&bars->tmem_full, 1);
    mbar_init(&bars->tmem_empty, 16);   // 8 epilogue warps in each CTA
    fence_barrier_init();
    tma_prefetch_desc(&tmLane);
    tma_prefetch_desc(&tmCol);
  }
  if (warp == 1) {
    tmem_alloc_pair(&bars->tmem_ptr, 512);
    tmem_relinquish_pair();
  }
  fence_before_sync();
  cluster_sync_all();
  fence_after_sync();
  const uint32_t tmem_base = bars->tmem_ptr;
  const uint32_t stage_tx = 2u * (uint32_t)(1 + p.nacc) * W2_BOX;   // both CTAs: lane box + one A box per accumulator

  if (warp == 0) {
    // ===================== TMA producer (one thread in each CTA) =====================
    if (elect_one()) {
      uint32_t stage = 0, phase = 0;
      for (int item = cluster_id; item < p.num_pairs * p.ksplit; item += num_clusters) {
        const int pair = item / p.ksplit, ks = item - pair * p.ksplit;
        const int lane_row0 = (2 * pair + (int)rank) * 128;
        const int kb_end = min(p.nkb, (ks + 1) * p.kb_per);
        for (int kb = ks * p.kb_per; kb < kb_end; ++kb) {
          mbar_wait(&bars->empty[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&bars->full[stage], stage_tx);
          uint8_t* st = smem + stage * W2_STAGE;
          tma_load_2d_pair(st, &tmLane, &bars->full[stage], kb * 64, lane_row0, kEvictFirst);
          for (int a = 0; a < p.nacc; ++a) {  // this CTA's half of accumulator a's columns (rows of A)
            if (p.col_mn) {
              const int n0 = a * 256 + (int)rank * 128;
              tma_load_2d_pair(st + (1 + a) * W2_BOX, &tmCol, &bars->full[stage], n0, kb * 64, kEvictLast);
              tma_load_2d_pair(st + (1 + a) * W2_BOX + 8192, &tmCol, &bars->full[stage], n0 + 64, kb * 64, kEvictLast);
            } else {
              tma_load_2d_pair(st + (1 + a) * W2_BOX, &tmCol, &bars->full[stage], kb * 64,
                               a * 256 + (int)rank * (p.acc_n[a] >> 1), kEvictLast);
            }
          }
          if (++stage == W2_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && elect_one()) {
      uint32_t stage = 0, phase = 0, it = 0;
      for (int item = cluster_id; item < p.num_pairs * p.ksplit; item += num_clusters, ++it) {
        const int ks = item % p.ksplit;
        const int kb0 = ks * p.kb_per, kb_end = min(p.nkb, (ks + 1) * p.kb_per);
        mbar_wait(&bars->tmem_empty, (it & 1) ^ 1);   // both CTAs' epilogues of the previous item are done
        fence_after_sync();
        for (int kb = kb0; kb < kb_end; ++kb) {
          mbar_wait(&bars->full[stage], phase);
          fence_after_sync();
          const uint8_t* st = smem + stage * W2_STAGE;
          const uint64_t adesc = make_smem_desc_sw128(smem_u32(st));
          for (int a = 0; a < p.nacc; ++a) {
            // K-major: 16 k = 32 bytes inside the swizzle atom; MN-major: 16 k-rows of 128 bytes, the second 64 columns 8 KB on
            const uint64_t bdesc = p.col_mn ? make_smem_desc_sw128_mn(smem_u32(st + (1 + a) * W2_BOX), 8192)
                                            : make_smem_desc_sw128(smem_u32(st + (1 + a) * W2_BOX));
            const uint32_t bstep = p.col_mn ? 2048u : 32u;
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              mma_f16_pair(tmem_base + a * 256, advance_desc_k(adesc, k4 * 32), advance_desc_k(bdesc, k4 * bstep),
                           p.idesc[a], ((kb - kb0) | k4) != 0 ? 1u : 0u);
          }
          mma_commit_pair(&bars->empty[stage]);
          if (++stage == W2_STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit_pair(&bars->tmem_full);
      }
    }
  } else {
    // ===================== epilogue warps (2..9) of both CTAs: group g stores accumulator g =====================
    const int quad = warp & 3;
    const int group = (warp - 2) >> 2;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    uint32_t it = 0;
    for (int item = cluster_id; item < p.num_pairs * p.ksplit; item += num_clusters, ++it) {
      const int pair = item / p.ksplit, ks = item - pair * p.ksplit;
      const int row = (2 * pair + (int)rank) * 128 + quad * 32 + lane;   // row of B = row of out owned by this thread
      const bool row_ok = row < p.n_rows;
      mbar_wait(&bars->tmem_full, it & 1);
      fence_after_sync();
      if (group < p.nacc) {
        const int ncols = min(256, p.m_rows - group * 256);
        const bool split = p.ksplit > 1;
        float* o = split ? p.slab + ks * p.slab_stride + (int64_t)row * p.m_rows + group * 256
                         : p.out + (int64_t)row * p.ldo + group * 256;
        const float alpha = split ? 1.f : p.alpha;
        const int accumulate = split ? 0 : p.accumulate;
        for (int c0 = 0; c0 < ncols; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tmem_base + lane_addr + group * 256 + c0, v);
          tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (c0 + j + 3 < ncols) {
                float4 r = make_float4(__uint_as_float(v[j]) * alpha, __uint_as_float(v[j + 1]) * alpha,
                                       __uint_as_float(v[j + 2]) * alpha, __uint_as_float(v[j + 3]) * alpha);
                float4* dst = reinterpret_cast<float4*>(o + c0 + j);
                if (accumulate) { const float4 q = *dst; r.x += q.x; r.y += q.y; r.z += q.z; r.w += q.w; }
                *dst = r;
              } else {
                for (int e = 0; e < 4; ++e)
                  if (c0 + j + e < ncols) {
                    const float r = __uint_as_float(v[j + e]) * alpha;
                    o[c0 + j + e] = accumulate ? o[c0 + j + e] + r : r;
                  }
              }
            }
          }
        }
      }
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&bars->tmem_empty, 0);
    }
  }

  fence_before_sync();
  cluster_sync_all();
  if (warp == 1) {
    fence_after_sync();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

bool gemm_wide2_enabled() {
  return tune_get(ILFDT_TUNE_GEMM_WIDE2) != 0;
}
// shapes the pair kernel takes: a short A (all of its rows fit two accumulators), many B rows, a long reduction
bool gemm_wide2_supported(int64_t M, int64_t N, int64_t Kr, const float* out, int64_t ldo) {
  return M >= 16 && M <= 512 && M % 16 == 0 && (M <= 256 || M % 32 == 0) && N >= 1024 && Kr >= 1024 && Kr % 8 == 0 &&
         ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
}

// out[n, m] (+)= alpha * sum_ks slab[ks][n, m]   (m_rows % 4 == 0)
__global__ void __launch_bounds__(256) w2_reduce_kernel(const float* __restrict__ slab, int ksplit, long long stride,
                                                        int n_rows, int m_rows, float alpha, int accumulate,
                                                        float* __restrict__ out, long long ldo) {
  const long long i4 = (long long)blockIdx.x * 256 + threadIdx.x;     // float4 index over [n_rows, m_rows]
  const int m4 = m_rows >> 2;
  if (i4 >= (long long)n_rows * m4) return;
  const int n = (int)(i4 / m4), m = (int)(i4 - (long long)n * m4) * 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int ks = 0; ks < ksplit; ++ks) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(slab + ks * stride + (long long)n * m_rows + m));
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  float4* dst = reinterpret_cast<float4*>(out + (long long)n * ldo + m);
  float4 r = make_float4(acc.x * alpha, acc.y * alpha, acc.z * alpha, acc.w * alpha);
  if (accumulate) { const float4 q = *dst; r.x += q.x; r.y += q.y; r.z += q.z; r.w += q.w; }
  *dst = r;
}

static int w2_clusters() {
  return std::max(1, sm_count() / 2);
}
// how the reduction is cut for a launch with few B rows: 1 = not at all
static int w2_ksplit(int64_t M, int64_t N, int64_t Kr) {
  const int64_t pairs = cdiv(N, 256), nkb = cdiv(Kr, 64);
  if (M % 4 != 0 || pairs * 2 > w2_clusters() || nkb < 16) return 1;
  return (int)std::max<int64_t>(1, std::min<int64_t>(w2_clusters() / pairs, nkb / 4));
}
bool gemm_wide2_split_supported(int64_t M, int64_t N, int64_t Kr, const float* out, int64_t ldo) {
  return M >= 16 && M <= 512 && M % 16 == 0 && (M <= 256 || M % 32 == 0) && N >= 1 && Kr >= 1024 && Kr % 8 == 0 && ldo % 4 == 0 &&
         (reinterpret_cast<uintptr_t>(out) & 15) == 0 && w2_ksplit(M, N, Kr) > 1;
}
int64_t gemm_wide2_ws_floats(int64_t M, int64_t N, int64_t Kr) {
  const int ks = w2_ksplit(M, N, Kr);
  return ks > 1 ? (int64_t)ks * round_up(N, 256) * M : 0;
}

// out[n, m] (+)= alpha * sum_k B[n,k] A[m,k]; A [M, Kr], B [N, Kr] bf16 row-major. ws != nullptr: split reduction.
// A given as [Kr, M] (reduction outermost): both halves of every accumulator must be full 128-column boxes
bool gemm_wide2_mn_supported(int64_t M, int64_t N, int64_t Kr, const float* out, int64_t ldo) {
  return (M == 256 || M == 512) && N >= 1 && Kr >= 1024 && ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
         (gemm_wide2_supported(M, N, Kr, out, ldo) || w2_ksplit(M, N, Kr) > 1);
}

int32_t launch_gemm_wide2(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha, int accumulate,
                          float* out, int64_t ldo, cudaStream_t st, float* ws, int col_mn) {
  W2Params p{};
  p.col_mn = col_mn;
  p.ksplit = ws != nullptr ? w2_ksplit(M, N, Kr) : 1;
  p.kb_per = (int)cdiv(cdiv(Kr, 64), p.ksplit);
  p.ksplit = (int)cdiv(cdiv(Kr, 64), p.kb_per);
  p.slab = ws;
  p.slab_stride = round_up(N, 256) * M;
  p.n_rows = (int)N;
  p.m_rows = (int)M;
  p.nkb = (int)cdiv(Kr, 64);
  p.num_pairs = (int)cdiv(N, 256);
  p.nacc = (int)cdiv(M, 256);
  for (int a = 0; a < p.nacc; ++a) {
    p.acc_n[a] = (int)round_up(std::min<int64_t>(256, M - a * 256), 32);   // halves stay multiples of 16 rows
    p.idesc[a] = make_idesc_bf16(256, p.acc_n[a], 0, col_mn);
  }
  p.alpha = alpha;
  p.accumulate = accumulate;
  p.out = out;
  p.ldo = ldo;
  CUtensorMap tmLane, tmCol;
  int32_t rc = make_tmap_bf16(&tmLane, B, N, Kr, Kr, 128);
  if (rc != ILFDT_OK) return rc;
  if (col_mn) rc = make_tmap_bf16(&tmCol, A, Kr, M, M, 64);    // [Kr, M] row-major, box = 64 k-rows x 64 columns
  else rc = make_tmap_bf16(&tmCol, A, M, Kr, Kr, 128);
  if (rc != ILFDT_OK) return rc;
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
  cfg.blockDim = dim3(W2_THREADS);
  cfg.dynamicSmemBytes = W2_SMEM_BYTES;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (configured_dev != dev) {
    ILFDT_CUDA_OK(cudaFuncSetAttribute(gemm_wide2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, W2_SMEM_BYTES));
    cfg.gridDim = dim3(2 * (sm_count() / 2));
    int n = 0;
    ILFDT_CUDA_OK(cudaOccupancyMaxActiveClusters(&n, gemm_wide2_kernel, &cfg));
    ILFDT_REQUIRE(n >= 1, ILFDT_E_CUDA, "gemm_wide2: no CTA pair can be resident");
    max_clusters = std::min(n, sm_count() / 2);
    configured_dev = dev;
  }
  cfg.gridDim = dim3(2 * std::min(p.num_pairs * p.ksplit, max_clusters));
  ILFDT_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_wide2_kernel, tmLane, tmCol, p));
  if (p.ksplit > 1) {
    const long long n4 = (long long)N * (M / 4);
    w2_reduce_kernel<<<(unsigned)cdiv(n4, 256), 256, 0, st>>>(ws, p.ksplit, p.slab_stride, (int)N, (int)M, alpha, accumulate, out,
                                                            ldo);
    ILFDT_LAUNCH_OK();
  }
  return ILFDT_OK;
}

}  // namespace ilfdt
