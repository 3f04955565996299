// extern "C" entry points of libilfdt.so (see include/ilfdt.h for the contract of each).
#include "common.cuh"
#include "launch.h"
#include <algorithm>
#include <atomic>

namespace ilfdt {

// tuning knobs: explicit API state instead of environment variables read into statics (defaults = the tuned values)
static std::atomic<int64_t> g_tune[ILFDT_TUNE_COUNT_] = {{1024}, {32}, {1}, {1}, {1}, {1}, {128}, {1}, {0}, {1}, {1}, {1024}};
int64_t tune_get(int key) { return (key >= 0 && key < ILFDT_TUNE_COUNT_) ? g_tune[key].load(std::memory_order_relaxed) : 0; }

int sm_count() {
  static thread_local int cached_dev = -1, cached = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess) cached = n;
    cached_dev = dev;
  }
  return cached;
}

int32_t require_sm100() {
  static thread_local int ok_dev = -1;
  int dev = 0;
  ILFDT_CUDA_OK(cudaGetDevice(&dev));
  if (dev == ok_dev) return ILFDT_OK;
  int major = 0, minor = 0;
  ILFDT_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  ILFDT_CUDA_OK(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  ILFDT_REQUIRE(major == 10, ILFDT_E_ARCH,
                "libilfdt is built for sm_100a only; device %d is sm_%d%d and there is no fallback path", dev, major,
                minor);
  ok_dev = dev;
  return ILFDT_OK;
}

// F32 route of the InfoNCE step: logits are streamed through the workspace in row chunks.
// F32 workspace: [chunk of logits R x B floats][2 b doubles: per-row d scale contributions of the two directions]
static int64_t nce_f32_logits_bytes(int64_t b, int64_t B);
static double* nce_f32_row_acc(void* ws, int64_t b, int64_t B) {
  return reinterpret_cast<double*>(reinterpret_cast<char*>(ws) + nce_f32_logits_bytes(b, B));
}
static int64_t nce_f32_chunk_rows(int64_t b, int64_t B) {
  const int64_t budget_elems = 1ll << 26;  // 256 MiB of fp32
  return std::max<int64_t>(1, std::min<int64_t>(b, budget_elems / std::max<int64_t>(B, 1)));
}

static int64_t nce_f32_logits_bytes(int64_t b, int64_t B) { return round_up(nce_f32_chunk_rows(b, B) * B * 4, 16); }
// fp32-grade logits on the tensor cores: L = scale * X . G^T as one bf16 contraction over the 3-term split of both
// operands (launch_split3_bf16; the reduction is the embedding width, a few dozen leading-term steps, see
// kernels_simt.cu). The products with a reduction over B (dX = P . G) stay on the SIMT kernel.
static bool nce_f32_x3(int64_t D) { return tune_get(ILFDT_TUNE_F32_X3) != 0 && D % 8 == 0 && D <= 1024; }
static int64_t nce_f32_x3_bytes(int64_t b, int64_t B, int64_t D) {
  return (D % 8 == 0 && D <= 1024) ? round_up((b + B) * 6 * D * 2, 256) + 256 : 0;
}
static int64_t nce_f32_ws_bytes(int64_t b, int64_t B, int64_t D = 0) {
  return round_up(nce_f32_logits_bytes(b, B) + 2 * b * 8, 256) + nce_f32_x3_bytes(b, B, D);
}
struct NceX3 {
  bool on = false;
  char* X6 = nullptr;   // [b, 6 D] bf16, B-operand order
  char* G6 = nullptr;   // [B, 6 D] bf16, A-operand order
};
static int32_t nce_x3_prepare(NceX3& x3, const float* X, const float* G, int64_t b, int64_t B, int64_t D, void* ws,
                              int64_t ws_bytes, cudaStream_t st) {
  x3.on = nce_f32_x3(D) && ws_bytes >= nce_f32_ws_bytes(b, B, D);
  if (!x3.on) return ILFDT_OK;
  x3.X6 = reinterpret_cast<char*>(ws) + round_up(nce_f32_logits_bytes(b, B) + 2 * b * 8, 256);
  x3.G6 = x3.X6 + round_up(b * 6 * D * 2, 256);
  int32_t rc = launch_split3_bf16(X, nullptr, b, D, 1, x3.X6, st);
  if (rc != ILFDT_OK) return rc;
  return launch_split3_bf16(G, nullptr, B, D, 0, x3.G6, st);
}
// L[rows, B] = scale * X[r0 : r0 + rows, :] . G^T
static int32_t nce_f32_logits(const NceX3& x3, const float* X, const float* G, int64_t r0, int64_t rows, int64_t B, int64_t D,
                              const float* scale, float* L, cudaStream_t st) {
  if (x3.on)
    return launch_gemm_nt_bf16(x3.G6, x3.X6 + r0 * 6 * D * 2, B, rows, 6 * D, 1.f, scale, 0, L, B, st);
  return launch_gemm_f32(X + r0 * D, D, 1, G, D, 1, L, B, 1, rows, B, D, 1.f, scale, 0, st);
}

static int32_t nce_lse_f32(const float* X, const float* G, int64_t b, int64_t B, int64_t D, const float* scale,
                           int64_t label_offset, float* lse, float* diag, void* ws, int64_t ws_bytes, cudaStream_t st) {
  const int64_t R = nce_f32_chunk_rows(b, B);
  ILFDT_REQUIRE(ws != nullptr && ws_bytes >= R * B * 4, ILFDT_E_INVALID_ARG, "nce_lse(f32): workspace too small");
  float* L = reinterpret_cast<float*>(ws);
  NceX3 x3;
  int32_t rc0 = nce_x3_prepare(x3, X, G, b, B, D, ws, ws_bytes, st);
  if (rc0 != ILFDT_OK) return rc0;
  for (int64_t r0 = 0; r0 < b; r0 += R) {
    const int64_t rows = std::min(R, b - r0);
    int32_t rc = nce_f32_logits(x3, X, G, r0, rows, B, D, scale, L, st);
    if (rc != ILFDT_OK) return rc;
    rc = launch_lse_rows(L, B, rows, B, label_offset + r0, lse + r0, diag + r0, st);
    if (rc != ILFDT_OK) return rc;
  }
  return ILFDT_OK;
}

static int32_t nce_fwd_f32(const float* x_img, const float* x_txt, const float* g_img, const float* g_txt, int64_t b,
                           int64_t B, int64_t D, const float* scale, int64_t label_offset, float* lse_i, float* lse_t,
                           float* diag_i, float* diag_t, void* ws, int64_t ws_bytes, cudaStream_t st) {
  // logits_per_image = x_img . g_txt^T ; logits_per_text = x_txt . g_img^T  (clip_fdt.py:421-422)
  int32_t rc = nce_lse_f32(x_img, g_txt, b, B, D, scale, label_offset, lse_i, diag_i, ws, ws_bytes, st);
  if (rc != ILFDT_OK) return rc;
  return nce_lse_f32(x_txt, g_img, b, B, D, scale, label_offset, lse_t, diag_t, ws, ws_bytes, st);
}

static int32_t nce_bwd_f32(const float* x_img, const float* x_txt, const float* g_img, const float* g_txt, int64_t b,
                           int64_t B, int64_t D, const float* scale, int64_t label_offset, const float* lse_i,
                           const float* lse_t, const float* grad_out, float* dx_img, float* dx_txt, float* dg_img,
                           float* dg_txt, float* dscale, void* ws, int64_t ws_bytes, cudaStream_t st) {
  const int64_t R = nce_f32_chunk_rows(b, B);
  ILFDT_REQUIRE(ws != nullptr && ws_bytes >= nce_f32_ws_bytes(b, B), ILFDT_E_INVALID_ARG, "nce_bwd(f32): workspace too small");
  float* L = reinterpret_cast<float*>(ws);
  double* row_acc = nce_f32_row_acc(ws, b, B);
  const float coef = 1.f / (2.f * (float)b);  // CE mean over the local rows, then /2 (loss.py:47-49)
  for (int dir = 0; dir < 2; ++dir) {
    const float* X = dir == 0 ? x_img : x_txt;
    const float* G = dir == 0 ? g_txt : g_img;
    const float* lse = dir == 0 ? lse_i : lse_t;
    float* dX = dir == 0 ? dx_img : dx_txt;
    float* dG = dir == 0 ? dg_txt : dg_img;
    NceX3 x3;
    int32_t rc0 = nce_x3_prepare(x3, X, G, b, B, D, ws, ws_bytes, st);
    if (rc0 != ILFDT_OK) return rc0;
    for (int64_t r0 = 0; r0 < b; r0 += R) {
      const int64_t rows = std::min(R, b - r0);
      int32_t rc = nce_f32_logits(x3, X, G, r0, rows, B, D, scale, L, st);
      if (rc != ILFDT_OK) return rc;
      rc = launch_softmax_grad_rows(L, L, B, rows, B, label_offset + r0, lse + r0, grad_out, coef, scale,
                                    row_acc + dir * b + r0, st);
      if (rc != ILFDT_OK) return rc;
      // dX[r,:] = scale * sum_c P[r,c] G[c,:]
      rc = launch_gemm_f32(L, B, 1, G, 1, D, dX + r0 * D, D, 1, rows, D, B, 1.f, scale, 0, st);
      if (rc != ILFDT_OK) return rc;
      // dG[c,:] (+)= scale * sum_r P[r,c] X[r,:]
      rc = launch_gemm_f32(L, 1, B, X + r0 * D, 1, D, dG, D, 1, B, D, rows, 1.f, scale, r0 > 0 ? 1 : 0, st);
      if (rc != ILFDT_OK) return rc;
    }
  }
  return launch_sum_rows(row_acc, 2 * b, dscale, st);
}

// Gathered form of the F32 route (see ilfdt_nce_bwd_gathered): per direction, for a chunk of local rows,
// L = scale * X . G^T, P = lane softmax + stream softmax terms, dX = scale * P . G.
static int32_t nce_bwd_gathered_f32(const float* x_img, const float* x_txt, const float* g_img, const float* g_txt,
                                    int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset,
                                    const float* lse_i_all, const float* lse_t_all, const float* go_all, float* dx_img,
                                    float* dx_txt, float* dscale, void* ws, int64_t ws_bytes, cudaStream_t st) {
  const int64_t R = nce_f32_chunk_rows(b, B);
  ILFDT_REQUIRE(ws != nullptr && ws_bytes >= nce_f32_ws_bytes(b, B), ILFDT_E_INVALID_ARG, "nce_bwd(f32): workspace too small");
  float* L = reinterpret_cast<float*>(ws);
  double* row_acc = nce_f32_row_acc(ws, b, B);
  const float coef = 1.f / (2.f * (float)b);
  for (int dir = 0; dir < 2; ++dir) {
    const float* X = dir == 0 ? x_img : x_txt;
    const float* G = dir == 0 ? g_txt : g_img;
    const float* lse_lane = (dir == 0 ? lse_i_all : lse_t_all) + label_offset;
    const float* lse_stream = dir == 0 ? lse_t_all : lse_i_all;
    float* dX = dir == 0 ? dx_img : dx_txt;
    NceX3 x3;
    int32_t rc0 = nce_x3_prepare(x3, X, G, b, B, D, ws, ws_bytes, st);
    if (rc0 != ILFDT_OK) return rc0;
    for (int64_t r0 = 0; r0 < b; r0 += R) {
      const int64_t rows = std::min(R, b - r0);
      int32_t rc = nce_f32_logits(x3, X, G, r0, rows, B, D, scale, L, st);
      if (rc != ILFDT_OK) return rc;
      rc = launch_softmax_grad_sym_rows(L, L, B, rows, B, label_offset + r0, lse_lane + r0, lse_stream,
                                        go_all + label_offset / b, go_all, b, coef, scale, row_acc + dir * b + r0, st);
      if (rc != ILFDT_OK) return rc;
      rc = launch_gemm_f32(L, B, 1, G, 1, D, dX + r0 * D, D, 1, rows, D, B, 1.f, scale, 0, st);
      if (rc != ILFDT_OK) return rc;
    }
  }
  return launch_sum_rows(row_acc, 2 * b, dscale, st);
}

}  // namespace ilfdt

using namespace ilfdt;

#define ST(s) reinterpret_cast<cudaStream_t>(s)
#define ARCH_GUARD()                        \
  do {                                      \
    int32_t rc__ = require_sm100();         \
    if (rc__ != ILFDT_OK) return rc__;      \
  } while (0)
#define DTYPE_GUARD(dt) \
  ILFDT_REQUIRE((dt) == ILFDT_F32 || (dt) == ILFDT_BF16, ILFDT_E_INVALID_ARG, "dtype must be ILFDT_F32 or ILFDT_BF16")

extern "C" {

int32_t ilfdt_abi_version(void) { return ILFDT_ABI_VERSION; }
const char* ilfdt_last_error(void) { return err_buf(); }

int32_t ilfdt_tune_set(int32_t key, int64_t value) {
  ILFDT_REQUIRE(key >= 0 && key < ILFDT_TUNE_COUNT_, ILFDT_E_INVALID_ARG, "tune_set: unknown key %d", key);
  ILFDT_REQUIRE(value >= 0, ILFDT_E_INVALID_ARG, "tune_set: negative value");
  ILFDT_REQUIRE(key != ILFDT_TUNE_DSD_TILE || value == 32 || value == 128, ILFDT_E_INVALID_ARG,
                "tune_set: ILFDT_TUNE_DSD_TILE is 32 or 128");
  g_tune[key].store(value, std::memory_order_relaxed);
  return ILFDT_OK;
}
int64_t ilfdt_tune_get(int32_t key) { return tune_get(key); }

int32_t ilfdt_device_check(int32_t device) {
  int major = 0;
  ILFDT_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  ILFDT_REQUIRE(major == 10, ILFDT_E_ARCH, "device %d is not an sm_100 part (major=%d)", device, major);
  return ILFDT_OK;
}

int32_t ilfdt_cast_bf16(const float* src, void* dst_bf16, int64_t n, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(n >= 0 && (n == 0 || (src && dst_bf16)), ILFDT_E_INVALID_ARG, "cast_bf16: null pointer");
  return launch_cast_bf16(src, dst_bf16, n, ST(stream));
}

int32_t ilfdt_cast_bf16_rows(const float* src, const float* rowmask, int64_t rows, int64_t cols, void* dst_bf16,
                             void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && cols >= 0 && (rows * cols == 0 || (src && dst_bf16)), ILFDT_E_INVALID_ARG,
                "cast_bf16_rows: bad arguments");
  return launch_cast_bf16_rows(src, rowmask, rows, cols, dst_bf16, ST(stream));
}

int32_t ilfdt_split3_bf16(const float* src, const float* rowmask, int64_t rows, int64_t cols, int32_t side, void* dst_bf16,
                          void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && cols >= 0 && (side == 0 || side == 1) && (rows * cols == 0 || (src && dst_bf16)),
                ILFDT_E_INVALID_ARG, "split3_bf16: bad arguments");
  return launch_split3_bf16(src, rowmask, rows, cols, side, dst_bf16, ST(stream));
}

int32_t ilfdt_mask_rows_bf16(const void* src_bf16, const float* rowmask, int64_t rows, int64_t cols, void* dst_bf16,
                             void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && cols >= 0 && (rows * cols == 0 || (src_bf16 && rowmask && dst_bf16)), ILFDT_E_INVALID_ARG,
                "mask_rows_bf16: bad arguments");
  return launch_mask_rows_bf16(src_bf16, rowmask, rows, cols, dst_bf16, ST(stream));
}

int32_t ilfdt_transpose_bf16(const void* src, int32_t src_dtype, int64_t rows, int64_t cols, void* dst_bf16,
                             void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(src_dtype);
  ILFDT_REQUIRE(rows >= 0 && cols >= 0 && (rows * cols == 0 || (src && dst_bf16)), ILFDT_E_INVALID_ARG,
                "transpose_bf16: bad arguments");
  return launch_transpose_bf16(src, src_dtype, rows, cols, dst_bf16, ST(stream));
}

int64_t ilfdt_ground_fwd_ws_bytes(int32_t dtype, int64_t b, int64_t n_tok, int64_t K, int64_t D) {
  (void)n_tok; (void)D;
  return dtype == ILFDT_F32 ? b * K * 8 : 0;
}

int32_t ilfdt_ground_fwd(const void* q, const void* sd, int32_t dtype, const float* mask, int64_t b, int64_t n_tok,
                         int64_t K, int64_t D, float sqrt_d, float temperature, int32_t pool_type, float* pooled,
                         int16_t* arg, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && K >= 1 && D >= 1, ILFDT_E_INVALID_ARG, "ground_fwd: bad shape");
  ILFDT_REQUIRE(b == 0 || (q && sd && pooled), ILFDT_E_INVALID_ARG, "ground_fwd: null pointer");
  ILFDT_REQUIRE(sqrt_d > 0.f && temperature != 0.f, ILFDT_E_INVALID_ARG, "ground_fwd: bad scale");
  ILFDT_REQUIRE(pool_type >= ILFDT_POOL_MAX && pool_type <= ILFDT_POOL_MEAN, ILFDT_E_INVALID_ARG,
                "ground_fwd: bad pool_type");
  ILFDT_REQUIRE(pool_type != ILFDT_POOL_MAX || arg != nullptr || b == 0, ILFDT_E_INVALID_ARG,
                "ground_fwd: arg required for max pooling");
  if (b == 0) return ILFDT_OK;
  if (dtype == ILFDT_F32)
    return launch_ground_fwd_f32(reinterpret_cast<const float*>(q), reinterpret_cast<const float*>(sd), mask, b,
                                 n_tok, K, D, sqrt_d, temperature, pool_type, pooled, arg, ws, ws_bytes, ST(stream));
  ILFDT_REQUIRE(pool_type == ILFDT_POOL_MAX, ILFDT_E_UNSUPPORTED_SHAPE,
                "ground_fwd(bf16): only max pooling is fused on the tensor-core path (use ILFDT_F32)");
  ILFDT_REQUIRE(temperature > 0.f, ILFDT_E_INVALID_ARG, "ground_fwd(bf16): temperature must be positive");
  return launch_ground_fwd_bf16(q, sd, mask, b, n_tok, K, D, sqrt_d, temperature, pooled, arg, ST(stream));
}

int64_t ilfdt_text_plan_ints(int64_t b) { return text_plan_ints(b); }
int32_t ilfdt_ground_fwd_text_supported(int64_t n_tok) { return ground_fwd_text_supported(n_tok) ? 1 : 0; }

int32_t ilfdt_text_compact(const void* q, int32_t q_dtype, const float* mask, int64_t b, int64_t n_tok, int64_t D,
                           void* groups_ws, int32_t* plan, void* q_compact_bf16, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(q_dtype);
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && D >= 1, ILFDT_E_INVALID_ARG, "text_compact: bad shape");
  ILFDT_REQUIRE(b == 0 || (q && mask && groups_ws && plan && q_compact_bf16), ILFDT_E_INVALID_ARG,
                "text_compact: null pointer");
  return launch_text_compact(q, q_dtype, mask, b, n_tok, D, reinterpret_cast<uint8_t*>(groups_ws), plan, q_compact_bf16,
                             ST(stream));
}

int32_t ilfdt_text_compact_x3(const float* q, const float* mask, int64_t b, int64_t n_tok, int64_t D, void* groups_ws,
                              int32_t* plan, void* q_compact_x3_bf16, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && D >= 1, ILFDT_E_INVALID_ARG, "text_compact_x3: bad shape");
  ILFDT_REQUIRE(b == 0 || (q && mask && groups_ws && plan && q_compact_x3_bf16), ILFDT_E_INVALID_ARG,
                "text_compact_x3: null pointer");
  return launch_text_compact(q, ILFDT_F32, mask, b, n_tok, D, reinterpret_cast<uint8_t*>(groups_ws), plan,
                             q_compact_x3_bf16, ST(stream), 1);
}

int32_t ilfdt_ground_fwd_text(const void* q_compact_bf16, const void* sd_bf16, const int32_t* plan, int64_t b,
                              int64_t n_tok, int64_t K, int64_t D, float sqrt_d, float temperature, float* pooled,
                              int16_t* arg, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && K >= 1 && D >= 1, ILFDT_E_INVALID_ARG, "ground_fwd_text: bad shape");
  ILFDT_REQUIRE(b == 0 || (q_compact_bf16 && sd_bf16 && plan && pooled && arg), ILFDT_E_INVALID_ARG,
                "ground_fwd_text: null pointer");
  return launch_ground_fwd_text_bf16(q_compact_bf16, sd_bf16, plan, b, n_tok, K, D, sqrt_d, temperature, pooled, arg,
                                     ST(stream));
}

int32_t ilfdt_sparsemax_fwd(const float* z, int64_t rows, int64_t K, float* w, void* w_bf16, float* tau,
                            int32_t* nnz, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && K >= 1 && (rows == 0 || (z && (w || w_bf16))), ILFDT_E_INVALID_ARG,
                "sparsemax_fwd: bad arguments");
  return launch_sparsemax_fwd(z, rows, K, w, w_bf16, tau, nnz, ST(stream));
}

int32_t ilfdt_sparsemax_bwd(const float* w, const float* g, int64_t rows, int64_t K, float* dz, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && K >= 1 && (rows == 0 || (w && g && dz)), ILFDT_E_INVALID_ARG,
                "sparsemax_bwd: bad arguments");
  return launch_sparsemax_bwd(w, g, rows, K, dz, ST(stream));
}

int32_t ilfdt_gemm_nt(const void* A, const void* B, int32_t dtype, int64_t M, int64_t N, int64_t Kr, float alpha,
                      int32_t accumulate, float* out, int64_t ldo, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(M >= 0 && N >= 0 && Kr >= 1 && ldo >= M, ILFDT_E_INVALID_ARG, "gemm_nt: bad shape");
  ILFDT_REQUIRE(M * N == 0 || (A && B && out), ILFDT_E_INVALID_ARG, "gemm_nt: null pointer");
  if (dtype == ILFDT_F32)
    return launch_gemm_f32(reinterpret_cast<const float*>(A), Kr, 1, reinterpret_cast<const float*>(B), Kr, 1, out, 1,
                           ldo, M, N, Kr, alpha, nullptr, accumulate, ST(stream));
  return launch_gemm_nt_bf16(A, B, M, N, Kr, alpha, nullptr, accumulate, out, ldo, ST(stream));
}

int64_t ilfdt_gemm_nt_splitk_ws_bytes(int64_t M, int64_t N, int64_t Kr) {
  return gemm_wide2_enabled() ? gemm_wide2_ws_floats(M, N, Kr) * 4 : 0;
}

int32_t ilfdt_gemm_nt_splitk(const void* A_bf16, const void* B_bf16, int64_t M, int64_t N, int64_t Kr, float alpha,
                             int32_t accumulate, float* out, int64_t ldo, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(M >= 0 && N >= 0 && Kr >= 1 && ldo >= M, ILFDT_E_INVALID_ARG, "gemm_nt_splitk: bad shape");
  ILFDT_REQUIRE(M == 0 || N == 0 || (A_bf16 && B_bf16 && out), ILFDT_E_INVALID_ARG, "gemm_nt_splitk: null pointer");
  if (M == 0 || N == 0) return ILFDT_OK;
  const int64_t need = ilfdt_gemm_nt_splitk_ws_bytes(M, N, Kr);
  if (need > 0 && gemm_wide2_split_supported(M, N, Kr, out, ldo)) {
    ILFDT_REQUIRE(ws != nullptr && ws_bytes >= need, ILFDT_E_INVALID_ARG, "gemm_nt_splitk: workspace too small");
    return launch_gemm_wide2(A_bf16, B_bf16, M, N, Kr, alpha, accumulate, out, ldo, ST(stream), reinterpret_cast<float*>(ws));
  }
  return launch_gemm_nt_bf16(A_bf16, B_bf16, M, N, Kr, alpha, nullptr, accumulate, out, ldo, ST(stream));
}

int32_t ilfdt_gemm_nn_supported(int64_t M, int64_t N, int64_t Kr) {
  return gemm_wide2_enabled() && gemm_wide2_mn_supported(M, N, Kr, nullptr, 4) ? 1 : 0;
}

int32_t ilfdt_gemm_nn(const void* A_bf16, const void* B_bf16, int64_t M, int64_t N, int64_t Kr, float alpha, int32_t accumulate,
                      float* out, int64_t ldo, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(M >= 0 && N >= 0 && Kr >= 1 && ldo >= M, ILFDT_E_INVALID_ARG, "gemm_nn: bad shape");
  ILFDT_REQUIRE(M == 0 || N == 0 || (A_bf16 && B_bf16 && out), ILFDT_E_INVALID_ARG, "gemm_nn: null pointer");
  if (M == 0 || N == 0) return ILFDT_OK;
  ILFDT_REQUIRE(gemm_wide2_enabled() && gemm_wide2_mn_supported(M, N, Kr, out, ldo), ILFDT_E_UNSUPPORTED_SHAPE,
                "gemm_nn: M in {256, 512}, Kr >= 1024, 16-byte aligned out with ldo %% 4 == 0 and either N >= 1024 or a split "
                "reduction required (got M=%lld N=%lld Kr=%lld); see ilfdt_gemm_nn_supported", (long long)M, (long long)N,
                (long long)Kr);
  ILFDT_REQUIRE((reinterpret_cast<uintptr_t>(A_bf16) & 15) == 0, ILFDT_E_INVALID_ARG, "gemm_nn: A must be 16-byte aligned");
  const int64_t need = ilfdt_gemm_nt_splitk_ws_bytes(M, N, Kr);
  float* slabs = nullptr;
  if (need > 0 && gemm_wide2_split_supported(M, N, Kr, out, ldo)) {
    ILFDT_REQUIRE(ws != nullptr && ws_bytes >= need, ILFDT_E_INVALID_ARG, "gemm_nn: workspace too small");
    slabs = reinterpret_cast<float*>(ws);
  }
  return launch_gemm_wide2(A_bf16, B_bf16, M, N, Kr, alpha, accumulate, out, ldo, ST(stream), slabs, 1);
}

int32_t ilfdt_gemm_tn(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha, int32_t accumulate,
                      float* out, int64_t ldo, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(M >= 0 && N >= 0 && Kr >= 1 && ldo >= M, ILFDT_E_INVALID_ARG, "gemm_tn: bad shape");
  ILFDT_REQUIRE((M == 0 || N == 0) || (A && B && out), ILFDT_E_INVALID_ARG, "gemm_tn: null pointer");
  return launch_gemm_tn_bf16(A, B, M, N, Kr, alpha, accumulate, out, ldo, ST(stream));
}

int32_t ilfdt_linear_bf16(const void* W_bf16, const void* X_bf16, int64_t M, int64_t N, int64_t Kr, const float* bias,
                          void* out_bf16, int64_t ldo, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(M >= 0 && N >= 0 && Kr >= 1 && ldo >= M, ILFDT_E_INVALID_ARG, "linear_bf16: bad shape");
  ILFDT_REQUIRE((M == 0 || N == 0) || (W_bf16 && X_bf16 && out_bf16), ILFDT_E_INVALID_ARG, "linear_bf16: null pointer");
  ILFDT_REQUIRE(M % 128 == 0 && Kr % 64 == 0, ILFDT_E_UNSUPPORTED_SHAPE,
                "linear_bf16: out features %% 128 and in features %% 64 required (got %lld, %lld)", (long long)M, (long long)Kr);
  return launch_linear_bf16(W_bf16, X_bf16, M, N, Kr, bias, out_bf16, ldo, ST(stream));
}

int64_t ilfdt_gemm_tn_splitk_ws_bytes(int64_t M, int64_t N, int64_t Kr) { return gemm_tn_splitk_ws_floats(M, N, Kr) * 4; }

int32_t ilfdt_gemm_tn_splitk(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha,
                             int32_t accumulate, float* out, int64_t ldo, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(M >= 0 && N >= 0 && Kr >= 1 && ldo >= M, ILFDT_E_INVALID_ARG, "gemm_tn_splitk: bad shape");
  ILFDT_REQUIRE((M == 0 || N == 0) || (A && B && out), ILFDT_E_INVALID_ARG, "gemm_tn_splitk: null pointer");
  ILFDT_REQUIRE(ws_bytes >= ilfdt_gemm_tn_splitk_ws_bytes(M, N, Kr), ILFDT_E_INVALID_ARG,
                "gemm_tn_splitk: workspace too small");
  return launch_gemm_tn_splitk_bf16(A, B, M, N, Kr, alpha, accumulate, out, ldo, (float*)ws, ST(stream));
}

int32_t ilfdt_gemm_f32_strided(const float* A, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_rs, int64_t b_cs,
                               float* C, int64_t c_rs, int64_t c_cs, int64_t M, int64_t N, int64_t Kr, float alpha,
                               int32_t accumulate, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(M >= 0 && N >= 0 && Kr >= 1, ILFDT_E_INVALID_ARG, "gemm_f32_strided: bad shape");
  ILFDT_REQUIRE(M * N == 0 || (A && B && C), ILFDT_E_INVALID_ARG, "gemm_f32_strided: null pointer");
  return launch_gemm_f32(A, a_rs, a_cs, B, b_rs, b_cs, C, c_rs, c_cs, M, N, Kr, alpha, nullptr, accumulate, ST(stream));
}

int32_t ilfdt_ground_bwd(float* g, const float* w, const int16_t* arg, const void* q, const void* sd, int32_t dtype,
                         const float* mask, const int32_t* q_row_base, const int32_t* nnz, int32_t* regime_ws, int64_t b,
                         int64_t n_tok, int64_t K, int64_t D, float sqrt_d, float temperature, void* dq, int32_t dq_dtype,
                         float* dsd, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  DTYPE_GUARD(dq_dtype);
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && K >= 1 && D >= 1, ILFDT_E_INVALID_ARG, "ground_bwd: bad shape");
  ILFDT_REQUIRE(b == 0 || (g && arg && q && sd && dq && dsd), ILFDT_E_INVALID_ARG, "ground_bwd: null pointer");
  return launch_ground_bwd(g, w, ILFDT_F32, arg, q, sd, dtype, mask, q_row_base, nnz, regime_ws, b, n_tok, K, D, sqrt_d,
                           temperature, dq, dq_dtype, dsd, ST(stream));
}

int32_t ilfdt_ground_bwd_w16(float* g, const void* w_bf16, const int16_t* arg, const void* q, const void* sd,
                             int32_t dtype, const float* mask, const int32_t* q_row_base, const int32_t* nnz,
                             int32_t* regime_ws, int64_t b, int64_t n_tok, int64_t K, int64_t D, float sqrt_d,
                             float temperature, void* dq, int32_t dq_dtype, float* dsd, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  DTYPE_GUARD(dq_dtype);
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && K >= 1 && D >= 1, ILFDT_E_INVALID_ARG, "ground_bwd_w16: bad shape");
  ILFDT_REQUIRE(b == 0 || (g && w_bf16 && arg && q && sd && dq && dsd), ILFDT_E_INVALID_ARG,
                "ground_bwd_w16: null pointer");
  return launch_ground_bwd(g, w_bf16, ILFDT_BF16, arg, q, sd, dtype, mask, q_row_base, nnz, regime_ws, b, n_tok, K, D, sqrt_d,
                           temperature, dq, dq_dtype, dsd, ST(stream));
}

int32_t ilfdt_att_fwd(const float* z, int64_t rows, int64_t K, int32_t kind, float* w, float* rowsum, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && K >= 1 && (rows == 0 || (z && w)), ILFDT_E_INVALID_ARG, "att_fwd: bad arguments");
  ILFDT_REQUIRE(kind == ILFDT_ATT_SOFTMAX || kind == ILFDT_ATT_SIGMOID, ILFDT_E_INVALID_ARG, "att_fwd: bad kind");
  return launch_att_fwd(z, rows, K, kind, w, rowsum, ST(stream));
}

int32_t ilfdt_att_bwd(const float* w, const float* g, const float* rowbias, int64_t rows, int64_t K, int32_t kind,
                      float* dz, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && K >= 1 && (rows == 0 || (w && g && dz)), ILFDT_E_INVALID_ARG, "att_bwd: bad arguments");
  ILFDT_REQUIRE(kind == ILFDT_ATT_SOFTMAX || kind == ILFDT_ATT_SIGMOID, ILFDT_E_INVALID_ARG, "att_bwd: bad kind");
  return launch_att_bwd(w, g, rowbias, rows, K, kind, dz, ST(stream));
}

int32_t ilfdt_rownorm_fwd(const float* u, const float* s, int64_t rows, int64_t D, float* f, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && D >= 1 && (rows == 0 || (u && s && f)), ILFDT_E_INVALID_ARG, "rownorm_fwd: bad arguments");
  return launch_rownorm_fwd(u, s, rows, D, f, ST(stream));
}

int32_t ilfdt_rownorm_bwd(const float* df, const float* f, const float* s, int64_t rows, int64_t D, float* du,
                          float* ds, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && D >= 1 && (rows == 0 || (df && f && s && du && ds)), ILFDT_E_INVALID_ARG,
                "rownorm_bwd: bad arguments");
  return launch_rownorm_bwd(df, f, s, rows, D, du, ds, ST(stream));
}

int32_t ilfdt_token_sum(const float* q, const float* mask, int64_t b, int64_t n_tok, int64_t D, float* out,
                        void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && D >= 1 && (b == 0 || (q && out)), ILFDT_E_INVALID_ARG,
                "token_sum: bad arguments");
  return launch_token_sum(q, mask, b, n_tok, D, out, ST(stream));
}

int32_t ilfdt_token_bcast(const float* v, const float* mask, int64_t b, int64_t n_tok, int64_t D, float scale,
                          float* out, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 0 && n_tok >= 1 && D >= 1 && (b == 0 || (v && out)), ILFDT_E_INVALID_ARG,
                "token_bcast: bad arguments");
  return launch_token_bcast(v, mask, b, n_tok, D, scale, out, ST(stream));
}

int32_t ilfdt_l2norm_fwd(const float* f, int64_t rows, int64_t D, float eps, float* x, void* x_bf16, float* norm,
                         void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && D >= 1 && (rows == 0 || f), ILFDT_E_INVALID_ARG, "l2norm_fwd: bad arguments");
  return launch_l2norm_fwd(f, rows, D, eps, x, x_bf16, norm, ST(stream));
}

int32_t ilfdt_l2norm_bwd(const float* x, const float* norm, const float* g, int64_t rows, int64_t D, float eps,
                         float* df, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(rows >= 0 && D >= 1 && (rows == 0 || (x && norm && g && df)), ILFDT_E_INVALID_ARG,
                "l2norm_bwd: bad arguments");
  return launch_l2norm_bwd(x, norm, g, rows, D, eps, df, ST(stream));
}

int32_t ilfdt_logits(const void* x, const void* g, int32_t dtype, int64_t b, int64_t B, int64_t D,
                     const float* scale, float* out, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(b >= 0 && B >= 0 && D >= 1 && (b * B == 0 || (x && g && out && scale)), ILFDT_E_INVALID_ARG,
                "logits: bad arguments");
  if (dtype == ILFDT_F32)
    return launch_gemm_f32(reinterpret_cast<const float*>(x), D, 1, reinterpret_cast<const float*>(g), D, 1, out, B, 1,
                           b, B, D, 1.f, scale, 0, ST(stream));
  // out[row, col]: lanes = gathered rows (col), columns = local rows -> coalesced stores
  return launch_gemm_nt_bf16(g, x, B, b, D, 1.f, scale, 0, out, B, ST(stream));
}

int64_t ilfdt_nce_ws_bytes(int32_t dtype, int64_t b, int64_t B, int64_t D) {
  if (dtype == ILFDT_F32) return nce_f32_ws_bytes(b, B, D);
  return std::max(nce_fwd_bf16_ws_bytes(b, B), nce_bwd_bf16_ws_bytes(b, B, D));
}

int32_t ilfdt_nce_fwd(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt, int32_t dtype,
                      int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset, float* lse_i,
                      float* lse_t,
                      float* diag_i, float* diag_t, float* loss, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(b >= 1 && B >= b && D >= 1, ILFDT_E_INVALID_ARG, "nce_fwd: need 1 <= b <= B");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B, ILFDT_E_INVALID_ARG, "nce_fwd: labels out of range");
  ILFDT_REQUIRE(x_img && x_txt && g_img && g_txt && lse_i && lse_t && diag_i && diag_t && loss && scale, ILFDT_E_INVALID_ARG,
                "nce_fwd: null pointer");
  int32_t rc;
  if (dtype == ILFDT_F32)
    rc = nce_fwd_f32(reinterpret_cast<const float*>(x_img), reinterpret_cast<const float*>(x_txt),
                     reinterpret_cast<const float*>(g_img), reinterpret_cast<const float*>(g_txt), b, B, D, scale,
                     label_offset, lse_i, lse_t, diag_i, diag_t, ws, ws_bytes, ST(stream));
  else
    rc = launch_nce_fwd_bf16(x_img, x_txt, g_img, g_txt, b, B, D, scale, label_offset, lse_i, lse_t, diag_i, diag_t,
                             ws, ws_bytes, ST(stream));
  if (rc != ILFDT_OK) return rc;
  return launch_nce_loss(lse_i, diag_i, lse_t, diag_t, b, loss, ST(stream));
}

int32_t ilfdt_nce_lse(const void* x, const void* g, int32_t dtype, int64_t b, int64_t B, int64_t D, const float* scale,
                      int64_t label_offset, float* lse, float* diag, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(b >= 1 && B >= b && D >= 1, ILFDT_E_INVALID_ARG, "nce_lse: need 1 <= b <= B");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B, ILFDT_E_INVALID_ARG, "nce_lse: labels out of range");
  ILFDT_REQUIRE(x && g && lse && diag && scale, ILFDT_E_INVALID_ARG, "nce_lse: null pointer");
  if (dtype == ILFDT_F32)
    return nce_lse_f32(reinterpret_cast<const float*>(x), reinterpret_cast<const float*>(g), b, B, D, scale, label_offset,
                       lse, diag, ws, ws_bytes, ST(stream));
  return launch_nce_lse_bf16(x, g, b, B, D, scale, label_offset, lse, diag, ws, ws_bytes, ST(stream));
}

int32_t ilfdt_nce_loss(const float* lse_i, const float* diag_i, const float* lse_t, const float* diag_t, int64_t b,
                       float* loss, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 1 && lse_i && diag_i && lse_t && diag_t && loss, ILFDT_E_INVALID_ARG, "nce_loss: bad arguments");
  return launch_nce_loss(lse_i, diag_i, lse_t, diag_t, b, loss, ST(stream));
}

int32_t ilfdt_nce_bwd(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt, int32_t dtype, int64_t b,
                      int64_t B, int64_t D, const float* scale, int64_t label_offset, const float* lse_i,
                      const float* lse_t, const float* grad_out, float* dx_img, float* dx_txt, float* dg_img, float* dg_txt,
                      float* dscale, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(b >= 1 && B >= b && D >= 1, ILFDT_E_INVALID_ARG, "nce_bwd: need 1 <= b <= B");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B, ILFDT_E_INVALID_ARG, "nce_bwd: labels out of range");
  ILFDT_REQUIRE(x_img && x_txt && g_img && g_txt && lse_i && lse_t && grad_out && dx_img && dx_txt && dg_img &&
                    dg_txt && dscale && scale,
                ILFDT_E_INVALID_ARG, "nce_bwd: null pointer");
  if (dtype == ILFDT_F32)
    return nce_bwd_f32(reinterpret_cast<const float*>(x_img), reinterpret_cast<const float*>(x_txt),
                       reinterpret_cast<const float*>(g_img), reinterpret_cast<const float*>(g_txt), b, B, D, scale,
                       label_offset, lse_i, lse_t, grad_out, dx_img, dx_txt, dg_img, dg_txt, dscale, ws, ws_bytes,
                       ST(stream));
  return launch_nce_bwd_bf16(x_img, x_txt, g_img, g_txt, b, B, D, scale, label_offset, lse_i, lse_t, grad_out, dx_img,
                             dx_txt, dg_img, dg_txt, dscale, ws, ws_bytes, ST(stream));
}

int32_t ilfdt_nce_bwd_gathered(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt, int32_t dtype,
                               int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset,
                               const float* lse_i_all, const float* lse_t_all, const float* grad_out_all, float* dx_img,
                               float* dx_txt, float* dscale, void* ws, int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(b >= 1 && B >= b && D >= 1 && B % b == 0, ILFDT_E_INVALID_ARG,
                "nce_bwd_gathered: need 1 <= b, B a multiple of b");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B && label_offset % b == 0, ILFDT_E_INVALID_ARG,
                "nce_bwd_gathered: label_offset must be rank*b");
  ILFDT_REQUIRE(x_img && x_txt && g_img && g_txt && lse_i_all && lse_t_all && grad_out_all && dx_img && dx_txt &&
                    dscale && scale,
                ILFDT_E_INVALID_ARG, "nce_bwd_gathered: null pointer");
  if (dtype == ILFDT_F32)
    return nce_bwd_gathered_f32(reinterpret_cast<const float*>(x_img), reinterpret_cast<const float*>(x_txt),
                                reinterpret_cast<const float*>(g_img), reinterpret_cast<const float*>(g_txt), b, B, D,
                                scale, label_offset, lse_i_all, lse_t_all, grad_out_all, dx_img, dx_txt, dscale, ws,
                                ws_bytes, ST(stream));
  return launch_nce_bwd_gathered_bf16(x_img, x_txt, g_img, g_txt, b, B, D, scale, label_offset, lse_i_all, lse_t_all,
                                      grad_out_all, dx_img, dx_txt, dscale, ST(stream));
}

int32_t ilfdt_ce_rows_fwd(const float* logits, int64_t b, int64_t B, int64_t label_offset, float* lse, float* diag,
                          void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 0 && B >= 1 && (b == 0 || (logits && lse && diag)), ILFDT_E_INVALID_ARG, "ce_rows_fwd: bad arguments");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B, ILFDT_E_INVALID_ARG, "ce_rows_fwd: labels out of range");
  return launch_lse_rows(logits, B, b, B, label_offset, lse, diag, ST(stream));
}

int32_t ilfdt_ce_rows_bwd(const float* logits, int64_t b, int64_t B, int64_t label_offset, const float* lse,
                          const float* grad_out, float coef, float* dlogits, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 0 && B >= 1 && (b == 0 || (logits && lse && grad_out && dlogits)), ILFDT_E_INVALID_ARG,
                "ce_rows_bwd: bad arguments");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B, ILFDT_E_INVALID_ARG, "ce_rows_bwd: labels out of range");
  return launch_softmax_grad_rows(logits, dlogits, B, b, B, label_offset, lse, grad_out, coef, nullptr, nullptr,
                                  ST(stream));
}

int32_t ilfdt_nce_topk_hits(const void* x, const void* g, int32_t dtype, int64_t b, int64_t B, int64_t D,
                            const float* scale, int64_t label_offset, const float* diag, int32_t* counts, void* ws,
                            int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dtype);
  ILFDT_REQUIRE(b >= 1 && B >= b && D >= 1, ILFDT_E_INVALID_ARG, "nce_topk_hits: need 1 <= b <= B");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B, ILFDT_E_INVALID_ARG, "nce_topk_hits: labels out of range");
  ILFDT_REQUIRE(x && g && scale && diag && counts, ILFDT_E_INVALID_ARG, "nce_topk_hits: null pointer");
  if (dtype == ILFDT_BF16)
    return launch_nce_topk_bf16(x, g, b, B, D, scale, label_offset, diag, counts, ws, ws_bytes, ST(stream));
  // F32: stream the logits through `ws` in row chunks (same chunking as ilfdt_nce_fwd) and count per chunk
  const int64_t R = nce_f32_chunk_rows(b, B);
  ILFDT_REQUIRE(ws != nullptr && ws_bytes >= R * B * 4, ILFDT_E_INVALID_ARG, "nce_topk_hits(f32): workspace too small");
  float* L = reinterpret_cast<float*>(ws);
  ILFDT_CUDA_OK(cudaMemsetAsync(counts, 0, 8, ST(stream)));
  for (int64_t r0 = 0; r0 < b; r0 += R) {
    const int64_t rows = std::min(R, b - r0);
    int32_t rc = launch_gemm_f32(reinterpret_cast<const float*>(x) + r0 * D, D, 1, reinterpret_cast<const float*>(g), D, 1,
                                 L, B, 1, rows, B, D, 1.f, scale, 0, ST(stream));
    if (rc != ILFDT_OK) return rc;
    if ((rc = launch_topk_hits_acc(L, rows, B, label_offset + r0, counts, ST(stream))) != ILFDT_OK) return rc;
  }
  return ILFDT_OK;
}

int32_t ilfdt_topk_hits(const float* logits, int64_t b, int64_t B, int64_t label_offset, int32_t* counts,
                        void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(b >= 0 && B >= b && counts && (b == 0 || logits), ILFDT_E_INVALID_ARG, "topk_hits: bad arguments");
  ILFDT_REQUIRE(label_offset >= 0 && label_offset + b <= B, ILFDT_E_INVALID_ARG, "topk_hits: labels out of range");
  return launch_topk_hits(logits, b, B, label_offset, counts, ST(stream));
}

int32_t ilfdt_randn_fill(float* dst, int64_t n, uint64_t seed, uint32_t generation, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(n >= 0 && (n == 0 || dst), ILFDT_E_INVALID_ARG, "randn_fill: bad arguments");
  return launch_randn_fill(dst, n, seed, generation, ST(stream));
}

int32_t ilfdt_copy(void* dst, const void* src, int64_t bytes, void* stream) {
  ARCH_GUARD();
  ILFDT_REQUIRE(bytes >= 0 && (bytes == 0 || (dst && src)), ILFDT_E_INVALID_ARG, "copy: bad arguments");
  if (bytes == 0) return ILFDT_OK;
  ILFDT_CUDA_OK(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, ST(stream)));
  return ILFDT_OK;
}

/* ---- q_map producer rows (clip_fdt.py:86-92) ---------------------------------------------------- */
int64_t ilfdt_ln_bwd_ws_bytes(int64_t rows, int64_t C) { return rows > 0 ? ln_bwd_ws_floats(rows, C) * 4 : 0; }
int64_t ilfdt_colsum_ws_bytes(int64_t rows, int64_t C) { return rows > 0 ? colsum_ws_floats(rows, C) * 4 : 0; }

int32_t ilfdt_ln_fwd(const void* src, int32_t src_dtype, const float* bias, const float* gamma, const float* beta,
                     int64_t rows, int64_t C, float eps, int32_t gelu, void* y_bf16, void* pre_bf16, float* mean,
                     float* rstd, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(src_dtype);
  ILFDT_REQUIRE(rows >= 0 && C >= 1 && eps >= 0.f, ILFDT_E_INVALID_ARG, "ln_fwd: bad shape");
  ILFDT_REQUIRE(rows == 0 || (src && gamma && beta && y_bf16 && mean && rstd), ILFDT_E_INVALID_ARG, "ln_fwd: null pointer");
  return launch_ln_fwd(src, src_dtype, bias, gamma, beta, rows, C, eps, gelu, y_bf16, pre_bf16, mean, rstd, ST(stream));
}

int32_t ilfdt_ln_bwd(const void* dy, int32_t dy_dtype, const void* src, int32_t src_dtype, const float* gamma,
                     const float* mean, const float* rstd, int64_t rows, int64_t C, int32_t gelu, void* dsrc,
                     int32_t dsrc_dtype, float* dgamma, float* dbeta, float* dbias, void* ws, int64_t ws_bytes,
                     void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dy_dtype);
  DTYPE_GUARD(src_dtype);
  DTYPE_GUARD(dsrc_dtype);
  ILFDT_REQUIRE(rows >= 1 && C >= 1, ILFDT_E_INVALID_ARG, "ln_bwd: bad shape");
  ILFDT_REQUIRE(dy && src && gamma && mean && rstd && dsrc && dgamma && dbeta && ws, ILFDT_E_INVALID_ARG,
                "ln_bwd: null pointer");
  ILFDT_REQUIRE(ws_bytes >= ilfdt_ln_bwd_ws_bytes(rows, C), ILFDT_E_INVALID_ARG, "ln_bwd: workspace too small");
  return launch_ln_bwd(dy, dy_dtype, src, src_dtype, gamma, mean, rstd, rows, C, gelu, dsrc, dsrc_dtype, dgamma, dbeta,
                       dbias, (float*)ws, ST(stream));
}

int32_t ilfdt_bias_cast(const float* src, const float* bias, int64_t rows, int64_t C, void* dst, int32_t dst_dtype,
                        void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(dst_dtype);
  ILFDT_REQUIRE(rows >= 0 && C >= 1 && (rows == 0 || (src && bias && dst)), ILFDT_E_INVALID_ARG, "bias_cast: bad arguments");
  return launch_bias_cast(src, bias, rows, C, dst, dst_dtype, ST(stream));
}

int32_t ilfdt_colsum(const void* src, int32_t src_dtype, int64_t rows, int64_t C, float* out, void* ws,
                     int64_t ws_bytes, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(src_dtype);
  ILFDT_REQUIRE(rows >= 1 && C >= 1 && src && out && ws, ILFDT_E_INVALID_ARG, "colsum: bad arguments");
  ILFDT_REQUIRE(ws_bytes >= ilfdt_colsum_ws_bytes(rows, C), ILFDT_E_INVALID_ARG, "colsum: workspace too small");
  return launch_colsum(src, src_dtype, rows, C, out, (float*)ws, ST(stream));
}

int32_t ilfdt_colsum_static(const void* src, int32_t src_dtype, int64_t rows, int64_t C, float* out, void* stream) {
  ARCH_GUARD();
  DTYPE_GUARD(src_dtype);
  ILFDT_REQUIRE(rows >= 1 && C >= 1 && src && out, ILFDT_E_INVALID_ARG, "colsum_static: bad arguments");
  return launch_colsum_static(src, src_dtype, rows, C, out, ST(stream));
}

}  // extern "C"
