/*
 * ilfdt.h -- C ABI of libilfdt.so: the B200 (sm_100a) kernels behind the FDT-grounding +
 * ClipInfoCELoss hot path of iterated-learning FDT-CLIP training.
 *
 * This is the drop-in boundary. The reference (hellomuffin/iterated-learning-for-vlm)
 * is pure PyTorch and has no FFI of its own; each entry point below names the reference lines
 * (paths relative to the reference root) whose arithmetic it replaces. The Python host side
 * (iterated-learning-for-vlm_b200/ilfdt) binds these with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - every function returns an int32 status: 0 = ok, negative = ILFDT_E_*; the message for the
 *     last failure on the calling thread is ilfdt_last_error().
 *   - the caller owns ALL memory. Pointers are device pointers unless named host_*. Tensors are
 *     dense row-major with the shapes given. Nothing here allocates user-visible buffers,
 *     synchronises the host, or keeps global mutable state (one small per-thread error string and the
 *     ilfdt_tune_* knobs).
 *   - `stream` is a cudaStream_t passed as void*. Work is enqueued on it and the call returns.
 *   - dtype arguments are ILFDT_F32 or ILFDT_BF16. F32 selects the fp32-exact SIMT kernels
 *     (parity mode, <=1e-5 rel); BF16 selects the tcgen05/TMEM/TMA tensor-core kernels with fp32
 *     accumulation (throughput mode, <=2e-2 rel).
 *   - `ws`/`ws_bytes`: caller-provided scratch; query the size with the matching *_ws_bytes().
 *   - there is no CPU fallback: on a device that is not sm_100 every compute call returns
 *     ILFDT_E_ARCH.
 */
#ifndef ILFDT_H_
#define ILFDT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ILFDT_ABI_VERSION 4

enum {
  ILFDT_OK = 0,
  ILFDT_E_INVALID_ARG = -1,
  ILFDT_E_UNSUPPORTED_SHAPE = -2,
  ILFDT_E_CUDA = -3,
  ILFDT_E_NCCL = -4,
  ILFDT_E_ARCH = -5
};

enum { ILFDT_F32 = 0, ILFDT_BF16 = 1 };
enum { ILFDT_POOL_MAX = 0, ILFDT_POOL_SUM = 1, ILFDT_POOL_MEAN = 2 };

/* ---- library ---------------------------------------------------------------------------- */
int32_t ilfdt_abi_version(void);
const char* ilfdt_last_error(void);
/* 0 if `device` is an sm_100 part and the kernels can run there, ILFDT_E_ARCH otherwise. */
int32_t ilfdt_device_check(int32_t device);

/* Tuning knobs: process-wide, explicit (never read from the environment), none changes a result beyond fp32
 * summation order. They exist so that tests and tools can A/B the routes in one process. */
enum {
  ILFDT_TUNE_DSD_SPLIT_ROWS = 0, /* ilfdt_ground_bwd: samples per range of the codebook-gradient sweep (default 1024);
                                    0 = one sweep, order-deterministic, no float reductions */
  ILFDT_TUNE_DSD_TILE = 1,       /* ilfdt_ground_bwd: 32 (default) or 128 samples per cooperative tile */
  ILFDT_TUNE_DSD_STREAM = 2,     /* ilfdt_ground_bwd: 1 (default) = evict-first loads of the once-read coef / arg */
  ILFDT_TUNE_GROUND_2CTA = 3,    /* ilfdt_ground_fwd*: 1 (default) = cta_group::2 kernel, 0 = single-CTA kernel */
  ILFDT_TUNE_GEMM_WIDE2 = 4,     /* ilfdt_gemm_nt: 1 (default) = CTA-pair kernel for short-operand / long-reduction shapes */
  ILFDT_TUNE_NCE_2CTA = 5,       /* ilfdt_nce_bwd*: 1 (default) = cta_group::2 kernel */
  ILFDT_TUNE_SPARSE_BWD_NNZ = 6, /* ilfdt_ground_bwd*: mean support (codes per row) up to which the sparse-regime kernel
                                    serves the launch (default 128; 0 = always the dense sweeps) */
  ILFDT_TUNE_DSD_PIPE = 7,       /* ilfdt_ground_bwd* (bf16 rows): 1 (default) = pipelined gather stream (dsd v4: 4 codes x 64
                                    samples per tile, groups of 2 rows in a two-deep register pipeline, packed FFMA2); 2 =
                                    groups of 4 rows; 3 = scalar FMAs; 0 = the v3 kernel. Same summation order: identical results */
  ILFDT_TUNE_DQ_PIPE = 8,        /* ilfdt_ground_bwd* (bf16 rows): per-sample kernel's gather loop: 0 = batches of 4 rows, 1 =
                                    pipelined groups of 2 (3 CTAs / SM), 2 = pipelined groups of 4 (2 CTAs / SM). Same order */
  ILFDT_TUNE_F32_X3 = 9,         /* F32 InfoNCE entry points: 1 (default) = the logits contractions run on the tensor cores as a
                                    3-term bf16 split (ilfdt_split3_bf16; fp32-grade: 7e-7 of fp64), 0 = SIMT FFMA kernels */
  ILFDT_TUNE_DQ_DEEP = 10,       /* ilfdt_ground_bwd* (bf16 rows): 1 (default) = launches with at most 2 x SM count samples keep 8
                                    codebook rows in flight per warp instead of 4 (2 CTAs / SM); 0 = always 4. Same order */
  ILFDT_TUNE_TEXT_MERGED_MAX_B = 11, /* ilfdt_ground_fwd_text: batches of at most this many samples (default 1024) run the tiles
                                    of ALL length buckets in one launch, larger ones one launch per bucket; 0 = never. Same bits */
  ILFDT_TUNE_COUNT_ = 12
};
int32_t ilfdt_tune_set(int32_t key, int64_t value);
int64_t ilfdt_tune_get(int32_t key);

/* ---- layout helpers (feed the bf16 tensor-core kernels) ----------------------------------- */
/* dst_bf16[i] = bf16(src[i]) */
int32_t ilfdt_cast_bf16(const float* src, void* dst_bf16, int64_t n, void* stream);
/* dst_bf16[r, :] = rowmask[r] == 0 ? bf16(src[r, :]) : 0.  Folding the pad mask into the operand makes the
 * tensor-core score of a pad token exactly the 0 that clip_fdt.py:125-130 produces by multiplying. */
int32_t ilfdt_cast_bf16_rows(const float* src, const float* rowmask, int64_t rows, int64_t cols, void* dst_bf16,
                             void* stream);
/* The same for a q that already is bf16 (a producer running under bf16 autocast, or host buffers kept in
 * bf16): dst_bf16[r, :] = rowmask[r] == 0 ? src_bf16[r, :] : 0. */
int32_t ilfdt_mask_rows_bf16(const void* src_bf16, const float* rowmask, int64_t rows, int64_t cols,
                             void* dst_bf16, void* stream);
/* fp32-grade contractions on the bf16 tensor cores: src [rows, cols] f32 -> dst_bf16 [rows, 6*cols], six
 * concatenated copies of the row's three bf16 terms hi + mid + lo (= src to 2^-24) in the order [2 1 0 1 0 0] (side 0, the
 * A operand) or [0 1 2 0 1 0] (side 1, the B operand; smallest partial products first, the leading one last): a single bf16 contraction over 6*cols then accumulates, in fp32,
 * every partial product down to 2^-16 of the leading one. rowmask (optional): rows with a non-zero entry are zeroed. */
int32_t ilfdt_split3_bf16(const float* src, const float* rowmask, int64_t rows, int64_t cols, int32_t side, void* dst_bf16,
                          void* stream);
/* src [rows, cols] f32 or bf16 -> dst_bf16 [cols, rows] */
int32_t ilfdt_transpose_bf16(const void* src, int32_t src_dtype, int64_t rows, int64_t cols,
                             void* dst_bf16, void* stream);

/* ---- a1: Query_model score block, prototype/model/clip_fdt.py:113-145 --------------------- */
/* q [b, n_tok, D], sd [K, D] (dtype = F32 or BF16, both the same), mask [b, n_tok] f32 with 0 for a
 * real token and -inf for a pad (text_transformer.py:182-194) or NULL.
 * pooled[b,K] = pool_t( ((q.sd^T) / sqrt_d) * (mask==0) / temperature )   (clip_fdt.py:118-145),
 * arg[b,K] = the token index the max came from (int16; pool MAX only, may be NULL otherwise).
 * The [b,n_tok,K] scores never exist in memory. */
int64_t ilfdt_ground_fwd_ws_bytes(int32_t dtype, int64_t b, int64_t n_tok, int64_t K, int64_t D);
int32_t ilfdt_ground_fwd(const void* q, const void* sd, int32_t dtype, const float* mask,
                         int64_t b, int64_t n_tok, int64_t K, int64_t D,
                         float sqrt_d, float temperature, int32_t pool_type,
                         float* pooled, int16_t* arg, void* ws, int64_t ws_bytes, void* stream);

/* ---- a2: Sparsemax, prototype/model/sparsemax.py:22-71 ------------------------------------ */
/* a1 for padded text in BF16 (clip_fdt.py:113-145 with a pad mask). Every pad token scores exactly 0
 * (clip_fdt.py:125-130), so only the rows [0, len_s] of sample s -- its valid tokens plus ONE pad -- can affect the
 * max over tokens. ilfdt_text_compact plans on the DEVICE (no host sync): samples are bucketed by
 * R = min(16*ceil((len+1)/16), n_tok), q (f32 or bf16) is cast into a compacted bf16 operand in bucket order with the
 * pad rows zeroed, and `plan` (ilfdt_text_plan_ints(b) int32) receives per bucket {samples, first row, first slot,
 * rows per sample} followed by slot_of[b], row_base[b], sample_of[b]. ilfdt_ground_fwd_text runs the tensor-core
 * kernel once per bucket; pooled / arg [b,K] are bit-identical to ilfdt_ground_fwd on the full-width operand.
 * groups_ws: b bytes of scratch; q_compact_bf16: room for b*n_tok rows. The backward takes plan's row_base
 * (ilfdt_ground_bwd q_row_base). Supported when every bucket size has a static epilogue (n_tok 77, 49, <= 64 ...):
 * ilfdt_ground_fwd_text_supported. */
int64_t ilfdt_text_plan_ints(int64_t b);
int32_t ilfdt_ground_fwd_text_supported(int64_t n_tok);
int32_t ilfdt_text_compact(const void* q, int32_t q_dtype, const float* mask, int64_t b, int64_t n_tok, int64_t D,
                           void* groups_ws, int32_t* plan, void* q_compact_bf16, void* stream);
/* The same compaction for the fp32-grade score contraction (ilfdt_split3_bf16, clip_fdt.py:118 in fp32 arithmetic): row
 * (row_base[s] + t) of q_compact_x3_bf16 (room for b*n_tok rows of 6*D bf16) holds the side-1 split of the fp32 source row,
 * zeros for a pad. ilfdt_ground_fwd_text then takes it with the side-0 split of the codebook, D = 6 * the model width and
 * sqrt_d = sqrt(model width). */
int32_t ilfdt_text_compact_x3(const float* q, const float* mask, int64_t b, int64_t n_tok, int64_t D, void* groups_ws,
                              int32_t* plan, void* q_compact_x3_bf16, void* stream);
int32_t ilfdt_ground_fwd_text(const void* q_compact_bf16, const void* sd_bf16, const int32_t* plan, int64_t b,
                              int64_t n_tok, int64_t K, int64_t D, float sqrt_d, float temperature, float* pooled,
                              int16_t* arg, void* stream);

/* z [rows, K] -> w [rows, K] (simplex projection of each row), tau[rows], nnz[rows];
 * w_bf16 (optional, may be NULL) receives a bf16 copy of w for the pooling GEMM; w itself may be NULL
 * when w_bf16 is given (the caller keeps only the bf16 copy, see ilfdt_ground_bwd_w16). */
int32_t ilfdt_sparsemax_fwd(const float* z, int64_t rows, int64_t K, float* w, void* w_bf16,
                            float* tau, int32_t* nnz, void* stream);
/* dz = nz * (g - mean_nz(g)),  nz = (w > 0): the gradient autograd derives through sort/cumsum. */
int32_t ilfdt_sparsemax_bwd(const float* w, const float* g, int64_t rows, int64_t K, float* dz,
                            void* stream);

/* ---- generic contraction used by a3 (pooling, clip_fdt.py:154) and its backward ------------ */
/* out[n, m] (+)= alpha * sum_k A[m,k] * B[n,k];  A [M,Kr], B [N,Kr] both with the reduction
 * index contiguous; out is f32 [N, M] with leading dimension ldo (elements).
 * BF16: tcgen05 kernel (M % 128 == 0, Kr % 64 == 0 required); F32: SIMT kernel, any shape. */
int32_t ilfdt_gemm_nt(const void* A, const void* B, int32_t dtype, int64_t M, int64_t N, int64_t Kr,
                      float alpha, int32_t accumulate, float* out, int64_t ldo, void* stream);
/* ilfdt_gemm_nt (BF16) for a SHORT A, FEW B rows and a long reduction -- the pooling f[b,D] = w[b,K] . sd[K,D] of
 * clip_fdt.py:154 at a per-GPU batch of a few hundred rows (run.sh: 256), where the unsplit launch has 1..16 tiles for 148
 * SMs: the reduction is cut over the CTA pairs into slabs of `ws`, added in a fixed order (deterministic, no atomics).
 * ws: ilfdt_gemm_nt_splitk_ws_bytes(M, N, Kr); 0 = this shape is not split (the call then equals ilfdt_gemm_nt). */
int64_t ilfdt_gemm_nt_splitk_ws_bytes(int64_t M, int64_t N, int64_t Kr);
int32_t ilfdt_gemm_nt_splitk(const void* A_bf16, const void* B_bf16, int64_t M, int64_t N, int64_t Kr, float alpha,
                             int32_t accumulate, float* out, int64_t ldo, void* ws, int64_t ws_bytes, void* stream);
/* The same product with A given as [Kr, M] row-major -- the reduction index OUTERMOST, i.e. the bf16 codebook sd [K, D] itself
 * for the pooling f[b,D] = w[b,K] . sd[K,D] of clip_fdt.py:154: no transposed copy of the codebook is made (tcgen05 reads the
 * staged 64 x 64 boxes as an MN-major operand). out[n, m] (+)= alpha * sum_k A[k,m] * B[n,k]; B [N,Kr] bf16. CTA-pair kernel
 * only: ilfdt_gemm_nn_supported(M, N, Kr) != 0 (M = 256 or 512, Kr >= 1024, and N >= 1024 or a split reduction), else
 * ILFDT_E_UNSUPPORTED_SHAPE -- transpose and call ilfdt_gemm_nt then. ws: ilfdt_gemm_nt_splitk_ws_bytes(M, N, Kr) bytes
 * (may be NULL when that is 0). Same k order per output element as ilfdt_gemm_nt(_splitk): identical results. */
int32_t ilfdt_gemm_nn_supported(int64_t M, int64_t N, int64_t Kr);
int32_t ilfdt_gemm_nn(const void* A_bf16, const void* B_bf16, int64_t M, int64_t N, int64_t Kr, float alpha,
                      int32_t accumulate, float* out, int64_t ldo, void* ws, int64_t ws_bytes, void* stream);
/* out[N, M] (+)= alpha * B[Kr,N]^T . A[Kr,M] for bf16 operands stored with the REDUCTION index outermost (tcgen05
 * MN-major operands: no transposed copies). dsd[K,D] += w[b,K]^T . df[b,D] of the pooling backward
 * (clip_fdt.py:154): A = df, B = w, M = D, N = K, Kr = b. M, N multiples of 8. */
int32_t ilfdt_gemm_tn(const void* A_bf16, const void* B_bf16, int64_t M, int64_t N, int64_t Kr, float alpha,
                      int32_t accumulate, float* out, int64_t ldo, void* stream);
/* nn.Linear on bf16 operands (q_map, clip_fdt.py:88,91): out_bf16[n, m] = bf16(sum_k W[m,k] * X[n,k] + bias[m]), W [M = out
 * features, Kr], X [N rows, Kr], bias [M] f32 or NULL. tcgen05 kernel with the bias add and the bf16 rounding in its
 * TMEM read-out (the fp32 product never reaches HBM). M % 128 == 0, Kr % 64 == 0. */
int32_t ilfdt_linear_bf16(const void* W_bf16, const void* X_bf16, int64_t M, int64_t N, int64_t Kr, const float* bias,
                          void* out_bf16, int64_t ldo, void* stream);
/* ilfdt_gemm_tn for weight-gradient shapes (short M and N, Kr in the millions: dW of q_map's Linear layers,
 * clip_fdt.py:88,91): the reduction is split over the SMs into slabs of `ws` that a second kernel sums in a fixed
 * order (deterministic, no atomics). ws: ilfdt_gemm_tn_splitk_ws_bytes(M, N, Kr) (0 = not split, plain launch). */
int64_t ilfdt_gemm_tn_splitk_ws_bytes(int64_t M, int64_t N, int64_t Kr);
int32_t ilfdt_gemm_tn_splitk(const void* A_bf16, const void* B_bf16, int64_t M, int64_t N, int64_t Kr, float alpha,
                             int32_t accumulate, float* out, int64_t ldo, void* ws, int64_t ws_bytes, void* stream);
/* fp32 only, arbitrary strides: C(m,n) (+)= alpha * sum_k A(m,k) B(n,k), X(i,j) = X[i*rs + j*cs] */
int32_t ilfdt_gemm_f32_strided(const float* A, int64_t a_rs, int64_t a_cs,
                               const float* B, int64_t b_rs, int64_t b_cs,
                               float* C, int64_t c_rs, int64_t c_cs,
                               int64_t M, int64_t N, int64_t Kr, float alpha, int32_t accumulate,
                               void* stream);

/* ---- a13: backward of the score block through max-pool + sparsemax ------------------------- */
/* Inputs: g [b,K] = dLoss/dw (f32, scratch: OVERWRITTEN), w [b,K], arg [b,K], q, sd, mask as in
 * ilfdt_ground_fwd.  coef[b,k] = nz*(g - mean_nz g) * keep[b,arg] / (sqrt_d*temperature);
 * dq[b,arg[b,k],:] += coef*sd[k,:]  (dq [b,n_tok,D], dq_dtype F32 or BF16 = the dtype q had at the caller,
 * fully written);
 * dsd[k,:] += coef*q[b,arg[b,k],:]   (dsd [K,D] f32, ACCUMULATED into).
 * Only support entries (w>0) are touched: cost scales with nnz, not with n_tok*K.
 * w == NULL: g already holds dLoss/dpooled (softmax / sigmoid attention): no support, no mean to
 * subtract, every non-zero entry is scattered.
 * q_row_base: NULL when q is the full [b,n_tok,D] operand; else row_base[b] of ilfdt_text_compact's plan
 * (q is then the compacted text operand: token t of sample s lives in row q_row_base[s] + t).
 * nnz (optional, NULL = off): nnz[b] from ilfdt_sparsemax_fwd, with regime_ws = one int32 of device scratch. When the
 * MEAN support is at most ILFDT_TUNE_SPARSE_BWD_NNZ codes per row (trained, peaked attention) the launch is served by a
 * per-sample kernel that touches only the support instead of the two dense sweeps over [b,K]; the decision is taken on
 * the device (no host synchronisation, CUDA-graph safe), results agree up to fp32 summation order. */
int32_t ilfdt_ground_bwd(float* g, const float* w, const int16_t* arg, const void* q, const void* sd,
                         int32_t dtype, const float* mask, const int32_t* q_row_base, const int32_t* nnz,
                         int32_t* regime_ws, int64_t b, int64_t n_tok, int64_t K, int64_t D, float sqrt_d,
                         float temperature, void* dq, int32_t dq_dtype, float* dsd, void* stream);
/* Same, with the support taken from the bf16 copy of the weights that ilfdt_sparsemax_fwd writes for the
 * tensor-core pooling GEMM (w_bf16 [b,K], > 0 exactly where the fp32 weights are): half the bytes read, and
 * a caller that discards att_weight (Clip_FDT.forward, clip_fdt.py:233-234) need not have the fp32 copy
 * written at all. Needs D in {256,512,768,1024}, n_tok <= 256, K % 4 == 0 (ILFDT_E_UNSUPPORTED_SHAPE otherwise). */
int32_t ilfdt_ground_bwd_w16(float* g, const void* w_bf16, const int16_t* arg, const void* q, const void* sd,
                             int32_t dtype, const float* mask, const int32_t* q_row_base, const int32_t* nnz,
                             int32_t* regime_ws, int64_t b, int64_t n_tok, int64_t K, int64_t D, float sqrt_d,
                             float temperature, void* dq, int32_t dq_dtype, float* dsd, void* stream);

/* ---- non-default Query_model variants, clip_fdt.py:67-78,140-157 ----------------------------------- */
enum { ILFDT_ATT_SOFTMAX = 1, ILFDT_ATT_SIGMOID = 2 };
/* kind SOFTMAX: w = softmax(z) per row (clip_fdt.py:72). kind SIGMOID: w = sigmoid(z) (:76), rowsum[r] = sum_k w
 * (the denominator of :156-157; may be NULL for softmax). z, w [rows, K] f32. */
int32_t ilfdt_att_fwd(const float* z, int64_t rows, int64_t K, int32_t kind, float* w, float* rowsum, void* stream);
/* SOFTMAX: dz = w * (g - sum_k w g).  SIGMOID: dz = (g + rowbias[r]) * w * (1 - w) (rowbias may be NULL). */
int32_t ilfdt_att_bwd(const float* w, const float* g, const float* rowbias, int64_t rows, int64_t K, int32_t kind,
                      float* dz, void* stream);
/* f = u / s[row]   (att_ft / att_weight.sum, clip_fdt.py:156-157) and its backward:
 * du = df / s, ds[r] = -(df[r,:] . f[r,:]) / s[r]. u, f, df, du [rows, D]; s, ds [rows]. */
int32_t ilfdt_rownorm_fwd(const float* u, const float* s, int64_t rows, int64_t D, float* f, void* stream);
int32_t ilfdt_rownorm_bwd(const float* df, const float* f, const float* s, int64_t rows, int64_t D, float* du,
                          float* ds, void* stream);
/* sum / mean token pooling goes through the token-summed query: out[b,:] = sum_t keep[b,t] q[b,t,:]  (q f32
 * [b,n_tok,D]) and, for the backward, out[b,t,:] = keep[b,t] * scale * v[b,:]. */
int32_t ilfdt_token_sum(const float* q, const float* mask, int64_t b, int64_t n_tok, int64_t D, float* out,
                        void* stream);
int32_t ilfdt_token_bcast(const float* v, const float* mask, int64_t b, int64_t n_tok, int64_t D, float scale,
                          float* out, void* stream);

/* ---- a5: L2 normalise, clip_fdt.py:411-412 / clip.py:133-134 ------------------------------ */
/* x = f / (||f|| + eps); x_bf16 optional copy; norm[rows] = ||f||. */
int32_t ilfdt_l2norm_fwd(const float* f, int64_t rows, int64_t D, float eps, float* x, void* x_bf16,
                         float* norm, void* stream);
/* df = (g - x * (x.g) * ||f||/(||f||+eps)) / (||f|| + eps) */
int32_t ilfdt_l2norm_bwd(const float* x, const float* norm, const float* g, int64_t rows, int64_t D,
                         float eps, float* df, void* stream);

/* ---- a8 + a10: logits and ClipInfoCELoss, clip_fdt.py:421-422, loss_functions/loss.py:40-50 --- */
/* logits[b, B] = scale[0] * x[b,:] . g[B,:]   (materialising route; accuracy()/visualisation).
 * `scale` is a DEVICE pointer to the (already clamped) exp(logit_scale), so that no call on this
 * path ever needs the host to read a parameter back. */
int32_t ilfdt_logits(const void* x, const void* g, int32_t dtype, int64_t b, int64_t B, int64_t D,
                     const float* scale, float* out, void* stream);

/* Fused InfoNCE forward. x_img/x_txt [b,D] local rows, g_img/g_txt [B,D] gathered (rank-major),
 * label of local row r is label_offset + r (loss.py:41-45). Produces per-row log-sum-exp and label
 * logit of logits_per_image = scale*x_img.g_txt^T and logits_per_text = scale*x_txt.g_img^T and
 * loss[0] = (mean(lse_i-diag_i) + mean(lse_t-diag_t))/2.  The [b,B] logits are never stored in
 * BF16 mode; F32 mode streams them through `ws` in row chunks. */
int64_t ilfdt_nce_ws_bytes(int32_t dtype, int64_t b, int64_t B, int64_t D);
int32_t ilfdt_nce_fwd(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt,
                      int32_t dtype, int64_t b, int64_t B, int64_t D, const float* scale,
                      int64_t label_offset, float* lse_i, float* lse_t, float* diag_i, float* diag_t,
                      float* loss, void* ws, int64_t ws_bytes, void* stream);
/* The two halves of ilfdt_nce_fwd as separate calls, for callers that overlap the all-gather of one modality with the
 * direction that does not need it: logits_per_text = x_txt.g_img^T only needs the IMAGE gather, so its log-sum-exp runs
 * while the text gather (clip_fdt.py:418-419) is still in flight. ilfdt_nce_lse: lse[r], diag[r] of the rows
 * scale * x[r,:].g^T with label label_offset + r (x [b,D], g [B,D]; ws: ilfdt_nce_ws_bytes). ilfdt_nce_loss:
 * loss[0] = (mean(lse_i - diag_i) + mean(lse_t - diag_t)) / 2  (loss.py:47-49), fp64 accumulation, deterministic. */
int32_t ilfdt_nce_lse(const void* x, const void* g, int32_t dtype, int64_t b, int64_t B, int64_t D, const float* scale,
                      int64_t label_offset, float* lse, float* diag, void* ws, int64_t ws_bytes, void* stream);
int32_t ilfdt_nce_loss(const float* lse_i, const float* diag_i, const float* lse_t, const float* diag_t, int64_t b,
                       float* loss, void* stream);
/* Fused InfoNCE backward. grad_out: device pointer to dLoss (scalar). P = (softmax - onehot) *
 * grad_out/(2b) is recomputed tile by tile from lse_*. Outputs (all f32, overwritten):
 * dx_img, dx_txt [b,D]; dg_img, dg_txt [B,D] (gradient w.r.t. the gathered copies, to be
 * reduce-scattered by the caller: AllGather.backward, clip_fdt.py:182-188); dscale[0] = dLoss/dscale.
 * For callers whose g_* are independent tensors; training uses ilfdt_nce_bwd_gathered. */
int32_t ilfdt_nce_bwd(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt,
                      int32_t dtype, int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset,
                      const float* lse_i, const float* lse_t, const float* grad_out,
                      float* dx_img, float* dx_txt, float* dg_img, float* dg_txt, float* dscale,
                      void* ws, int64_t ws_bytes, void* stream);
/* Fused InfoNCE backward when g_img/g_txt ARE the all-gather of every rank's x_img/x_txt (rank-major,
 * this rank's rows at label_offset = rank*b; clip_fdt.py:418-419). With the all-gathered per-row
 * log-sum-exps lse_i_all/lse_t_all [B] (ilfdt_nce_fwd outputs of all ranks) and every rank's upstream
 * dLoss grad_out_all [B/b], each rank forms the TOTAL gradient of its own rows directly,
 *   dx_img[i] = scale * sum_j ( P_img[i,j] + P_txt[j,i] ) g_txt[j],   dx_txt likewise,
 * which equals the reference's dX + (all_reduce(dG))[rank] (clip_fdt.py:182-188) with the [B,D]
 * gradient exchange replaced by a [B]-float all-gather and half the contractions. dscale[0] is the
 * rank-local dLoss/dscale of this rank's own logits rows, as in ilfdt_nce_bwd. */
int32_t ilfdt_nce_bwd_gathered(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt,
                               int32_t dtype, int64_t b, int64_t B, int64_t D, const float* scale,
                               int64_t label_offset, const float* lse_i_all, const float* lse_t_all,
                               const float* grad_out_all, float* dx_img, float* dx_txt, float* dscale,
                               void* ws, int64_t ws_bytes, void* stream);

/* Cross-entropy over STORED logits [b,B] (the route taken when a caller hands ClipInfoCELoss real
 * tensors instead of the fused stash): lse[r], diag[r] = logits[r, label_offset + r]. */
int32_t ilfdt_ce_rows_fwd(const float* logits, int64_t b, int64_t B, int64_t label_offset, float* lse,
                          float* diag, void* stream);
/* dlogits[r,c] = (exp(logits[r,c] - lse[r]) - [c == label_offset + r]) * grad_out[0] * coef */
int32_t ilfdt_ce_rows_bwd(const float* logits, int64_t b, int64_t B, int64_t label_offset, const float* lse,
                          const float* grad_out, float coef, float* dlogits, void* stream);

/* ---- N1: accuracy(logits_per_image, labels, topk=(1,5)), prototype/utils/misc.py:464-477 ------ */
/* counts[0] = #rows whose label logit is the row max, counts[1] = #rows where fewer than 5 entries
 * exceed it.  logits f32 [b,B]. */
int32_t ilfdt_topk_hits(const float* logits, int64_t b, int64_t B, int64_t label_offset,
                        int32_t* counts, void* stream);
/* The same counts WITHOUT the [b,B] matrix: logits = scale[0] * x[b,:] . g[B,:]^T is recomputed tile by
 * tile on the tensor cores and each entry compared with the row's label logit `diag` (ilfdt_nce_fwd's diag_i for
 * logits_per_image) -- the comparison accuracy() makes at train_solver.py:421, entry for entry. counts[2] int32;
 * ws: ilfdt_nce_ws_bytes(dtype, b, B, D) bytes (BF16 uses b*4 of it; F32 streams row chunks of logits through it). */
int32_t ilfdt_nce_topk_hits(const void* x, const void* g, int32_t dtype, int64_t b, int64_t B, int64_t D,
                            const float* scale, int64_t label_offset, const float* diag, int32_t* counts, void* ws,
                            int64_t ws_bytes, void* stream);

/* ---- q_map producer (clip_fdt.py:86-92: LayerNorm -> Linear -> GELU -> LayerNorm -> Linear) ------- */
/* The two Linear layers are ilfdt_gemm_nt / ilfdt_gemm_tn calls on bf16 operands; these are the rows between them.
 * Widths C in {256, 512, 768, 1024}. Column sums are two-stage and deterministic (no float atomics).
 *
 * y_bf16[r,:] = LayerNorm(act(src[r,:] + bias)) * gamma + beta with the biased variance and eps of nn.LayerNorm
 * (clip_fdt.py:87,90); mean / rstd [rows] are saved for the backward. gelu = 0: act = identity, bias ignored. gelu = 1
 * (nn.GELU(), exact erf, clip_fdt.py:89): either src is the first Linear's pre-activation itself (bf16 from
 * ilfdt_linear_bf16; bias and pre_bf16 NULL), or src is the fp32 product without its bias and pre_bf16 [rows,C]
 * receives src + bias, the pre-activation that the backward differentiates. src f32 or bf16 in both forms. */
int32_t ilfdt_ln_fwd(const void* src, int32_t src_dtype, const float* bias, const float* gamma, const float* beta,
                     int64_t rows, int64_t C, float eps, int32_t gelu, void* y_bf16, void* pre_bf16, float* mean,
                     float* rstd, void* stream);
/* Backward of the above for dy [rows,C] (f32, or bf16 as ilfdt_linear_bf16 writes it). src = the forward's src
 * (gelu = 0) or the bf16 pre-activation (gelu = 1). dsrc: f32 (gelu = 0) or bf16 (gelu = 1, the operand of the next
 * contraction); dgamma, dbeta [C]; dbias [C] (gelu = 1, may be NULL) = column sum of dsrc.
 * ws: ilfdt_ln_bwd_ws_bytes(rows, C). */
int64_t ilfdt_ln_bwd_ws_bytes(int64_t rows, int64_t C);
int32_t ilfdt_ln_bwd(const void* dy, int32_t dy_dtype, const void* src, int32_t src_dtype, const float* gamma,
                     const float* mean, const float* rstd, int64_t rows, int64_t C, int32_t gelu, void* dsrc,
                     int32_t dsrc_dtype, float* dgamma, float* dbeta, float* dbias, void* ws, int64_t ws_bytes,
                     void* stream);
/* dst[r,:] = src[r,:] + bias  (the second Linear's bias, clip_fdt.py:91), written as bf16 (the score kernel's
 * operand) or f32. */
int32_t ilfdt_bias_cast(const float* src, const float* bias, int64_t rows, int64_t C, void* dst, int32_t dst_dtype,
                        void* stream);
/* out[c] = sum_r src[r,c]  (bias gradient of a Linear); ws: ilfdt_colsum_ws_bytes(rows, C). */
int64_t ilfdt_colsum_ws_bytes(int64_t rows, int64_t C);
int32_t ilfdt_colsum(const void* src, int32_t src_dtype, int64_t rows, int64_t C, float* out, void* ws,
                     int64_t ws_bytes, void* stream);
/* ilfdt_colsum for callers that cannot allocate a workspace (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same kernels, same grid and therefore the same bits as ilfdt_colsum; the per-CTA partial sums go to a file-scope
 * `static __device__` array of ILFDT_COLSUM_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until
 * the process exits. The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine); the
 * caller passes src and out as usual, and `stream` must belong to the CURRENT device (the array's address is looked
 * up once per device, on the first call with that device current -- make that first call outside stream capture).
 * Limit: (number of CTAs, at most 4 per SM) * C * 4 bytes <= ILFDT_COLSUM_STATIC_SCRATCH_BYTES -- on a 148-SM B200
 * every rows x C with C <= 884 fits (the q_map widths 512 and 768 always do); C up to 1024 fits while
 * rows <= 512 * 8. Beyond it: ILFDT_E_INVALID_ARG with a message naming the sizes -- use ilfdt_colsum.
 * The array is SHARED by all calls on a device and keeps the last call's contents (nothing in it needs to start at
 * zero: each call overwrites every entry it later reads). Calls that overlap on the GPU -- different streams, or
 * different host threads without stream ordering between them -- RACE on it; serialise them on one stream. */
#define ILFDT_COLSUM_STATIC_SCRATCH_BYTES (2 * 1024 * 1024)
int32_t ilfdt_colsum_static(const void* src, int32_t src_dtype, int64_t rows, int64_t C, float* out, void* stream);

/* ---- a12: codebook reset / store / keep, clip_fdt.py:281-283, train_solver.py:211-215,549-553 -- */
/* dst[i] ~ N(0,1): Philox4x32-10 keyed by seed, counter (i/4, generation), Box-Muller. Same values
 * on every rank for the same (seed, generation). */
int32_t ilfdt_randn_fill(float* dst, int64_t n, uint64_t seed, uint32_t generation, void* stream);
int32_t ilfdt_copy(void* dst, const void* src, int64_t bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ILFDT_H_ */
