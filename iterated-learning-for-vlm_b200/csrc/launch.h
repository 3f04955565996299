// Internal launch entry points (host side). Each returns an ILFDT_* status.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ilfdt {

// kernels_simt.cu
int32_t launch_gemm_f32(const float* A, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_rs, int64_t b_cs,
                        float* C, int64_t c_rs, int64_t c_cs, int64_t M, int64_t N, int64_t Kr, float alpha,
                        const float* alpha_dev, int accumulate, cudaStream_t st);
int32_t launch_ground_fwd_f32(const float* q, const float* sd, const float* mask, int64_t b, int64_t n_tok,
                              int64_t K, int64_t D, float sqrt_d, float temperature, int pool_type, float* pooled,
                              int16_t* arg, void* ws, int64_t ws_bytes, cudaStream_t st);
int32_t launch_l2norm_fwd(const float* f, int64_t rows, int64_t D, float eps, float* x, void* xb, float* norm,
                          cudaStream_t st);
int32_t launch_l2norm_bwd(const float* x, const float* norm, const float* g, int64_t rows, int64_t D, float eps,
                          float* df, cudaStream_t st);
int32_t launch_lse_rows(const float* L, int64_t ld, int64_t rows, int64_t cols, int64_t label0, float* lse,
                        float* diag, cudaStream_t st);
int32_t launch_softmax_grad_rows(const float* L, float* out, int64_t ld, int64_t rows, int64_t cols, int64_t label0, const float* lse,
                                 const float* go, float coef, const float* scale_dev, double* row_acc, cudaStream_t st);
int32_t launch_sum_rows(const double* acc, int64_t n, float* out, cudaStream_t st);
int32_t launch_softmax_grad_sym_rows(const float* L, float* out, int64_t ld, int64_t rows, int64_t cols, int64_t label0,
                                     const float* lse_lane, const float* lse_stream, const float* go_lane,
                                     const float* go_stream, int64_t rows_per_go, float coef, const float* scale_dev,
                                     double* row_acc, cudaStream_t st);
int32_t launch_nce_loss(const float* lse_i, const float* diag_i, const float* lse_t, const float* diag_t, int64_t b,
                        float* loss, cudaStream_t st);
int32_t launch_topk_hits(const float* L, int64_t b, int64_t B, int64_t label0, int32_t* counts, cudaStream_t st);
int32_t launch_cast_bf16(const float* src, void* dst, int64_t n, cudaStream_t st);
int32_t launch_cast_bf16_rows(const float* src, const float* rowmask, int64_t rows, int64_t cols, void* dst,
                              cudaStream_t st);
int32_t launch_split3_bf16(const float* src, const float* rowmask, int64_t rows, int64_t cols, int side, void* dst,
                           cudaStream_t st);
int32_t launch_mask_rows_bf16(const void* src, const float* rowmask, int64_t rows, int64_t cols, void* dst,
                              cudaStream_t st);
int32_t launch_transpose_bf16(const void* src, int src_dtype, int64_t rows, int64_t cols, void* dst, cudaStream_t st);
int32_t launch_randn_fill(float* dst, int64_t n, uint64_t seed, uint32_t gen, cudaStream_t st);

// sparsemax.cu
int32_t launch_sparsemax_fwd(const float* z, int64_t rows, int64_t K, float* w, void* wb, float* tau, int32_t* nnz,
                             cudaStream_t st);
int32_t launch_sparsemax_bwd(const float* w, const float* g, int64_t rows, int64_t K, float* dz, cudaStream_t st);

// ground_bwd.cu
int32_t launch_ground_bwd(float* g, const void* w, int w_dtype, const int16_t* arg, const void* q, const void* sd, int dtype,
                          const float* mask, const int* row_base, const int* nnz, int* regime_ws, int64_t b, int64_t n_tok,
                          int64_t K, int64_t D, float sqrt_d, float temperature, void* dq, int dq_dtype, float* dsd,
                          cudaStream_t st);

// gemm_umma.cu (tcgen05)
int32_t launch_gemm_nt_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha,
                            const float* alpha_dev, int accumulate, float* out, int64_t ldo, cudaStream_t st);
int32_t launch_gemm_tn_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha, int accumulate,
                            float* out, int64_t ldo, cudaStream_t st);
int32_t launch_linear_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, const float* bias,
                           void* out_bf16, int64_t ldo, cudaStream_t st);
int64_t gemm_tn_splitk_ws_floats(int64_t M, int64_t N, int64_t Kr);
int32_t launch_gemm_tn_splitk_bf16(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha,
                                   int accumulate, float* out, int64_t ldo, float* ws, cudaStream_t st);
int32_t launch_ground_fwd_bf16(const void* q, const void* sd, const float* mask, int64_t b, int64_t n_tok, int64_t K,
                               int64_t D, float sqrt_d, float temperature, float* pooled, int16_t* arg,
                               cudaStream_t st);
int64_t nce_fwd_bf16_ws_bytes(int64_t b, int64_t B);
bool ground_fwd_text_supported(int64_t n_tok);
int32_t launch_ground_fwd_text_bf16(const void* q_compact, const void* sd, const int* plan, int64_t b, int64_t n_tok,
                                    int64_t K, int64_t D, float sqrt_d, float temperature, float* pooled, int16_t* arg,
                                    cudaStream_t st);
// ground_umma2.cu (tcgen05 cta_group::2)
bool ground2_enabled();
bool ground2_supported(int64_t n_tok, int64_t D);
int32_t launch_ground2(const void* q, const void* sd, int64_t q_rows, int64_t n_samples, int64_t n_tok, int64_t K, int64_t D,
                       float pool_scale, float* pooled, int16_t* arg, const int* dyn, const int* sample_map,
                       int64_t est_samples, cudaStream_t st);
int32_t launch_nce_lse2(const void* x, const void* g, int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset,
                        float* part_m, float* part_s, float* diag, int* nchunk_out, cudaStream_t st);
bool ground2_text_merged_supported(int64_t n_tok, int64_t D);
int32_t launch_ground2_text_merged(const void* q, const void* sd, int64_t q_rows, int64_t b, int64_t n_tok, int64_t K, int64_t D,
                                   float pool_scale, float* pooled, int16_t* arg, const int* plan, const int* sample_map,
                                   cudaStream_t st);
// text_buckets.cu
int64_t text_plan_ints(int64_t b);
int32_t launch_text_compact(const void* q, int q_dtype, const float* mask, int64_t b, int64_t n_tok, int64_t D,
                            uint8_t* groups_ws, int* plan, void* q_out, cudaStream_t st, int x3 = 0);
int32_t launch_nce_lse_bf16(const void* x, const void* g, int64_t b, int64_t B, int64_t D, const float* scale,
                            int64_t label_offset, float* lse, float* diag, void* ws, int64_t ws_bytes, cudaStream_t st);
int32_t launch_nce_fwd_bf16(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt, int64_t b,
                            int64_t B, int64_t D, const float* scale, int64_t label_offset, float* lse_i, float* lse_t,
                            float* diag_i, float* diag_t, void* ws, int64_t ws_bytes, cudaStream_t st);

// rowops.cu
int32_t launch_att_fwd(const float* z, int64_t rows, int64_t K, int kind, float* w, float* rowsum, cudaStream_t st);
int32_t launch_att_bwd(const float* w, const float* g, const float* rowbias, int64_t rows, int64_t K, int kind,
                       float* dz, cudaStream_t st);
int32_t launch_rownorm_fwd(const float* u, const float* s, int64_t rows, int64_t D, float* f, cudaStream_t st);
int32_t launch_rownorm_bwd(const float* df, const float* f, const float* s, int64_t rows, int64_t D, float* du,
                           float* ds, cudaStream_t st);
int32_t launch_token_sum(const float* q, const float* mask, int64_t b, int64_t n_tok, int64_t D, float* out,
                         cudaStream_t st);
int32_t launch_token_bcast(const float* v, const float* mask, int64_t b, int64_t n_tok, int64_t D, float scale,
                           float* out, cudaStream_t st);

// nce_bwd_umma.cu (tcgen05)
int64_t nce_bwd_bf16_ws_bytes(int64_t b, int64_t B, int64_t D);
int32_t launch_nce_bwd_bf16(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt,
                            int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset, const float* lse_i,
                            const float* lse_t, const float* grad_out, float* dx_img, float* dx_txt, float* dg_img,
                            float* dg_txt, float* dscale, void* ws, int64_t ws_bytes, cudaStream_t st);
// gemm_wide2.cu
bool gemm_wide2_enabled();
bool gemm_wide2_supported(int64_t M, int64_t N, int64_t Kr, const float* out, int64_t ldo);
int32_t launch_gemm_wide2(const void* A, const void* B, int64_t M, int64_t N, int64_t Kr, float alpha, int accumulate,
                          float* out, int64_t ldo, cudaStream_t st, float* ws = nullptr, int col_mn = 0);
bool gemm_wide2_mn_supported(int64_t M, int64_t N, int64_t Kr, const float* out, int64_t ldo);
bool gemm_wide2_split_supported(int64_t M, int64_t N, int64_t Kr, const float* out, int64_t ldo);
int64_t gemm_wide2_ws_floats(int64_t M, int64_t N, int64_t Kr);
int32_t launch_nce_topk_bf16(const void* x, const void* g, int64_t b, int64_t B, int64_t D, const float* scale,
                             int64_t label_offset, const float* diag, int32_t* counts, void* ws, int64_t ws_bytes,
                             cudaStream_t st);
int32_t launch_topk_hits_acc(const float* L, int64_t rows, int64_t B, int64_t label0, int32_t* counts, cudaStream_t st);
int32_t launch_nce_bwd_gathered_bf16(const void* x_img, const void* x_txt, const void* g_img, const void* g_txt,
                                     int64_t b, int64_t B, int64_t D, const float* scale, int64_t label_offset,
                                     const float* lse_i_all, const float* lse_t_all, const float* go_all,
                                     float* dx_img, float* dx_txt, float* dscale, cudaStream_t st);

// qmap_rows.cu
int64_t ln_bwd_ws_floats(int64_t rows, int64_t C);
int64_t colsum_ws_floats(int64_t rows, int64_t C);
int32_t launch_ln_fwd(const void* src, int src_dtype, const float* bias, const float* gamma, const float* beta,
                      int64_t rows, int64_t C, float eps, int gelu, void* y, void* pre, float* mean, float* rstd,
                      cudaStream_t st);
int32_t launch_ln_bwd(const void* dy, int dy_dtype, const void* src, int src_dtype, const float* gamma, const float* mean,
                      const float* rstd, int64_t rows, int64_t C, int gelu, void* dsrc, int dsrc_dtype, float* dgamma,
                      float* dbeta, float* dbias, float* ws, cudaStream_t st);
int32_t launch_bias_cast(const float* src, const float* bias, int64_t rows, int64_t C, void* dst, int dst_dtype,
                         cudaStream_t st);
int32_t launch_colsum(const void* src, int src_dtype, int64_t rows, int64_t C, float* out, float* ws, cudaStream_t st);
int32_t launch_colsum_static(const void* src, int src_dtype, int64_t rows, int64_t C, float* out, cudaStream_t st);

}  // namespace ilfdt
