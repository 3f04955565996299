// Row-wise kernels for the non-default Query_model variants (clip_fdt.py:67-78,140-157):
// softmax / sigmoid attention over the codes, sigmoid's 1/sum(w) normalisation, and sum / mean token
// pooling expressed through the token-summed query. All HBM-bound, one CTA (or warp) per row.
#include "common.cuh"
#include "launch.h"

namespace ilfdt {

constexpr int RO_THREADS = 256;

__device__ __forceinline__ float block_sum_f(float v, float* sm) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < RO_THREADS / 32; ++w) t += sm[w];
  __syncthreads();
  return t;
}
__device__ __forceinline__ float block_max_f(float v, float* sm) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = sm[0];
#pragma unroll
  for (int w = 1; w < RO_THREADS / 32; ++w) t = fmaxf(t, sm[w]);
  __syncthreads();
  return t;
}

// kind 1: w = softmax(z) over the row (nn.Softmax(dim=-1), clip_fdt.py:72)
// kind 2: w = sigmoid(z), rowsum = sum_k w        (nn.Sigmoid, clip_fdt.py:76; sum for :156-157)
__global__ void __launch_bounds__(RO_THREADS)
att_fwd_kernel(const float* __restrict__ z, int K, int kind, float* __restrict__ w, float* __restrict__ rowsum) {
  __shared__ float sm[RO_THREADS / 32];
  const int64_t r = blockIdx.x;
  const float* zr = z + r * K;
  float* wr = w + r * K;
  if (kind == 1) {
    float mx = -INFINITY;
    for (int i = threadIdx.x; i < K; i += RO_THREADS) mx = fmaxf(mx, zr[i]);
    mx = block_max_f(mx, sm);
    float s = 0.f;
    for (int i = threadIdx.x; i < K; i += RO_THREADS) s += expf(zr[i] - mx);
    s = block_sum_f(s, sm);
    const float inv = 1.f / s;
    for (int i = threadIdx.x; i < K; i += RO_THREADS) wr[i] = expf(zr[i] - mx) * inv;
    if (rowsum && threadIdx.x == 0) rowsum[r] = 1.f;
  } else {
    float s = 0.f;
    for (int i = threadIdx.x; i < K; i += RO_THREADS) {
      float v = 1.f / (1.f + expf(-zr[i]));
      wr[i] = v;
      s += v;
    }
    s = block_sum_f(s, sm);
    if (rowsum && threadIdx.x == 0) rowsum[r] = s;
  }
}

// kind 1: dz = w * (g - sum_k w g);  kind 2: dz = (g + rowbias[r]) * w * (1 - w)
__global__ void __launch_bounds__(RO_THREADS)
att_bwd_kernel(const float* __restrict__ w, const float* __restrict__ g, const float* __restrict__ rowbias, int K,
               int kind, float* __restrict__ dz) {
  __shared__ float sm[RO_THREADS / 32];
  const int64_t r = blockIdx.x;
  const float* wr = w + r * K;
  const float* gr = g + r * K;
  float* dr = dz + r * K;
  if (kind == 1) {
    float s = 0.f;
    for (int i = threadIdx.x; i < K; i += RO_THREADS) s = fmaf(wr[i], gr[i], s);
    s = block_sum_f(s, sm);
    for (int i = threadIdx.x; i < K; i += RO_THREADS) dr[i] = wr[i] * (gr[i] - s);
  } else {
    const float bias = rowbias ? rowbias[r] : 0.f;
    for (int i = threadIdx.x; i < K; i += RO_THREADS) {
      float v = wr[i];
      dr[i] = (gr[i] + bias) * v * (1.f - v);
    }
  }
}

// f = u / s[row]   (clip_fdt.py:156-157), one warp per row
__global__ void rownorm_fwd_kernel(const float* __restrict__ u, const float* __restrict__ s, int64_t rows, int D,
                                   float* __restrict__ f) {
  int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const float den = s[r];
  for (int i = threadIdx.x & 31; i < D; i += 32) f[r * D + i] = u[r * D + i] / den;
}
// du = df / s ; ds = -(df . f) / s
__global__ void rownorm_bwd_kernel(const float* __restrict__ df, const float* __restrict__ f,
                                   const float* __restrict__ s, int64_t rows, int D, float* __restrict__ du,
                                   float* __restrict__ ds) {
  int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const float den = s[r];
  float dot = 0.f;
  for (int i = lane; i < D; i += 32) {
    float d = df[r * D + i];
    dot = fmaf(d, f[r * D + i], dot);
    du[r * D + i] = d / den;
  }
  dot = warp_sum(dot);
  if (lane == 0) ds[r] = -dot / den;
}

// out[b,:] = sum_t keep[b,t] * q[b,t,:]
__global__ void token_sum_kernel(const float* __restrict__ q, const float* __restrict__ mask, int n_tok, int D,
                                 float* __restrict__ out) {
  const int64_t b = blockIdx.x;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float s = 0.f;
    for (int t = 0; t < n_tok; ++t)
      if (mask == nullptr || mask[b * n_tok + t] == 0.f) s += q[(b * n_tok + t) * D + d];
    out[b * D + d] = s;
  }
}
// out[b,t,:] = keep[b,t] * scale * v[b,:]
__global__ void token_bcast_kernel(const float* __restrict__ v, const float* __restrict__ mask, int n_tok, int D,
                                   float scale, float* __restrict__ out) {
  const int64_t bt = blockIdx.x;
  const int64_t b = bt / n_tok;
  const float k = (mask == nullptr || mask[bt] == 0.f) ? scale : 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) out[bt * D + d] = k * v[b * D + d];
}

int32_t launch_att_fwd(const float* z, int64_t rows, int64_t K, int kind, float* w, float* rowsum, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  att_fwd_kernel<<<(unsigned)rows, RO_THREADS, 0, st>>>(z, (int)K, kind, w, rowsum);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_att_bwd(const float* w, const float* g, const float* rowbias, int64_t rows, int64_t K, int kind,
                       float* dz, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  att_bwd_kernel<<<(unsigned)rows, RO_THREADS, 0, st>>>(w, g, rowbias, (int)K, kind, dz);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_rownorm_fwd(const float* u, const float* s, int64_t rows, int64_t D, float* f, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  rownorm_fwd_kernel<<<(unsigned)cdiv(rows, 8), 256, 0, st>>>(u, s, rows, (int)D, f);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_rownorm_bwd(const float* df, const float* f, const float* s, int64_t rows, int64_t D, float* du,
                           float* ds, cudaStream_t st) {
  if (rows == 0) return ILFDT_OK;
  rownorm_bwd_kernel<<<(unsigned)cdiv(rows, 8), 256, 0, st>>>(df, f, s, rows, (int)D, du, ds);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_token_sum(const float* q, const float* mask, int64_t b, int64_t n_tok, int64_t D, float* out,
                         cudaStream_t st) {
  if (b == 0) return ILFDT_OK;
  token_sum_kernel<<<(unsigned)b, 256, 0, st>>>(q, mask, (int)n_tok, (int)D, out);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}
int32_t launch_token_bcast(const float* v, const float* mask, int64_t b, int64_t n_tok, int64_t D, float scale,
                           float* out, cudaStream_t st) {
  if (b == 0) return ILFDT_OK;
  ILFDT_REQUIRE(b * n_tok < (1ll << 31), ILFDT_E_UNSUPPORTED_SHAPE, "token_bcast: b*n_tok too large");
  token_bcast_kernel<<<(unsigned)(b * n_tok), 128, 0, st>>>(v, mask, (int)n_tok, (int)D, scale, out);
  ILFDT_LAUNCH_OK();
  return ILFDT_OK;
}

}  // namespace ilfdt
