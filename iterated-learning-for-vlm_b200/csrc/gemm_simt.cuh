// fp32 SIMT tile contraction (parity mode). 128x128x16 CTA tile, 256 threads, 8x8 per thread,
// register-prefetch double buffering. C(m,n) = sum_k A(m,k) * B(n,k) with arbitrary strides.
// Used by every fp32-mode op so that the reference's true-fp32 matmuls (torch default,
// allow_tf32=False) are matched to <=1e-5.
#pragma once
#include "common.cuh"

namespace ilfdt {

struct StridedMat {
  const float* p;
  int64_t rs, cs;  // element (i, k) at p[i*rs + k*cs]
  int rows;        // extent of i
};

constexpr int SG_BM = 128, SG_BN = 128, SG_BK = 16, SG_THREADS = 256;

struct SimtSmem {
  float a[SG_BK][SG_BM + 4];
  float b[SG_BK][SG_BN + 4];
};

// Load this thread's two "quads" (4 elements each) of a 128 x 16 operand tile.
// kcontig: quad = 4 consecutive k of one row; else quad = 4 consecutive rows at one k.
__device__ __forceinline__ void simt_load_tile(const StridedMat& X, int row0, int k0, int Kr,
                                               bool kcontig, bool vec, float4 (&r)[2]) {
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    int qd = threadIdx.x + i * SG_THREADS;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (kcontig) {
      int row = row0 + (qd >> 2), k = k0 + (qd & 3) * 4;
      if (row < X.rows) {
        const float* p = X.p + (int64_t)row * X.rs + k;
        if (vec && k + 3 < Kr) {
          float4 t = __ldg(reinterpret_cast<const float4*>(p));
          v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (k + j < Kr) v[j] = __ldg(p + j * X.cs);
        }
      }
    } else {
      int k = k0 + (qd >> 5), row = row0 + (qd & 31) * 4;
      if (k < Kr) {
        const float* p = X.p + (int64_t)k * X.cs + (int64_t)row * X.rs;
        if (vec && row + 3 < X.rows) {
          float4 t = __ldg(reinterpret_cast<const float4*>(p));
          v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (row + j < X.rows) v[j] = __ldg(p + j * X.rs);
        }
      }
    }
    r[i] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

__device__ __forceinline__ void simt_store_tile(float (*s)[SG_BM + 4], bool kcontig,
                                                const float4 (&r)[2]) {
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    int qd = threadIdx.x + i * SG_THREADS;
    if (kcontig) {
      int row = qd >> 2, k = (qd & 3) * 4;
      s[k + 0][row] = r[i].x; s[k + 1][row] = r[i].y; s[k + 2][row] = r[i].z; s[k + 3][row] = r[i].w;
    } else {
      int k = qd >> 5, row = (qd & 31) * 4;
      *reinterpret_cast<float4*>(&s[k][row]) = r[i];
    }
  }
}

__device__ __forceinline__ bool simt_vec_ok(const StridedMat& X, int Kr, bool kcontig) {
  if ((reinterpret_cast<uintptr_t>(X.p) & 15) != 0) return false;
  if (kcontig) return (X.rs % 4 == 0) && (Kr % 4 == 0);
  return (X.cs % 4 == 0) && (X.rows % 4 == 0);
}

// Thread (tx = tid%16, ty = tid/16) owns rows {ty*4+i, 64+ty*4+i} and cols {tx*4+j, 64+tx*4+j}.
__device__ __forceinline__ int simt_row(int i) { return (i < 4 ? 0 : 60) + (threadIdx.x >> 4) * 4 + i; }
__device__ __forceinline__ int simt_col(int j) { return (j < 4 ? 0 : 60) + (threadIdx.x & 15) * 4 + j; }

__device__ __forceinline__ void simt_mainloop(const StridedMat& A, const StridedMat& B, int Kr,
                                              int m0, int n0, SimtSmem& sm, float (&acc)[8][8]) {
  const bool a_kc = (A.cs == 1), b_kc = (B.cs == 1);
  const bool a_vec = simt_vec_ok(A, Kr, a_kc), b_vec = simt_vec_ok(B, Kr, b_kc);
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float4 ra[2], rb[2];
  simt_load_tile(A, m0, 0, Kr, a_kc, a_vec, ra);
  simt_load_tile(B, n0, 0, Kr, b_kc, b_vec, rb);
  const int ty4 = (threadIdx.x >> 4) * 4, tx4 = (threadIdx.x & 15) * 4;
  for (int k0 = 0; k0 < Kr; k0 += SG_BK) {
    __syncthreads();
    simt_store_tile(sm.a, a_kc, ra);
    simt_store_tile(sm.b, b_kc, rb);
    __syncthreads();
    if (k0 + SG_BK < Kr) {
      simt_load_tile(A, m0, k0 + SG_BK, Kr, a_kc, a_vec, ra);
      simt_load_tile(B, n0, k0 + SG_BK, Kr, b_kc, b_vec, rb);
    }
#pragma unroll
    for (int k = 0; k < SG_BK; ++k) {
      float4 a0 = *reinterpret_cast<const float4*>(&sm.a[k][ty4]);
      float4 a1 = *reinterpret_cast<const float4*>(&sm.a[k][64 + ty4]);
      float4 b0 = *reinterpret_cast<const float4*>(&sm.b[k][tx4]);
      float4 b1 = *reinterpret_cast<const float4*>(&sm.b[k][64 + tx4]);
      float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
  }
}

}  // namespace ilfdt
