// Shared host/device helpers for libilfdt (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <math.h>

#include "../../include/ilfdt.h"

namespace ilfdt {

// ---- per-thread error string (the only global state of the library) ------------------------
inline char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}
inline int32_t fail(int32_t code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define ILFDT_CUDA_OK(expr)                                                                    \
  do {                                                                                         \
    cudaError_t e__ = (expr);                                                                  \
    if (e__ != cudaSuccess) {                                                                  \
      (void)cudaGetLastError(); /* reported here: do not leave it for the next launch check */ \
      return ::ilfdt::fail(ILFDT_E_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,          \
                           cudaGetErrorString(e__));                                           \
    }                                                                                          \
  } while (0)

#define ILFDT_LAUNCH_OK()                                                                      \
  do {                                                                                         \
    cudaError_t e__ = cudaGetLastError();                                                      \
    if (e__ != cudaSuccess)                                                                    \
      return ::ilfdt::fail(ILFDT_E_CUDA, "%s:%d launch -> %s", __FILE__, __LINE__,             \
                           cudaGetErrorString(e__));                                           \
  } while (0)

#define ILFDT_REQUIRE(cond, code, ...)                                                         \
  do {                                                                                         \
    if (!(cond)) return ::ilfdt::fail(code, __VA_ARGS__);                                      \
  } while (0)

// Refuse to run anywhere but on an sm_100 part: there is deliberately no other code path.
int32_t require_sm100();
int sm_count();
// process-wide tuning knobs (ilfdt_tune_set / ilfdt_tune_get, include/ilfdt.h); key = ILFDT_TUNE_*
int64_t tune_get(int key);

inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline int64_t round_up(int64_t a, int64_t b) { return cdiv(a, b) * b; }

// ---- device helpers ----------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Order-preserving map float -> uint32 (so that unsigned compare == float compare; 0 is below all).
__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t o) {
  uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  return __uint_as_float(u);
}

// (a0, a1) += c * (v0, v1) as ONE packed fp32x2 FMA (Blackwell FFMA2): two independent round-to-nearest FMAs, so results
// equal two fmaf calls bit for bit, at half the issue slots. The scatter kernels are ~68 % issue-bound (ncu r02).
__device__ __forceinline__ void ffma2(float& a0, float& a1, float c, float v0, float v1) {
  asm("{\n .reg .b64 rc, rv, ra;\n mov.b64 rc, {%2, %2};\n mov.b64 rv, {%3, %4};\n mov.b64 ra, {%0, %1};\n"
      " fma.rn.f32x2 ra, rc, rv, ra;\n mov.b64 {%0, %1}, ra;\n}\n"
      : "+f"(a0), "+f"(a1)
      : "f"(c), "f"(v0), "f"(v1));
}

// 8 consecutive elements -> fp32, from either fp32 or bf16 storage (16-byte aligned for bf16,
// 32-byte for fp32).
template <typename T>
struct Load8;
template <>
struct Load8<float> {
  __device__ __forceinline__ static void ld(const float* p, float (&v)[8]) {
    float4 a = __ldg(reinterpret_cast<const float4*>(p));
    float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <>
struct Load8<__nv_bfloat16> {
  __device__ __forceinline__ static void ld(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
    uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};

}  // namespace ilfdt
