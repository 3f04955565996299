// L2 -> SM gather probe (B200): how fast can the SMs pull 1 KB rows at random from an L2-RESIDENT table with no math?
// That is the ceiling of the scatter-backward kernels (csrc/ground_bwd.cu), which gather one 1 KB bf16 row of the
// codebook (dq) or of q (dsd) per support entry.     build + run: bash tools/l2_probe.sh
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

// one warp = one gatherer: U rows of 1 KB in flight (two 16-byte loads per lane and row), like the kernels' inner loops
template <int U>
__global__ void __launch_bounds__(256) gather_probe(const uint4* __restrict__ table, uint32_t rows, int iters, uint32_t* sink) {
  const int lane = threadIdx.x & 31;
  const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
    uint4 v[U][2];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t r = mix(gw * 9781u + it * 6271u + u * 17u) % rows;
      const uint4* row = table + (size_t)r * 64;
      v[u][0] = __ldg(row + lane);
      v[u][1] = __ldg(row + 32 + lane);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) acc ^= v[u][0].x ^ v[u][0].w ^ v[u][1].y ^ v[u][1].z;
  }
  if (acc == 0x12345u) sink[0] = acc;
}

template <int U>
static void run(const uint4* table, size_t bytes, int ctas_per_sm, int sms, uint32_t* sink) {
  const uint32_t rows = (uint32_t)(bytes / 1024);
  const int grid = sms * ctas_per_sm, iters = 2048 / U;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  gather_probe<U><<<grid, 256>>>(table, rows, iters, sink);   // warm: pulls the table into L2
  cudaEventRecord(e0);
  for (int rep = 0; rep < 5; ++rep) gather_probe<U><<<grid, 256>>>(table, rows, iters, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  ms /= 5;
  const double gathered = (double)grid * 8 * iters * U * 1024.0;
  printf("{\"table_MB\": %.0f, \"rows_in_flight_per_warp\": %d, \"ctas_per_sm\": %d, \"threads_per_sm\": %d, \"ms\": %.3f, "
         "\"gathered_GB\": %.1f, \"TB_per_s\": %.2f}\n", bytes / 1048576.0, U, ctas_per_sm, ctas_per_sm * 256, ms, gathered / 1e9,
         gathered / ms / 1e9);
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const size_t big = (size_t)2048 << 20;
  uint4* table; uint32_t* sink;
  cudaMalloc(&table, big); cudaMalloc(&sink, 4);
  cudaMemset(table, 1, big);
  // 16 MB = the bf16 codebook (dq gathers), 50 MB = one 1024-sample range of q (dsd gathers), 2048 MB = not L2-resident
  for (size_t mb : {16, 50, 100, 2048})
    for (int cps : {3, 8}) {
      run<2>(table, mb << 20, cps, sms, sink);
      run<4>(table, mb << 20, cps, sms, sink);
      run<8>(table, mb << 20, cps, sms, sink);
    }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
