// Epilogue body shared by the single-CTA and the CTA-pair grounding kernels.
#pragma once
#include "umma.cuh"

namespace ilfdt {

using namespace umma;

// Max/argmax over the token columns of each sample with everything static: NTOK tokens per sample,
// SPT = 256 / NTOK samples per tile. ~3 instructions per accumulator element (FMNMX, FSETP, SEL), no
// loop-carried chain longer than one sample, TMEM loads double-buffered against the arithmetic.
template <int NTOK>
__device__ __forceinline__ void poolmax_static(uint32_t taddr, int s0, int n_samples, int row, bool row_ok,
                                               int m_rows, float pool_scale, float* __restrict__ pooled,
                                               int16_t* __restrict__ arg, const int* __restrict__ sample_map, int slot0) {
  constexpr int SPT = 256 / NTOK, NCOLS = SPT * NTOK, NCH = (NCOLS + 31) / 32;
  float best[SPT];
  int bidx[SPT];
  uint32_t v[2][32];
  tmem_ld32(taddr, v[0]);
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    tmem_ld_wait();
    if (c + 1 < NCH) tmem_ld32(taddr + (c + 1) * 32, v[(c + 1) & 1]);
#pragma unroll
    for (int s = 0; s < SPT; ++s) {
      const int lo = s * NTOK > c * 32 ? s * NTOK : c * 32;
      const int hi = (s + 1) * NTOK < c * 32 + 32 ? (s + 1) * NTOK : c * 32 + 32;
      if (lo < hi) {
        float cm = __uint_as_float(v[c & 1][lo - c * 32]);
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (c * 32 + j > lo && c * 32 + j < hi) cm = fmaxf(cm, __uint_as_float(v[c & 1][j]));
        int ci = lo - s * NTOK;
#pragma unroll
        for (int j = 31; j >= 0; --j)  // descending: the lowest token index among ties wins
          if (c * 32 + j >= lo && c * 32 + j < hi && __uint_as_float(v[c & 1][j]) == cm) ci = c * 32 + j - s * NTOK;
        if (lo == s * NTOK) { best[s] = cm; bidx[s] = ci; }
        else if (cm > best[s]) { best[s] = cm; bidx[s] = ci; }
        if (hi == (s + 1) * NTOK && row_ok && s0 + s < n_samples) {
          const int smp = sample_map ? __ldg(sample_map + slot0 + s0 + s) : s0 + s;
          const int64_t o = (int64_t)smp * m_rows + row;
          pooled[o] = best[s] * pool_scale;
          arg[o] = (int16_t)bidx[s];
        }
      }
    }
  }
}


}  // namespace ilfdt
