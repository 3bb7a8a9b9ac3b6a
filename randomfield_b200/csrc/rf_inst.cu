// rf_inst.cu -- one instantiation unit of the tile kernels: compile with
//   -DRF_INST_T=float|double -DRF_INST_N=<power of two in [64, 4096]>
#include "rf_launch.h"

#ifndef RF_INST_T
#error "define RF_INST_T and RF_INST_N"
#endif

namespace rf {

template <class K>
static int set_smem(K kernel, size_t bytes) {
  if (bytes > 48 * 1024)
    return (int)cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  return 0;
}

using T = RF_INST_T;
constexpr int N = RF_INST_N;

template <>
int Launch<T, N>::k1(bool injected, const K1Params<T>& p, int grid, cudaStream_t s) {
  constexpr size_t smem = strided_smem_bytes<T, N>();
  using P = FftPlan<T, N>;
  int e;
  if (injected) {
    if ((e = set_smem(k1_generate_xfft<T, N, true>, smem))) return e;
    k1_generate_xfft<T, N, true><<<grid, P::NT, smem, s>>>(p);
  } else {
    if ((e = set_smem(k1_generate_xfft<T, N, false>, smem))) return e;
    k1_generate_xfft<T, N, false><<<grid, P::NT, smem, s>>>(p);
  }
  return (int)cudaGetLastError();
}

template <>
int Launch<T, N>::k2(bool inverse, const K2Params<T>& p, int grid, cudaStream_t s) {
  constexpr size_t smem = strided_smem_bytes<T, N>();
  using P = FftPlan<T, N>;
  int e;
  if (inverse) {
    if ((e = set_smem(k2_strided_fft<T, N, true>, smem))) return e;
    k2_strided_fft<T, N, true><<<grid, P::NT, smem, s>>>(p);
  } else {
    if ((e = set_smem(k2_strided_fft<T, N, false>, smem))) return e;
    k2_strided_fft<T, N, false><<<grid, P::NT, smem, s>>>(p);
  }
  return (int)cudaGetLastError();
}

template <>
int Launch<T, N>::k3(const K3Params<T>& p, int grid, cudaStream_t s) {
  constexpr size_t smem = lines_smem_bytes<T, N>();
  using P = LinePlan<T, N>;
  int e;
  if ((e = set_smem(k3_c2r_lines<T, N>, smem))) return e;
  k3_c2r_lines<T, N><<<grid, P::NT, smem, s>>>(p);
  return (int)cudaGetLastError();
}

template <>
int Launch<T, N>::k3_lines_per_cta() {
  return LinePlan<T, N>::TL;
}

}  // namespace rf
