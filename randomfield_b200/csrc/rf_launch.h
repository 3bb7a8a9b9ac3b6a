// rf_launch.h -- host-side launch entry points of the templated tile kernels.
// Each (T, N) pair is compiled in its own translation unit (rf_inst.cu with
// -DRF_INST_T=... -DRF_INST_N=...) so the build parallelises; rf_api.cu dispatches on N.
#pragma once
#include "rf_fast.cuh"

namespace rf {

template <typename T, int N>
struct Launch {
  // returns cudaError_t as int
  static int k1(bool injected, const K1Params<T>& p, int grid, cudaStream_t s);
  static int k2(bool inverse, const K2Params<T>& p, int grid, cudaStream_t s);
  static int k3(const K3Params<T>& p, int grid, cudaStream_t s);
  static int k3_lines_per_cta();
};

}  // namespace rf
