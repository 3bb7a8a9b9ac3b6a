// rf_fused_inst.cu -- instantiations of the fused y+z kernel k23q for one NY (all NZ):
//   -DRF_PIPE_N=<64..1024>
#include "rf_launch.h"

#ifndef RF_PIPE_N
#error "define RF_PIPE_N"
#endif

namespace rf {

constexpr int NY = RF_PIPE_N;

template <int NZ>
static int launch23q(const CUtensorMap& in, const CUtensorMap& out, const pipe::K23Params& p, int grid,
                     cudaStream_t s) {
  constexpr int smem = pipe::k23q_smem_bytes<NY, NZ>();
  if constexpr (smem <= 227 * 1024) {
    int e = (int)cudaFuncSetAttribute(pipe::k23q_fused_yz<NY, NZ>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e) return e;
    pipe::k23q_fused_yz<NY, NZ><<<grid, pipe::kThreadsQ, smem, s>>>(in, out, p);
    return (int)cudaGetLastError();
  } else {
    return (int)cudaErrorInvalidConfiguration;   // the ring does not fit (nz = 64): rf_create never selects it
  }
}

template <>
int PipeLaunch<NY>::k23q(int nz, const CUtensorMap& in, const CUtensorMap& out, const pipe::K23Params& p,
                         int grid, cudaStream_t s) {
  switch (nz) {
    case 64: return launch23q<64>(in, out, p, grid, s);
    case 128: return launch23q<128>(in, out, p, grid, s);
    case 256: return launch23q<256>(in, out, p, grid, s);
    case 512: return launch23q<512>(in, out, p, grid, s);
    case 1024: return launch23q<1024>(in, out, p, grid, s);
    case 2048: return launch23q<2048>(in, out, p, grid, s);
  }
  return (int)cudaErrorInvalidValue;
}

}  // namespace rf
