// rf_fuse12_inst.cu -- instantiation of the fused pass-1 + pass-2 kernel for one N (nx == ny):
//   -DRF_PIPE_N=<512 | 1024>
#include "rf_launch.h"

#ifndef RF_PIPE_N
#error "define RF_PIPE_N"
#endif

namespace rf {

constexpr int N = RF_PIPE_N;

template <bool POT>
static int launch_k12(const pipe::K1PMaps& maps, const CUtensorMap& in, const CUtensorMap& out,
                      const fuse12::K12Params& p, int grid, cudaStream_t s) {
  const int smem = fuse12::k12_smem_bytes<N>(p.hinge_cells);
  int e = (int)cudaFuncSetAttribute(fuse12::k12_generate_xyfft<N, POT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e) return e;
  (void)out;   // pass 2 stores from registers
  fuse12::k12_generate_xyfft<N, POT><<<grid, fuse12::kThreads, smem, s>>>(maps, in, p);
  return (int)cudaGetLastError();
}

template <>
int PipeLaunch<N>::k12(bool potential, const pipe::K1PMaps& maps, const CUtensorMap& in, const CUtensorMap& out,
                       const fuse12::K12Params& p, int grid, cudaStream_t s) {
  return potential ? launch_k12<true>(maps, in, out, p, grid, s) : launch_k12<false>(maps, in, out, p, grid, s);
}
template <>
bool PipeLaunch<N>::k12_fits(int hinge_cells) {
  return fuse12::k12_smem_bytes<N>(hinge_cells) <= 227 * 1024;
}

}  // namespace rf
