// rf_pipe_inst.cu -- one instantiation unit of the TMA-pipelined kernels: compile with
//   -DRF_PIPE_N=<power of two in [64, 2048]>
#include "rf_launch.h"

#ifndef RF_PIPE_N
#error "define RF_PIPE_N"
#endif

namespace rf {

constexpr int N = RF_PIPE_N;

template <class K>
static int set_smem_pipe(K kernel, int bytes) {
  return (int)cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

template <>
int PipeLaunch<N>::k3p(const pipe::K3PParams& p, int grid, cudaStream_t s) {
  constexpr int smem = pipe::k3p_smem_bytes<N>();
  int e;
  if ((e = set_smem_pipe(pipe::k3p_c2r_lines<N>, smem))) return e;
  pipe::k3p_c2r_lines<N><<<grid, pipe::kThreads, smem, s>>>(p);
  return (int)cudaGetLastError();
}
template <>
int PipeLaunch<N>::k3p_lines_per_tile() {
  return pipe::LPlan<N>::TL;
}

#if RF_PIPE_N <= 1024
template <>
int PipeLaunch<N>::k2p(const CUtensorMap& in, const CUtensorMap& out, const pipe::K2PParams& p, int grid,
                       cudaStream_t s) {
  constexpr int smem = pipe::k2p_smem_bytes<N>();
  int e;
  if ((e = set_smem_pipe(pipe::k2p_strided_fft<N>, smem))) return e;
  pipe::k2p_strided_fft<N><<<grid, pipe::kThreads, smem, s>>>(in, out, p);
  return (int)cudaGetLastError();
}
template <bool INJ, bool POT>
static int launch_k1p(const pipe::K1PMaps& out, const pipe::K1PParams& p, int grid, cudaStream_t s) {
  constexpr int smem = pipe::k1p_smem_bytes<N>();
  int e;
  if ((e = set_smem_pipe(pipe::k1p_generate_xfft<N, INJ, POT>, smem))) return e;
  pipe::k1p_generate_xfft<N, INJ, POT><<<grid, pipe::kConsumers, smem, s>>>(out, p);
  return (int)cudaGetLastError();
}
template <>
int PipeLaunch<N>::k1p(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p,
                       int grid, cudaStream_t s) {
  if (injected) return potential ? launch_k1p<true, true>(out, p, grid, s) : launch_k1p<true, false>(out, p, grid, s);
  return potential ? launch_k1p<false, true>(out, p, grid, s) : launch_k1p<false, false>(out, p, grid, s);
}
template <bool INJ, bool POT>
static int launch_k1q(const pipe::K1PMaps& out, const pipe::K1PParams& p, int grid, cudaStream_t s) {
  const int smem = pipe::k1q_smem_bytes<N>(INJ ? 0 : p.tab.lut_n);
  int e;
  if ((e = set_smem_pipe(pipe::k1q_generate_xfft<N, INJ, POT>, smem))) return e;
  pipe::k1q_generate_xfft<N, INJ, POT><<<grid, pipe::kConsumers, smem, s>>>(out, p);
  return (int)cudaGetLastError();
}
template <>
int PipeLaunch<N>::k1q(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p,
                       int grid, cudaStream_t s) {
  if (injected) return potential ? launch_k1q<true, true>(out, p, grid, s) : launch_k1q<true, false>(out, p, grid, s);
  return potential ? launch_k1q<false, true>(out, p, grid, s) : launch_k1q<false, false>(out, p, grid, s);
}
template <>
bool PipeLaunch<N>::k1q_fits(int hinge_cells) {
  // two CTAs (+1 KB reserved each) inside the 228 KB of an SM
  return pipe::k1q_smem_bytes<N>(hinge_cells) <= 112 * 1024;
}
template <>
int PipeLaunch<N>::tile_cols() {
  return pipe::SPlan<N>::TZ;
}
template <>
int PipeLaunch<N>::box_rows() {
  return pipe::SPlan<N>::kBoxRows;
}
#endif

}  // namespace rf
