// rf_pipe_inst.cu -- one instantiation unit of the TMA-pipelined kernels: compile with
//   -DRF_PIPE_N=<power of two in [64, 4096]>
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
int PipeLaunch<N>::k3q(const pipe::K3PParams& p, int grid, cudaStream_t s) {
  constexpr int smem = pipe::k3q_smem_bytes<N>();
  if constexpr (smem <= 227 * 1024) {
    int e;
    if ((e = set_smem_pipe(pipe::k3q_c2r_lines<N>, smem))) return e;
    pipe::k3q_c2r_lines<N><<<grid, pipe::kThreads, smem, s>>>(p);
    return (int)cudaGetLastError();
  } else {
    return k3p(p, grid, s);   // short lines: the padded buffers of three tiles do not fit
  }
}
template <>
int PipeLaunch<N>::k3f(const pipe::K3PParams& p, int grid, cudaStream_t s) {
  constexpr int smem = pipe::k3q_smem_bytes<N>();
  if constexpr (smem <= 227 * 1024) {
    int e;
    if ((e = set_smem_pipe(pipe::k3f_r2c_lines<N>, smem))) return e;
    pipe::k3f_r2c_lines<N><<<grid, pipe::kThreads, smem, s>>>(p);
    return (int)cudaGetLastError();
  } else {
    return (int)cudaErrorInvalidConfiguration;   // nz = 64: the estimator uses the generic line transform
  }
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
template <bool INJ, bool POT>
static int launch_k1r(const pipe::K1PMaps& out, const pipe::K1PParams& p, int grid, cudaStream_t s) {
  const int smem = pipe::k1r_smem_bytes<N>(INJ ? 0 : p.tab.lut_n);
  int e;
  if ((e = set_smem_pipe(pipe::k1r_generate_xfft<N, INJ, POT>, smem))) return e;
  pipe::k1r_generate_xfft<N, INJ, POT><<<grid, pipe::kThreadsR, smem, s>>>(out, p);
  return (int)cudaGetLastError();
}
template <>
int PipeLaunch<N>::k1r(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p,
                       int grid, cudaStream_t s) {
  if (injected) return potential ? launch_k1r<true, true>(out, p, grid, s) : launch_k1r<true, false>(out, p, grid, s);
  return potential ? launch_k1r<false, true>(out, p, grid, s) : launch_k1r<false, false>(out, p, grid, s);
}
template <>
bool PipeLaunch<N>::k1r_fits(int hinge_cells) {
  return pipe::k1r_smem_bytes<N>(hinge_cells) <= 226 * 1024;
}
template <>
int PipeLaunch<N>::tile_cols() {
  return pipe::SPlan<N>::TZ;
}
template <>
int PipeLaunch<N>::box_rows() {
  return pipe::SPlan<N>::kBoxRows;
}
template <>
bool PipeLaunch<N>::k1p_fits(int) {
  return true;
}
template <>
int PipeLaunch<N>::ctas_per_sm() {
  return 1;
}
#else
// ---- axes of 2048 / 4096 points: single-buffer in-place kernels of rf_big.cuh ----
template <>
int PipeLaunch<N>::k2p(const CUtensorMap& in, const CUtensorMap& out, const pipe::K2PParams& p, int grid,
                       cudaStream_t s) {
  constexpr int smem = big::k2b_smem_bytes<N>();
  static_assert(smem <= 227 * 1024, "k2b shared memory");
  int e;
  if ((e = set_smem_pipe(big::k2b_strided_fft<N>, smem))) return e;
  (void)out;   // results are stored from registers
  big::k2b_strided_fft<N><<<grid, big::BPlan<N>::NT, smem, s>>>(in, p);
  return (int)cudaGetLastError();
}
template <bool INJ, bool POT>
static int launch_k1b(const pipe::K1PMaps& out, const pipe::K1PParams& p, int grid, cudaStream_t s) {
  const int smem = big::kb_smem_bytes<N>(INJ ? 0 : p.tab.lut_n * 16);
  int e;
  if ((e = set_smem_pipe(big::k1b_generate_xfft<N, INJ, POT>, smem))) return e;
  big::k1b_generate_xfft<N, INJ, POT><<<grid, big::BPlan<N>::NT, smem, s>>>(out, p);
  return (int)cudaGetLastError();
}
template <>
int PipeLaunch<N>::k1p(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p,
                       int grid, cudaStream_t s) {
  if (injected) return potential ? launch_k1b<true, true>(out, p, grid, s) : launch_k1b<true, false>(out, p, grid, s);
  return potential ? launch_k1b<false, true>(out, p, grid, s) : launch_k1b<false, false>(out, p, grid, s);
}
template <>
int PipeLaunch<N>::k1q(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p,
                       int grid, cudaStream_t s) {
  return k1p(injected, potential, out, p, grid, s);
}
template <>
bool PipeLaunch<N>::k1q_fits(int) {
  return false;
}
template <>
int PipeLaunch<N>::k1r(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p,
                       int grid, cudaStream_t s) {
  return k1p(injected, potential, out, p, grid, s);
}
template <>
bool PipeLaunch<N>::k1r_fits(int) {
  return false;
}
template <>
bool PipeLaunch<N>::k1p_fits(int hinge_cells) {
  return big::kb_smem_bytes<N>(hinge_cells * 16) <= (big::BPlan<N>::CTAS == 2 ? 112 : 226) * 1024;
}
template <>
int PipeLaunch<N>::ctas_per_sm() {
  return big::BPlan<N>::CTAS;
}
template <>
int PipeLaunch<N>::tile_cols() {
  return big::BPlan<N>::TZ;
}
template <>
int PipeLaunch<N>::box_rows() {
  return big::BGeom<N>::kBoxRows;
}
#endif

}  // namespace rf
