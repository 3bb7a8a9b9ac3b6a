// rf_launch.h -- host-side launch entry points of the templated tile kernels.
// Each (T, N) pair is compiled in its own translation unit (rf_inst.cu with
// -DRF_INST_T=... -DRF_INST_N=...) so the build parallelises; rf_api.cu dispatches on N.
#pragma once
#include "rf_fast.cuh"
#include "rf_pipe.cuh"
#include "rf_big.cuh"
#include "rf_k3q.cuh"
#include "rf_k23q.cuh"
#include "rf_k3f.cuh"

namespace rf {

template <typename T, int N>
struct Launch {
  // returns cudaError_t as int
  static int k1(bool injected, const K1Params<T>& p, int grid, cudaStream_t s);
  static int k2(bool inverse, const K2Params<T>& p, int grid, cudaStream_t s);
  static int k3(const K3Params<T>& p, int grid, cudaStream_t s);
  static int k3_lines_per_cta();
};

// TMA-pipelined kernels (rf_pipe.cuh), float32, one unit per N (rf_pipe_inst.cu)
template <int N>
struct PipeLaunch {
  static int k3p(const pipe::K3PParams& p, int grid, cudaStream_t s);
  static int k3p_lines_per_tile();
  static int k3q(const pipe::K3PParams& p, int grid, cudaStream_t s);   // three-deep ring, in-place tiles
  // forward direction of the same axis: r2c lines (inverse sign; rf_k3f.cuh), N >= 128
  static int k3f(const pipe::K3PParams& p, int grid, cudaStream_t s);
  // N <= 1024: rf_pipe.cuh kernels; N = 2048, 4096: the single-buffer kernels of rf_big.cuh
  static int k2p(const CUtensorMap& in, const CUtensorMap& out, const pipe::K2PParams& p, int grid,
                 cudaStream_t s);
  static int k1p(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p, int grid, cudaStream_t s);
  // fused y + z pass (this N = ny <= 1024; nz in 128..2048): in-place tile bodies, three-deep ring (rf_k23q.cuh)
  static int k23q(int nz, const CUtensorMap& in, const CUtensorMap& out, const pipe::K23Params& p, int grid,
                  cudaStream_t s);
  // two-CTAs-per-SM variant of k1p; hinge_cells = entries of the sigma cell table staged in
  // shared memory.  k1q_fits() says whether that table fits next to the work buffer.
  static int k1q(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p, int grid, cudaStream_t s);
  static bool k1q_fits(int hinge_cells);
  // ping-pong variant: one 512-thread CTA per SM, two groups alternating generate / transform (N <= 1024)
  static int k1r(bool injected, bool potential, const pipe::K1PMaps& out, const pipe::K1PParams& p, int grid, cudaStream_t s);
  static bool k1r_fits(int hinge_cells);
  static bool k1p_fits(int hinge_cells);   // false: the sigma cell table does not fit (rf_big.cuh kernels)
  static int ctas_per_sm();  // resident CTAs per SM of k2p / k1p (grid = this * SM count)
  static int tile_cols();   // TZ
  static int box_rows();    // rows per TMA box of the store map
};

}  // namespace rf
