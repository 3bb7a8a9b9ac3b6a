// rf_lens.cuh -- lensing potential psi(r) along the line of sight.
//
// Reference: Generator.calculate_lensing_potential (randomfield/generate.py:353-418).  For
// every slice i the reference integrates, with scipy.integrate.simps (composite Simpson on
// non-uniform abscissae, `even='avg'` for an even number of samples), the integrand
//     y_j = -2 (cotK_j - cotK_{i-1}) dPhi_j,   j in [i_min, i)
// over the comoving distances DC_j: an O(nz^2) loop per line of sight in numpy.
//
// The integral is linear in y and the Simpson weights of a point depend only on its
// neighbours' spacing and on the parity of its distance from i_min, so with f = dPhi and
// g = cotK * dPhi
//     psi_j = -2 ( I_j[g] - cotK_j * I_j[f] ),
//     I_j   = E_j                                    (j - i_min even: Simpson panels from i_min)
//     I_j   = (E_j + T_{j-1} + T_{i_min} + O_j) / 2   (j - i_min odd: scipy's 'avg' rule)
// where E_j / O_j are the running sums of the three-point panels ending at or before j whose
// first point is an even / odd number of samples from i_min, and T_s is the trapezoid on
// [s, s+1].  That is two parity-split prefix sums per function: O(nz) per line, 8 B/cell.
//
// One warp per line: the line is staged in shared memory (one pad word per lane block so the
// lanes' sequential sweeps are bank-conflict free), every lane sweeps a contiguous block once
// for its parity totals, one warp scan of four doubles turns them into prefixes, a second
// sweep writes psi in place.  Accumulation is in float64 (the reference's simps promotes to
// float64 through the float64 abscissae).  Coefficient tables (host-built, float64, nz each):
//   cotk[j]; (pa, pb, pd)[j]: weights of the panel (j-2, j-1, j); pt[j] = (DC_j - DC_{j-1}) / 2.
#pragma once
#include <cuda_runtime.h>

namespace rf {

struct LensTables {
  const double* cotk;
  const double* pa;
  const double* pb;
  const double* pd;
  const double* pt;
};

__device__ __forceinline__ double warp_excl_scan(double v, int lane) {
  double incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    double up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  return incl - v;
}

// field: [nlines][nz] real, psi overwrites dPhi.  blockDim.x = 32 * warps; dynamic shared
// memory: 5 * (nz + 32) doubles (the coefficient tables, in the same blocked + padded layout as
// the lines: a lane reads its own block's coefficients without touching global memory) followed
// by warps * (nz + 32) * sizeof(T).
__host__ __device__ inline size_t lens_smem_bytes(int nz, int warps, size_t elem) {
  return (size_t)5 * (nz + 32) * sizeof(double) + (size_t)warps * (nz + 32) * elem;
}
template <typename T>
__global__ void lensing_scan_kernel(T* __restrict__ field, long long nlines, int nz, int i_min, LensTables tb) {
  extern __shared__ __align__(16) unsigned char lens_smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int lb = (nz + 31) / 32;                      // samples per lane block
  const int pitch = nz + 32;
  double* const tab = reinterpret_cast<double*>(lens_smem);
  T* const sm = reinterpret_cast<T*>(tab + 5 * pitch) + (size_t)wid * pitch;
  // one pad word per lane block; lb is a power of two for the usual nz (shift instead of divide)
  const int lbs = (lb & (lb - 1)) == 0 ? __ffs(lb) - 1 : -1;
  auto blk = [&](int j) { return lbs >= 0 ? (j >> lbs) : j / lb; };
  auto at = [&](int j) -> T& { return sm[j + blk(j)]; };
  for (int k = threadIdx.x; k < nz; k += blockDim.x) {
    const int pk = k + blk(k);
    tab[pk] = __ldg(tb.cotk + k);
    tab[pitch + pk] = __ldg(tb.pa + k);
    tab[2 * pitch + pk] = __ldg(tb.pb + k);
    tab[3 * pitch + pk] = __ldg(tb.pd + k);
    tab[4 * pitch + pk] = __ldg(tb.pt + k);
  }
  __syncthreads();
  auto cotk = [&](int j) { return tab[j + blk(j)]; };
  const int b0 = lane * lb, b1 = min(nz, b0 + lb);
  const double t0w = (i_min + 1 < nz) ? tab[4 * pitch + (i_min + 1) + blk(i_min + 1)] : 0.0;

  for (long long line = (long long)blockIdx.x * nwarps + wid; line < nlines;
       line += (long long)gridDim.x * nwarps) {
    T* const gl = field + line * nz;
    // stage the line: 16 independent loads in flight per lane before the first shared store
    for (int k0 = lane; k0 < nz; k0 += 32 * 16) {
      T r[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const int k = k0 + 32 * u;
        r[u] = k < nz ? __ldcs(gl + k) : (T)0;
      }
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const int k = k0 + 32 * u;
        if (k < nz) at(k) = r[u];
      }
    }
    __syncwarp();
    // ---- sweep 1: parity totals of the panels ending inside this lane's block ----
    // (lanes whose block is empty -- nz < 32 * lb -- touch nothing)
    const bool act = b0 < nz;
    double f2 = (act && b0 >= 2) ? (double)at(b0 - 2) : 0.0, f1 = (act && b0 >= 1) ? (double)at(b0 - 1) : 0.0;
    double g2 = (act && b0 >= 2) ? cotk(b0 - 2) * f2 : 0.0, g1 = (act && b0 >= 1) ? cotk(b0 - 1) * f1 : 0.0;
    const double lf2 = f2, lf1 = f1, lg2 = g2, lg1 = g1;   // kept for sweep 2
    double sEf = 0, sEg = 0, sOf = 0, sOg = 0;
    for (int j = b0; j < b1; ++j) {
      const int pj = j + lane;   // j / lb == lane inside this lane's block
      const double f0 = (double)sm[pj], g0 = tab[pj] * f0;
      if (j >= i_min + 2) {
        const double a = tab[pitch + pj], b = tab[2 * pitch + pj], d = tab[3 * pitch + pj];
        const double qf = a * f2 + b * f1 + d * f0;
        const double qg = a * g2 + b * g1 + d * g0;
        if ((j - i_min) & 1) { sOf += qf; sOg += qg; } else { sEf += qf; sEg += qg; }
      }
      f2 = f1; f1 = f0; g2 = g1; g1 = g0;
    }
    double aEf = warp_excl_scan(sEf, lane), aEg = warp_excl_scan(sEg, lane);
    double aOf = warp_excl_scan(sOf, lane), aOg = warp_excl_scan(sOg, lane);
    // first trapezoid [i_min, i_min + 1] (scipy's 'avg' rule for even sample counts)
    double T0f = 0, T0g = 0;
    if (i_min + 1 < nz) {
      const double fa = (double)at(i_min), fb = (double)at(i_min + 1);
      T0f = t0w * (fa + fb);
      T0g = t0w * (cotk(i_min) * fa + cotk(i_min + 1) * fb);
    }
    __syncwarp();   // every lane has read its neighbours' samples: psi may overwrite dPhi
    // ---- sweep 2 ----
    f2 = lf2; f1 = lf1; g2 = lg2; g1 = lg1;
    for (int j = b0; j < b1; ++j) {
      const int pj = j + lane;
      const double f0 = (double)sm[pj], c0 = tab[pj], g0 = c0 * f0;
      double psi = 0.0;
      if (j >= i_min) {
        const bool odd = (j - i_min) & 1;
        if (j >= i_min + 2) {
          const double a = tab[pitch + pj], b = tab[2 * pitch + pj], d = tab[3 * pitch + pj];
          const double qf = a * f2 + b * f1 + d * f0;
          const double qg = a * g2 + b * g1 + d * g0;
          if (odd) { aOf += qf; aOg += qg; } else { aEf += qf; aEg += qg; }
        }
        double If, Ig;
        if (!odd) {
          If = aEf; Ig = aEg;
        } else {
          const double tw = tab[4 * pitch + pj];
          If = 0.5 * (aEf + tw * (f1 + f0)) + 0.5 * (T0f + aOf);
          Ig = 0.5 * (aEg + tw * (g1 + g0)) + 0.5 * (T0g + aOg);
        }
        psi = -2.0 * (Ig - c0 * If);
      }
      sm[pj] = (T)psi;
      f2 = f1; f1 = f0; g2 = g1; g1 = g0;
    }
    __syncwarp();
    for (int k = lane; k < nz; k += 32) gl[k] = at(k);
    __syncwarp();
  }
}

}  // namespace rf
