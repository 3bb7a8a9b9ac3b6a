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

// field: [nlines][nz] real dPhi; psi goes to `out` (same layout; may alias field: every read of a
// line precedes its writes).  blockDim.x = 32 * warps; dynamic shared
// memory: 5 * (nz + 32) doubles (the coefficient tables, in the same blocked + padded layout as
// the lines: a lane reads its own block's coefficients without touching global memory) followed
// by warps * (nz + 32) * sizeof(T).
__host__ __device__ inline size_t lens_smem_bytes(int nz, int warps, size_t elem) {
  return (size_t)5 * (nz + 32) * sizeof(double) + (size_t)warps * (nz + 32) * elem;
}
template <typename T>
__global__ void lensing_scan_kernel(const T* field, T* out, long long nlines, int nz, int i_min, LensTables tb) {
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
    const T* const gl = field + line * nz;
    T* const ol = out + line * nz;
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
    for (int k = lane; k < nz; k += 32) ol[k] = at(k);
    __syncwarp();
  }
}


// ------------------------------------------------------------------------------------------
// lensing_scan4_kernel: the same scan with FIXED sample ownership (nz % 128 == 0, nz <= 2048).
//
// The kernel above is work-efficient but reads five float64 coefficients per sample and sweep
// out of shared memory: 24 shared-memory wavefronts per 32 cells bound it at ~3 ms for 1024^3,
// and 16 warps per SM leave its dependent chains exposed.  Here thread t of a line group owns
// samples S t .. S t + S - 1 of EVERY line the group processes (S = 8 for float32, 4 for
// float64), so its coefficients live in registers, loaded once per thread: the j < i_min + 2
// mask folded in as zero weights, the first trapezoid T_{i_min} folded in as the "panel" of
// sample i_min + 1, zero trapezoid weights below i_min so that psi = 0 there, and the factor
// -2 folded into every weight (exact).  Lines are read and written straight from / to global
// memory as 16-byte vectors, the next line's loads in flight during the current line's scan,
// and the parity classes of a thread's samples are compile-time (S t even, P = i_min & 1).
// Running sums: E' (even-class panels) and M' = (E' + O') / 2 (half of every panel), so
//   even-class sample:  -2 I = E'          odd-class:  -2 I = M' - pt_j (y_{j-1} + y_j).
// Prefix sums: thread-local over S samples, a shuffle scan over the warp's 32 S samples in T,
// and float64 carries across the warps of a line through shared memory (one barrier per line,
// double-buffered).  All reads of a line precede the barrier, all writes follow it: in place.
// ------------------------------------------------------------------------------------------
template <typename T> struct LVec;
template <> struct LVec<float> {
  float v[4];
  __device__ __forceinline__ void load(const float* p) {
    float4 r = __ldcs(reinterpret_cast<const float4*>(p));
    v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
  }
  __device__ __forceinline__ static void store(float* p, float a, float b, float c, float d) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(a, b, c, d));
  }
  __device__ __forceinline__ static void load2(const float* p, float& a, float& b) {
    float2 r = __ldcs(reinterpret_cast<const float2*>(p));
    a = r.x; b = r.y;
  }
};
template <> struct LVec<double> {
  double v[4];
  __device__ __forceinline__ void load(const double* p) {
    double2 r0 = __ldcs(reinterpret_cast<const double2*>(p)), r1 = __ldcs(reinterpret_cast<const double2*>(p) + 1);
    v[0] = r0.x; v[1] = r0.y; v[2] = r1.x; v[3] = r1.y;
  }
  __device__ __forceinline__ static void store(double* p, double a, double b, double c, double d) {
    __stcs(reinterpret_cast<double2*>(p), make_double2(a, b));
    __stcs(reinterpret_cast<double2*>(p) + 1, make_double2(c, d));
  }
  __device__ __forceinline__ static void load2(const double* p, double& a, double& b) {
    double2 r = __ldcs(reinterpret_cast<const double2*>(p));
    a = r.x; b = r.y;
  }
};

template <typename T, int TPB, int S, bool P>
__device__ __forceinline__ void lens4_body(const T* field, T* out, long long nlines, int nz, int i_min,
                                           const LensTables& tb, double* tot) {
  constexpr int W = TPB / 32, V = S / 4;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = nz / S;                  // threads per line
  const int lpc = TPB / G;               // lines per CTA iteration
  const int grp = tid / G, t = tid - grp * G;
  const int j0 = S * t;
  const int w0 = (grp * G) >> 5;         // first warp of this line group
  // ---- per-thread coefficients (float64 products, rounded once to T), all scaled by -2:
  //      q' = -2 q, so the running sums are -2 I and psi = I'[g] - cotK_j I'[f] ----
  // even-class sample:  I' = E' (inclusive);  odd-class:  I' = M' - pt_j (f_{j-1} + f_j)  with
  // M' = (E' + O') / 2 the running sum of q' / 2 over BOTH classes (inclusive)
  T a[S], b[S], d[S], c[S], th[S];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const int j = j0 + s;
    // (odd-class samples below i_min: no panels and no trapezoid, so psi = 0 there)
    const double ptj = (j >= 1 && j > i_min) ? __ldg(tb.pt + j) : 0.0;
    double A = 0.0, B = 0.0, D = 0.0;
    if (j >= i_min + 2) {
      A = __ldg(tb.pa + j); B = __ldg(tb.pb + j); D = __ldg(tb.pd + j);
    } else if (j == i_min + 1) {
      B = ptj; D = ptj;                  // T_{i_min}: enters every later odd-class sum once
    }
    a[s] = (T)(-2.0 * A); b[s] = (T)(-2.0 * B); d[s] = (T)(-2.0 * D);
    c[s] = (T)__ldg(tb.cotk + j);
    th[s] = (T)(-ptj);
  }
  // cotK of the previous thread's last two samples (g = cotK * f of the neighbours)
  const T cn2 = j0 >= 2 ? (T)__ldg(tb.cotk + j0 - 2) : (T)0, cn1 = j0 >= 1 ? (T)__ldg(tb.cotk + j0 - 1) : (T)0;
  const long long lstride = (long long)gridDim.x * lpc;
  const long long base0 = (long long)blockIdx.x * lpc, line0 = base0 + grp;
  // iterations of the CTA (uniform: barriers inside) and of this thread's line group
  const int nit = base0 < nlines ? (int)((nlines - 1 - base0) / lstride) + 1 : 0;
  const int my_nit = line0 < nlines ? (int)((nlines - 1 - line0) / lstride) + 1 : 0;
  const T* gp = field + line0 * nz + j0;
  T* op = out + line0 * nz + j0;
  const long long gstep = lstride * nz;
  LVec<T> cur[V];
  T nb2 = 0, nb3 = 0;
  auto fetch = [&](const T* gl, bool ok) {
    if (ok) {
#pragma unroll
      for (int v = 0; v < V; ++v) cur[v].load(gl + 4 * v);
      if (lane == 0 && t > 0) LVec<T>::load2(gl - 2, nb2, nb3);
    }
  };
  fetch(gp, my_nit > 0);
  int buf = 0;
  for (int it = 0; it < nit; ++it, buf ^= 1, gp += gstep, op += gstep) {
    T f[S];
#pragma unroll
    for (int s = 0; s < S; ++s) f[s] = cur[s >> 2].v[s & 3];
    T p2 = __shfl_up_sync(0xffffffffu, f[S - 2], 1), p3 = __shfl_up_sync(0xffffffffu, f[S - 1], 1);
    if (lane == 0) { p2 = nb2; p3 = nb3; }
    fetch(gp + gstep, it + 1 < my_nit);  // next line's samples, in flight during this line's scan
    // ---- panels (zero weights where the reference has none) ----
    T qf[S], qg[S], g[S];
    const T g2 = cn2 * p2, g3 = cn1 * p3;
#pragma unroll
    for (int s = 0; s < S; ++s) g[s] = c[s] * f[s];
#pragma unroll
    for (int s = 0; s < S; ++s) {
      const T x2 = s >= 2 ? f[s >= 2 ? s - 2 : 0] : (s == 0 ? p2 : p3);
      const T x1 = s >= 1 ? f[s >= 1 ? s - 1 : 0] : p3;
      const T y2 = s >= 2 ? g[s >= 2 ? s - 2 : 0] : (s == 0 ? g2 : g3);
      const T y1 = s >= 1 ? g[s >= 1 ? s - 1 : 0] : g3;
      qf[s] = a[s] * x2 + b[s] * x1 + d[s] * f[s];
      qg[s] = a[s] * y2 + b[s] * y1 + d[s] * g[s];
    }
    constexpr int e0 = P ? 1 : 0;        // even-class samples: e0, e0 + 2, ...
    T ef = qf[e0], eg = qg[e0], af = qf[1 - e0], agg = qg[1 - e0];
#pragma unroll
    for (int s = 2; s < S; s += 2) {
      ef += qf[e0 + s]; eg += qg[e0 + s];
      af += qf[1 - e0 + s]; agg += qg[1 - e0 + s];
    }
    T loc[4] = {ef, eg, (T)0.5 * (ef + af), (T)0.5 * (eg + agg)};
    T inc[4] = {loc[0], loc[1], loc[2], loc[3]};
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        T up = __shfl_up_sync(0xffffffffu, inc[k], o);
        if (lane >= o) inc[k] += up;
      }
    }
    double* const tb_w = tot + (size_t)buf * W * 4;
    if (lane == 31) {
      reinterpret_cast<double2*>(tb_w + 4 * warp)[0] = make_double2((double)inc[0], (double)inc[1]);
      reinterpret_cast<double2*>(tb_w + 4 * warp)[1] = make_double2((double)inc[2], (double)inc[3]);
    }
    __syncthreads();
    double car[4] = {0.0, 0.0, 0.0, 0.0};
    for (int w = w0; w < warp; ++w) {
      const double2 u0 = reinterpret_cast<const double2*>(tb_w + 4 * w)[0];
      const double2 u1 = reinterpret_cast<const double2*>(tb_w + 4 * w)[1];
      car[0] += u0.x; car[1] += u0.y; car[2] += u1.x; car[3] += u1.y;
    }
    T Ef = (T)car[0] + (inc[0] - loc[0]), Eg = (T)car[1] + (inc[1] - loc[1]);
    T Mf = (T)car[2] + (inc[2] - loc[2]), Mg = (T)car[3] + (inc[3] - loc[3]);
    T psi[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
      const bool odd = ((s & 1) != 0) != P;
      Mf = fma((T)0.5, qf[s], Mf);
      Mg = fma((T)0.5, qg[s], Mg);
      T If, Ig;
      if (!odd) {
        Ef += qf[s]; Eg += qg[s];
        If = Ef; Ig = Eg;
      } else {
        const T fm1 = s == 0 ? p3 : f[s == 0 ? 0 : s - 1];
        const T gm1 = s == 0 ? g3 : g[s == 0 ? 0 : s - 1];
        If = fma(th[s], fm1 + f[s], Mf);
        Ig = fma(th[s], gm1 + g[s], Mg);
      }
      psi[s] = Ig - c[s] * If;
    }
    if (it < my_nit) {
#pragma unroll
      for (int v = 0; v < V; ++v) LVec<T>::store(op + 4 * v, psi[4 * v], psi[4 * v + 1], psi[4 * v + 2], psi[4 * v + 3]);
    }
  }
}

// resident threads per SM the register budget is set for (85 registers at 768, 128 at 512)
template <typename T, int S> struct LensOcc { static constexpr int value = (sizeof(T) == 4 && S == 4) ? 768 : 512; };
// S samples per thread: 8 for float32 (the scan and the carries amortise over twice the
// samples; 36 coefficient registers), 4 for float64
template <typename T> struct LensS { static constexpr int value = sizeof(T) == 4 ? 8 : 4; };
template <typename T, int TPB, int S>
__global__ void __launch_bounds__(TPB, LensOcc<T, S>::value / TPB)
lensing_scan4_kernel(const T* field, T* out, long long nlines, int nz, int i_min, LensTables tb) {
  __shared__ __align__(16) double tot[2 * (TPB / 32) * 4];
  if (i_min & 1) lens4_body<T, TPB, S, true>(field, out, nlines, nz, i_min, tb, tot);
  else lens4_body<T, TPB, S, false>(field, out, nlines, nz, i_min, tb, tot);
}

// whole warps per line and whole lines per CTA: nz / S a power of two in [32, 512]
__host__ inline bool lens4_supported(int nz, int S) {
  const int G = nz / S;
  return nz % S == 0 && G >= 32 && G <= 512 && (G & (G - 1)) == 0;
}

}  // namespace rf
