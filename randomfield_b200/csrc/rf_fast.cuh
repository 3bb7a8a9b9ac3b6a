// rf_fast.cuh -- tile FFT kernels for the seed -> delta(x) hot path (sm_100a).
//
//   K1  k1_generate_xfft : Philox + Box-Muller + sigma(k) + Hermitian pairing generated in
//                          registers, inverse FFT along x, one HBM write (4 B/cell).
//                          Replaces fill_with_log10k / tabulate_sigmas / randomize /
//                          symmetrize + the axis-0 pass of irfftn
//                          (generate.py:191-199,218; transform.py:303-315).
//   K2  k2_strided_fft   : inverse (or forward) complex FFT along a strided axis on
//                          [rows][TZ] tiles (axis-1 pass; also the slab-transposed read
//                          after the multi-GPU all-to-all).
//   K3  k3_c2r_lines     : half-length complex FFT + Hermitian unpack along the contiguous
//                          axis, in place, with the population mean/variance partial sums
//                          (generate.py:219) and an optional per-z scale (light-cone factor
//                          of generate.py:341-343) fused into the store.
//
// Internal k-space layout ("packed-Nyquist"): complex [nx][ny][nz/2]; column 0 holds
// P0 + i*PN where P0 / PN are the iz=0 and iz=nz/2 planes of the public (nx,ny,nz/2+1)
// array.  Both planes are Hermitian in (x,y) so their 2-D inverse transforms are real and
// separate again as the real and imaginary part after the x and y passes.  This makes the
// field exactly nx*ny*nz*sizeof(T) bytes at every stage and every row 128-byte aligned.
#pragma once
#include "rf_common.cuh"

namespace rf {

// ----------------------------------------------------------------------------------------
// shared-memory layouts
// ----------------------------------------------------------------------------------------
// strided axis: tile is [pos][TZ], lanes run along the TZ contiguous columns
template <int TZ>
struct LayoutStrided {
  static constexpr int kCols = TZ;
  __device__ static __forceinline__ int smidx(int pos, int c) { return pos * TZ + c; }
  __device__ static __forceinline__ void split(int w, int count, int& idx, int& c) {
    c = w % TZ;
    idx = w / TZ;
  }
};
// contiguous axis: tile is [line][pos] with one pad element every 32, lanes run along pos
template <int LINE_LEN>
struct LayoutLines {
  static constexpr int kPitch = LINE_LEN + LINE_LEN / 32 + 1;
  __device__ static __forceinline__ int smidx(int pos, int l) { return l * kPitch + pos + (pos >> 5); }
  __device__ static __forceinline__ void split(int w, int count, int& idx, int& l) {
    idx = w % count;
    l = w / count;
  }
};

template <typename T>
__device__ __forceinline__ cpx<T> ldg_cpx(const cpx<T>* p) {
  if constexpr (sizeof(T) == 4) {
    float2 v = __ldg(reinterpret_cast<const float2*>(p));
    return {v.x, v.y};
  } else {
    double2 v = __ldg(reinterpret_cast<const double2*>(p));
    return {v.x, v.y};
  }
}
template <typename T>
__device__ __forceinline__ cpx<T> ld_cpx(const cpx<T>* p) {
  if constexpr (sizeof(T) == 4) {
    float2 v = *reinterpret_cast<const float2*>(p);
    return {v.x, v.y};
  } else {
    double2 v = *reinterpret_cast<const double2*>(p);
    return {v.x, v.y};
  }
}
template <typename T>
__device__ __forceinline__ void st_cpx(cpx<T>* p, cpx<T> v) {
  if constexpr (sizeof(T) == 4) {
    *reinterpret_cast<float2*>(p) = make_float2(v.x, v.y);
  } else {
    *reinterpret_cast<double2*>(p) = make_double2(v.x, v.y);
  }
}

// twiddle multiply: table holds exp(+2*pi*i*k/N); the forward transform conjugates it.
template <bool INV, typename T>
__device__ __forceinline__ cpx<T> twmul(cpx<T> v, const cpx<T>* __restrict__ tw, int i) {
  cpx<T> t = ldg_cpx(tw + i);
  if constexpr (!INV) t.y = -t.y;
  return cmul(v, t);
}

// ----------------------------------------------------------------------------------------
// N-point FFT over a tile, N = R1*R2*R3 (R3 may be 1; R2 may be 1 => single step).
//   loadA(v, j, c)   fills v[m] with input element (j + (N/R1)*m) of column/line c
//   storeF(k, c, x)  receives output element k of column/line c
// `items_c` = number of columns/lines in the tile.  Work items are spread over NT threads.
// ----------------------------------------------------------------------------------------
struct CtaSync {
  __device__ __forceinline__ void operator()() const { __syncthreads(); }
};

template <typename T, int N, int R1, int R2, int R3, int NT, bool INV, class L, class LoadA, class StoreF,
          class Sync = CtaSync>
__device__ __forceinline__ void tile_fft(cpx<T>* __restrict__ sm, const cpx<T>* __restrict__ tw,
                                         int items_c, LoadA&& loadA, StoreF&& storeF,
                                         Sync sync = Sync()) {
  static_assert(R1 * R2 * R3 == N, "factorisation");
  constexpr int M = N / R1;
  const int tid = threadIdx.x;

  // ---- step A: R1-point DFTs over stride-M samples --------------------------------
  for (int w = tid; w < M * items_c; w += NT) {
    int j, c;
    L::split(w, M, j, c);
    cpx<T> v[R1];
    loadA(v, j, c);
    fft_dif<R1, INV>(v);
    if constexpr (R2 == 1) {
      sfor<0, R1>([&](auto ki) {
        constexpr int k1 = decltype(ki)::value;
        storeF(k1, c, v[brev<R1>(k1)]);
      });
    } else {
      sfor<0, R1>([&](auto ki) {
        constexpr int k1 = decltype(ki)::value;
        cpx<T> x = v[brev<R1>(k1)];
        if constexpr (k1 > 0) x = twmul<INV>(x, tw, j * k1);
        st_cpx(sm + L::smidx(k1 * M + j, c), x);
      });
    }
  }
  if constexpr (R2 == 1) return;
  sync();

  if constexpr (R3 == 1) {
    // ---- step B (final): R2-point DFTs, R2 == M ---------------------------------------
    for (int w = tid; w < R1 * items_c; w += NT) {
      int k1, c;
      L::split(w, R1, k1, c);
      cpx<T> v[R2];
      sfor<0, R2>([&](auto ji) {
        constexpr int j = decltype(ji)::value;
        v[j] = ld_cpx(sm + L::smidx(k1 * M + j, c));
      });
      fft_dif<R2, INV>(v);
      sfor<0, R2>([&](auto ki) {
        constexpr int k2 = decltype(ki)::value;
        storeF(k1 + R1 * k2, c, v[brev<R2>(k2)]);
      });
    }
  } else {
    constexpr int M2 = M / R2;  // == R3
    // ---- step B: R2-point DFTs over stride-M2 samples inside each k1 block -----------
    for (int w = tid; w < R1 * M2 * items_c; w += NT) {
      int q, c;
      L::split(w, R1 * M2, q, c);
      const int jp = q % M2, k1 = q / M2;
      cpx<T> v[R2];
      sfor<0, R2>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[m] = ld_cpx(sm + L::smidx(k1 * M + jp + M2 * m, c));
      });
      fft_dif<R2, INV>(v);
      sfor<0, R2>([&](auto ki) {
        constexpr int k2 = decltype(ki)::value;
        cpx<T> x = v[brev<R2>(k2)];
        if constexpr (k2 > 0) x = twmul<INV>(x, tw, R1 * jp * k2);
        st_cpx(sm + L::smidx(k1 * M + k2 * M2 + jp, c), x);
      });
    }
    sync();
    // ---- step C (final): R3-point DFTs --------------------------------------------------
    for (int w = tid; w < R1 * R2 * items_c; w += NT) {
      int q, c;
      L::split(w, R1 * R2, q, c);
      const int k1 = q % R1, k2 = q / R1;
      cpx<T> v[R3];
      sfor<0, R3>([&](auto ji) {
        constexpr int jp = decltype(ji)::value;
        v[jp] = ld_cpx(sm + L::smidx(k1 * M + k2 * M2 + jp, c));
      });
      fft_dif<R3, INV>(v);
      sfor<0, R3>([&](auto ki) {
        constexpr int k3 = decltype(ki)::value;
        storeF(k1 + R1 * (k2 + R2 * k3), c, v[brev<R3>(k3)]);
      });
    }
  }
}

// ----------------------------------------------------------------------------------------
// compile-time plan: factorisation, tile width and thread count per (T, N)
// ----------------------------------------------------------------------------------------
template <typename T, int N>
struct FftPlan;
#define RF_PLAN(TT, NN, A, B, C, TZV, NTV)                 \
  template <>                                              \
  struct FftPlan<TT, NN> {                                 \
    static constexpr int R1 = A, R2 = B, R3 = C, TZ = TZV, NT = NTV; \
    static constexpr int MINB = (NTV <= 128 ? 4 : NTV <= 256 ? 2 : 1); \
  };
// strided-axis plans (K1, K2): tile = N x TZ complex
RF_PLAN(float, 32, 8, 4, 1, 16, 128)
RF_PLAN(float, 64, 8, 8, 1, 16, 128)
RF_PLAN(float, 128, 16, 8, 1, 16, 256)
RF_PLAN(float, 256, 16, 16, 1, 16, 256)
RF_PLAN(float, 512, 16, 32, 1, 16, 512)
RF_PLAN(float, 1024, 32, 32, 1, 8, 256)
RF_PLAN(float, 2048, 16, 16, 8, 8, 512)
RF_PLAN(float, 4096, 16, 16, 16, 4, 512)
RF_PLAN(double, 32, 8, 4, 1, 8, 64)
RF_PLAN(double, 64, 8, 8, 1, 8, 64)
RF_PLAN(double, 128, 16, 8, 1, 8, 128)
RF_PLAN(double, 256, 16, 16, 1, 8, 128)
RF_PLAN(double, 512, 8, 8, 8, 8, 512)
RF_PLAN(double, 1024, 16, 8, 8, 4, 256)
RF_PLAN(double, 2048, 16, 16, 8, 4, 512)
RF_PLAN(double, 4096, 16, 16, 16, 2, 512)
#undef RF_PLAN

template <typename T, int N>
constexpr size_t strided_smem_bytes() {
  return (size_t)N * FftPlan<T, N>::TZ * sizeof(cpx<T>);
}

// ----------------------------------------------------------------------------------------
// K1: generate + inverse FFT along x
// ----------------------------------------------------------------------------------------
template <typename T>
struct K1Params {
  ModeGeom<T> geom;
  SigmaTable<T> tab;
  cpx<T>* out;            // [nx][ny_local][nz/2]
  const cpx<T>* tw;       // exp(+2 pi i k / nx), k < nx
  int ny_local, y0;       // y-slab owned by this rank (multi-GPU); y0 = first global iy
  // direct exchange (slab decomposition over peer memory): the x line is scattered straight
  // into the receive buffers of the ranks that own each x range.  peer[d] points at this
  // rank's source block [nx/G][ny_local][nz/2] inside rank d's receive buffer (peer[own] is
  // local memory); nxl_shift = log2(nx/G).  direct == 0: store to `out`.
  cpx<T>* peer[8];
  int direct, nxl_shift;
};

template <typename T, int NX, bool INJ>
__global__ void __launch_bounds__(FftPlan<T, NX>::NT, FftPlan<T, NX>::MINB)
k1_generate_xfft(const K1Params<T> p) {
  using P = FftPlan<T, NX>;
  using L = LayoutStrided<P::TZ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cpx<T>* sm = reinterpret_cast<cpx<T>*>(smem_raw);

  const int nzc = p.geom.nz / 2;
  const int ztiles = nzc / P::TZ;
  const int zt = blockIdx.x % ztiles;
  const int iyl = blockIdx.x / ztiles;
  const int iy = p.y0 + iyl;
  const int z0 = zt * P::TZ;
  constexpr int M = NX / P::R1;
  const ModeGeom<T>& g = p.geom;
  const int nzh = g.nz / 2 + 1;

  auto loadA = [&](cpx<T>(&v)[P::R1], int j, int c) {
    const int iz = z0 + c;
    if (iz == 0) {
      // packed column: (iz=0 plane) + i * (iz=nz/2 plane), each with the Hermitian rule
#pragma unroll 1
      for (int m = 0; m < P::R1; ++m) {
        const int ix = j + M * m;
        cpx<T> a = sym_mode(g, p.tab, ix, iy, 0);
        cpx<T> b = sym_mode(g, p.tab, ix, iy, g.nz / 2);
        cpx<T> r = {a.x - b.y, a.y + b.x};
        // static register indexing only: select with a compile-time sweep
        sfor<0, P::R1>([&](auto mi) {
          if (decltype(mi)::value == m) v[decltype(mi)::value] = r;
        });
      }
    } else if constexpr (INJ) {
      sfor<0, P::R1>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[m] = raw_mode(g, p.tab, j + M * m, iy, iz);
      });
    } else {
      const T jy2 = mode_sq<T>(iy, g.ny), jz2 = mode_sq<T>(iz, g.nz);
      sfor<0, P::R1 / 2>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        const int ix = j + M * m;                 // < nx/2; partner is ix + nx/2
        T ks1 = ksq_canon(mode_sq<T>(ix, NX), jy2, jz2, g.k0x2, g.k0y2, g.k0z2);
        T ks2 = ksq_canon(mode_sq<T>(ix + NX / 2, NX), jy2, jz2, g.k0x2, g.k0y2, g.k0z2);
        T s1 = sigma_eval(p.tab, ks1), s2 = sigma_eval(p.tab, ks2);
        if (g.potential) {
          s1 = (ks1 > (T)0) ? s1 / ks1 : (T)0;
          s2 = (ks2 > (T)0) ? s2 / ks2 : (T)0;
        }
        uint64_t q = ((uint64_t)ix * g.ny + iy) * nzh + iz;
        philox_out o = philox4x32_10((uint32_t)q, (uint32_t)(q >> 32), 0u, 0u, g.key0, g.key1);
        cpx<T> n1 = box_muller<T>(o.w[0], o.w[1]);
        cpx<T> n2 = box_muller<T>(o.w[2], o.w[3]);
        v[m] = {s1 * n1.x, s1 * n1.y};
        v[m + P::R1 / 2] = {s2 * n2.x, s2 * n2.y};
      });
    }
  };
  cpx<T>* outp = p.out + (size_t)iyl * nzc + z0;
  const size_t xstride = (size_t)p.ny_local * nzc;
  if (p.direct) {
    const int shift = p.nxl_shift, mask = (1 << shift) - 1;
    const size_t col = (size_t)iyl * nzc + z0;
    auto storeP = [&](int k, int c, cpx<T> x) {
      st_cpx(p.peer[k >> shift] + (size_t)(k & mask) * xstride + col + c, x);
    };
    tile_fft<T, NX, P::R1, P::R2, P::R3, P::NT, true, L>(sm, p.tw, P::TZ, loadA, storeP);
    return;
  }
  auto storeF = [&](int k, int c, cpx<T> x) { st_cpx(outp + (size_t)k * xstride + c, x); };
  tile_fft<T, NX, P::R1, P::R2, P::R3, P::NT, true, L>(sm, p.tw, P::TZ, loadA, storeF);
}

// ----------------------------------------------------------------------------------------
// K2: complex FFT along a strided axis (rows), tiles of TZ contiguous columns
// ----------------------------------------------------------------------------------------
template <typename T>
struct K2Params {
  const cpx<T>* in;
  cpx<T>* out;
  const cpx<T>* tw;        // exp(+2 pi i k / N)
  long long in_outer;      // element stride between outer indices (input)
  long long in_row;        // element stride between rows inside one source block
  long long in_src;        // element stride between source blocks (multi-GPU receive buffer)
  int rows_per_src;        // rows per source block (== N when single GPU)
  long long out_outer, out_row;
  int ncols;               // columns per row (multiple of TZ)
  T scale;                 // applied on store (1 for the hot path)
};

template <typename T, int N, bool INV>
__global__ void __launch_bounds__(FftPlan<T, N>::NT, FftPlan<T, N>::MINB)
k2_strided_fft(const K2Params<T> p) {
  using P = FftPlan<T, N>;
  using L = LayoutStrided<P::TZ>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cpx<T>* sm = reinterpret_cast<cpx<T>*>(smem_raw);
  const int ztiles = p.ncols / P::TZ;
  const int zt = blockIdx.x % ztiles;
  const long long o = blockIdx.x / ztiles;
  const int z0 = zt * P::TZ;
  constexpr int M = N / P::R1;
  const cpx<T>* inp = p.in + o * p.in_outer + z0;
  cpx<T>* outp = p.out + o * p.out_outer + z0;
  const int rps = p.rows_per_src;

  auto loadA = [&](cpx<T>(&v)[P::R1], int j, int c) {
    sfor<0, P::R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      const int r = j + M * m;
      const int src = r / rps, rl = r - src * rps;
      v[m] = ld_cpx(inp + src * p.in_src + rl * p.in_row + c);
    });
  };
  const T sc = p.scale;
  auto storeF = [&](int k, int c, cpx<T> x) {
    st_cpx(outp + (long long)k * p.out_row + c, cpx<T>{x.x * sc, x.y * sc});
  };
  tile_fft<T, N, P::R1, P::R2, P::R3, P::NT, INV, L>(sm, p.tw, P::TZ, loadA, storeF);
}

// ----------------------------------------------------------------------------------------
// K3: c2r along the contiguous axis, packed-Nyquist input, in place
// ----------------------------------------------------------------------------------------
template <typename T, int NZ>
struct LinePlan;
#define RF_LPLAN(TT, NN, A, B, C, TLV, NTV)                \
  template <>                                              \
  struct LinePlan<TT, NN> {                                \
    static constexpr int R1 = A, R2 = B, R3 = C, TL = TLV, NT = NTV; \
    static constexpr int MINB = (NTV <= 128 ? 4 : NTV <= 256 ? 2 : 1); \
  };
// complex length H = NZ/2 = R1*R2*R3; TL lines per CTA
RF_LPLAN(float, 64, 8, 4, 1, 32, 128)
RF_LPLAN(float, 128, 8, 8, 1, 16, 128)
RF_LPLAN(float, 256, 16, 8, 1, 16, 128)
RF_LPLAN(float, 512, 16, 16, 1, 8, 128)
RF_LPLAN(float, 1024, 16, 32, 1, 8, 256)
RF_LPLAN(float, 2048, 32, 32, 1, 4, 128)
RF_LPLAN(float, 4096, 16, 16, 8, 2, 256)
RF_LPLAN(double, 64, 8, 4, 1, 32, 128)
RF_LPLAN(double, 128, 8, 8, 1, 16, 128)
RF_LPLAN(double, 256, 16, 8, 1, 16, 128)
RF_LPLAN(double, 512, 16, 16, 1, 8, 128)
RF_LPLAN(double, 1024, 8, 8, 8, 4, 256)
RF_LPLAN(double, 2048, 16, 8, 8, 2, 128)
RF_LPLAN(double, 4096, 16, 16, 8, 2, 256)
#undef RF_LPLAN

template <typename T>
struct K3Params {
  T* buf;                  // in: complex [nlines][nz/2] ; out: real [nlines][nz] (same bytes)
  long long nlines;
  const cpx<T>* twH;       // exp(+2 pi i k / (nz/2))
  const cpx<T>* twN;       // exp(+2 pi i k / nz), k < nz/2
  const T* zscale;         // optional per-z factor (nz entries) or nullptr
  double* partials;        // optional [gridDim.x][2] (sum, sum of squares) or nullptr
};

template <typename T, int NZ>
constexpr size_t lines_smem_bytes() {
  return (size_t)LinePlan<T, NZ>::TL * LayoutLines<NZ / 2>::kPitch * sizeof(cpx<T>);
}

template <typename T, int NZ>
__global__ void __launch_bounds__(LinePlan<T, NZ>::NT, LinePlan<T, NZ>::MINB)
k3_c2r_lines(const K3Params<T> p) {
  using P = LinePlan<T, NZ>;
  constexpr int H = NZ / 2;
  using L = LayoutLines<H>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cpx<T>* sm = reinterpret_cast<cpx<T>*>(smem_raw);
  constexpr int M = H / P::R1;
  const long long line0 = (long long)blockIdx.x * P::TL;
  const int nl = (int)min((long long)P::TL, p.nlines - line0);
  cpx<T>* base = reinterpret_cast<cpx<T>*>(p.buf) + line0 * H;

  auto loadA = [&](cpx<T>(&v)[P::R1], int j, int l) {
    const cpx<T>* lp = base + (long long)l * H;
    sfor<0, P::R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      const int k = j + M * m;
      cpx<T> a = ld_cpx(lp + k);
      if (m == 0 && k == 0) {
        // X[0] = a.x (real), X[H] = a.y (real)
        v[m] = {a.x + a.y, a.x - a.y};
      } else {
        cpx<T> b = ld_cpx(lp + (H - k));
        cpx<T> e = {a.x + b.x, a.y - b.y};       // X[k] + conj(X[H-k])
        cpx<T> d = {a.x - b.x, a.y + b.y};       // X[k] - conj(X[H-k])
        cpx<T> w = ldg_cpx(p.twN + k);           // exp(+2 pi i k / NZ)
        cpx<T> o = cmul(d, w);
        v[m] = {e.x - o.y, e.y + o.x};           // E + i*O
      }
    });
  };
  double s1 = 0.0, s2 = 0.0;
  T a1 = 0, a2 = 0;
  const T* zs = p.zscale;
  auto storeF = [&](int k, int l, cpx<T> x) {
    a1 += x.x + x.y;
    a2 += x.x * x.x + x.y * x.y;
    if (zs) {
      x.x *= __ldg(zs + 2 * k);
      x.y *= __ldg(zs + 2 * k + 1);
    }
    st_cpx(base + (long long)l * H + k, x);
  };
  tile_fft<T, H, P::R1, P::R2, P::R3, P::NT, true, L>(sm, p.twH, nl, loadA, storeF);

  if (p.partials) {
    s1 = (double)a1;
    s2 = (double)a2;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    __shared__ double red[2][32];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) { red[0][wid] = s1; red[1][wid] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double t1 = 0, t2 = 0;
      for (int i = 0; i < P::NT / 32; ++i) { t1 += red[0][i]; t2 += red[1][i]; }
      p.partials[2 * (size_t)blockIdx.x] = t1;
      p.partials[2 * (size_t)blockIdx.x + 1] = t2;
    }
  }
}

}  // namespace rf
