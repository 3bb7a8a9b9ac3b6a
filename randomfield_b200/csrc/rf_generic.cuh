// rf_generic.cuh -- size-agnostic kernels: stage-level operators on the public layout,
// the any-length line FFT used when a grid is outside the tile-kernel list (and by
// transform.Plan on host arrays), the density epilogue and the statistics reduction.
#pragma once
#include "rf_common.cuh"

namespace rf {

// ----------------------------------------------------------------------------------------
// k-space field in the public (nx, ny, nz/2+1) layout, reference normalisation
// (value = N3 * internal sigma * deviate).  One thread per mode.
// ----------------------------------------------------------------------------------------
template <typename T>
__global__ void gen_kspace_public_kernel(ModeGeom<T> g, SigmaTable<T> tab, cpx<T>* out,
                                         int32_t* bins, T renorm, int y0, int ny_local) {
  const int nzh = g.nz / 2 + 1;
  const long long total = (long long)g.nx * ny_local * nzh;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int iz = (int)(i % nzh);
    long long r = i / nzh;
    int iyl = (int)(r % ny_local);
    int ix = (int)(r / ny_local);
    int iy = y0 + iyl;
    cpx<T> v = sym_mode(g, tab, ix, iy, iz);
    out[i] = {v.x * renorm, v.y * renorm};
    if (bins) {
      T ksq = ksq_canon(mode_sq<T>(ix, g.nx), mode_sq<T>(iy, g.ny), mode_sq<T>(iz, g.nz),
                        g.k0x2, g.k0y2, g.k0z2);
      bins[i] = sigma_bin(tab, ksq);
    }
  }
}

// ----------------------------------------------------------------------------------------
// The packed-Nyquist column of the internal layout: out[ix][iy_local] = P0 + i*PN where P0 / PN
// are the symmetrised modes on the planes iz = 0 and iz = nz/2 (internal normalisation).
// 2*nx*ny modes; the fused generator (k1p) reads this instead of evaluating the Hermitian
// rule inside its hot loop.
// ----------------------------------------------------------------------------------------
template <typename T>
__global__ void gen_packed_plane_kernel(ModeGeom<T> g, SigmaTable<T> tab, cpx<T>* out, int y0,
                                        int ny_local) {
  const int total = g.nx * ny_local;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int iyl = i % ny_local, ix = i / ny_local;
    cpx<T> a = sym_mode(g, tab, ix, y0 + iyl, 0);
    cpx<T> b = sym_mode(g, tab, ix, y0 + iyl, g.nz / 2);
    out[i] = {a.x - b.y, a.y + b.x};
  }
}

// ----------------------------------------------------------------------------------------
// any-length FFT of one line held in shared memory.  x holds the input, y is scratch,
// tw holds exp(+2 pi i k / n).  Returns the buffer that holds the result.
// Power-of-two lengths: Stockham radix-2; otherwise a direct O(n^2) DFT.
// ----------------------------------------------------------------------------------------
template <typename T>
__device__ cpx<T>* smem_line_fft(cpx<T>* x, cpx<T>* y, const cpx<T>* tw, int n, bool inverse) {
  const int tid = threadIdx.x, nt = blockDim.x;
  if ((n & (n - 1)) == 0) {
    for (int ns = 1; ns < n; ns <<= 1) {
      for (int j = tid; j < n / 2; j += nt) {
        int k = j & (ns - 1);
        cpx<T> w = tw[k * (n / (2 * ns))];
        if (!inverse) w.y = -w.y;
        cpx<T> a = x[j], b = cmul(x[j + n / 2], w);
        int j0 = ((j - k) << 1) + k;
        y[j0] = a + b;
        y[j0 + ns] = a - b;
      }
      __syncthreads();
      cpx<T>* t = x; x = y; y = t;
    }
    return x;
  }
  for (int k = tid; k < n; k += nt) {
    double sr = 0.0, si = 0.0;
    int idx = 0;
    for (int j = 0; j < n; ++j) {
      cpx<T> w = tw[idx];
      double wy = inverse ? (double)w.y : -(double)w.y;
      sr += (double)x[j].x * w.x - (double)x[j].y * wy;
      si += (double)x[j].x * wy + (double)x[j].y * w.x;
      idx += k;
      if (idx >= n) idx -= n;
    }
    y[k] = {(T)sr, (T)si};
  }
  __syncthreads();
  return y;
}

template <typename T>
__device__ void smem_fill_twiddles(cpx<T>* tw, int n) {
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    double s, c;
    sincospi(2.0 * (double)k / (double)n, &s, &c);
    tw[k] = {(T)c, (T)s};
  }
}

// lines are enumerated as (a, b): offset = a*sa + b*sb ; points along the line at `stride`.
struct LineSet {
  long long na, nb, sa_in, sb_in, sa_out, sb_out, stride_in, stride_out;
  int n;
};

// mode 0: c2c.  mode 1: c2r (input n/2+1 complex, output n real).  mode 2: r2c.
template <typename T>
__global__ void generic_line_fft_kernel(const void* in_, void* out_, LineSet ls, int mode,
                                        int inverse, T scale) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n = ls.n;
  cpx<T>* x = reinterpret_cast<cpx<T>*>(smem_raw);
  cpx<T>* y = x + n;
  cpx<T>* tw = y + n;
  smem_fill_twiddles(tw, n);
  const long long nlines = ls.na * ls.nb;
  for (long long line = blockIdx.x; line < nlines; line += gridDim.x) {
    const long long a = line / ls.nb, b = line % ls.nb;
    const long long oin = a * ls.sa_in + b * ls.sb_in, oout = a * ls.sa_out + b * ls.sb_out;
    __syncthreads();
    if (mode == 0) {
      const cpx<T>* in = reinterpret_cast<const cpx<T>*>(in_) + oin;
      for (int j = threadIdx.x; j < n; j += blockDim.x) x[j] = in[j * ls.stride_in];
    } else if (mode == 1) {
      const cpx<T>* in = reinterpret_cast<const cpx<T>*>(in_) + oin;
      const int h = n / 2;
      for (int j = threadIdx.x; j <= h; j += blockDim.x) {
        cpx<T> v = in[j * ls.stride_in];
        if (j == 0 || j == h) v.y = (T)0;
        x[j] = v;
        if (j > 0 && j < h) x[n - j] = {v.x, -v.y};
      }
    } else {
      const T* in = reinterpret_cast<const T*>(in_) + oin;
      for (int j = threadIdx.x; j < n; j += blockDim.x) x[j] = {in[j * ls.stride_in], (T)0};
    }
    __syncthreads();
    cpx<T>* r = smem_line_fft(x, y, tw, n, inverse != 0);
    if (mode == 0) {
      cpx<T>* out = reinterpret_cast<cpx<T>*>(out_) + oout;
      for (int j = threadIdx.x; j < n; j += blockDim.x)
        out[j * ls.stride_out] = {r[j].x * scale, r[j].y * scale};
    } else if (mode == 1) {
      T* out = reinterpret_cast<T*>(out_) + oout;
      for (int j = threadIdx.x; j < n; j += blockDim.x) out[j * ls.stride_out] = r[j].x * scale;
    } else {
      cpx<T>* out = reinterpret_cast<cpx<T>*>(out_) + oout;
      for (int j = threadIdx.x; j <= n / 2; j += blockDim.x)
        out[j * ls.stride_out] = {r[j].x * scale, r[j].y * scale};
    }
    // make x/y reusable: smem_line_fft may have swapped them
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------------
// density epilogue (generate.py:266-273, cosmotools.py:214-221), in place.
//   lognormal: rho = exp(delta * a[z]) * b[z],  a = sqrt(ln t)/sigma, b = rho_mean/sqrt(t)
//   linear   : rho = (1 + a[z] * delta) * b[z], a = growth, b = rho_mean
// ----------------------------------------------------------------------------------------
template <typename T>
__global__ void density_kernel(T* field, long long n, int nz, const T* a, const T* b,
                               int lognormal) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    int z = (int)(i % nz);
    T d = field[i];
    T az = __ldg(a + z), bz = __ldg(b + z);
    field[i] = lognormal ? exp(d * az) * bz : ((T)1 + az * d) * bz;
  }
}
template <typename T>
__global__ void density_kernel_v4(T* field, long long n4, int nz, const T* a, const T* b,
                                  int lognormal) {
  // nz % 4 == 0: each thread handles 4 consecutive z
  using V = typename std::conditional<sizeof(T) == 4, float4, double4>::type;
  V* f4 = reinterpret_cast<V*>(field);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    int z = (int)((i * 4) % nz);
    V d = f4[i];
    T az[4], bz[4], v[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) { az[q] = __ldg(a + z + q); bz[q] = __ldg(b + z + q); }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      v[q] = lognormal ? exp(v[q] * az[q]) * bz[q] : ((T)1 + az[q] * v[q]) * bz[q];
    d.x = v[0]; d.y = v[1]; d.z = v[2]; d.w = v[3];
    f4[i] = d;
  }
}

// ----------------------------------------------------------------------------------------
// statistics: per-CTA (sum, sumsq) partials and a fixed-order final reduction
// ----------------------------------------------------------------------------------------
template <typename T>
__global__ void stats_partial_kernel(const T* field, long long n, double* partials) {
  double s1 = 0, s2 = 0;
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long b = blockIdx.x * per, e = min(n, b + per);
  for (long long i = b + threadIdx.x; i < e; i += blockDim.x) {
    double v = (double)field[i];
    s1 += v;
    s2 += v * v;
  }
  __shared__ double r1[256], r2[256];
  r1[threadIdx.x] = s1;
  r2[threadIdx.x] = s2;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      r1[threadIdx.x] += r1[threadIdx.x + o];
      r2[threadIdx.x] += r2[threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    partials[2 * blockIdx.x] = r1[0];
    partials[2 * blockIdx.x + 1] = r2[0];
  }
}

// single CTA of 256 threads; deterministic for a given nparts
__global__ void stats_final_kernel(const double* partials, long long nparts, double* out) {
  double s1 = 0, s2 = 0;
  for (long long i = threadIdx.x; i < nparts; i += 256) {
    s1 += partials[2 * i];
    s2 += partials[2 * i + 1];
  }
  __shared__ double r1[256], r2[256];
  r1[threadIdx.x] = s1;
  r2[threadIdx.x] = s2;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      r1[threadIdx.x] += r1[threadIdx.x + o];
      r2[threadIdx.x] += r2[threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out[0] = r1[0];
    out[1] = r2[0];
  }
}

// ----------------------------------------------------------------------------------------
// stage-level operators on public-layout arrays (module-level functions of the reference)
// ----------------------------------------------------------------------------------------
// powertools.fill_with_log10k (powertools.py:40-61): real <- 0.5*log10(kx^2+ky^2+kz^2) with
// the reference's rounding sequence (f64 adds, each stored to T), imag <- 0.
template <typename T>
__global__ void stage_fill_log10k_kernel(cpx<T>* data, int nx, int ny, int nz, int ncols,
                                         double lambda0) {
  const long long total = (long long)nx * ny * ncols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int iz = (int)(i % ncols);
    long long r = i / ncols;
    int iy = (int)(r % ny), ix = (int)(r / ny);
    // np.fft.fftfreq(n, d)[i] = j / (n*d), j the signed mode number
    auto freq = [&](int idx, int n) {
      int j = (idx < (n + 1) / 2) ? idx : idx - n;   // fftfreq: positive for idx < ceil(n/2)
      return (double)j / ((double)n * lambda0);
    };
    double kx = freq(ix, nx), ky = freq(iy, ny), kz = freq(iz, nz);
    T s1 = (T)(__dadd_rn(__dmul_rn(kx, kx), __dmul_rn(ky, ky)));
    T s2 = (T)(__dadd_rn((double)s1, __dmul_rn(kz, kz)));
    T l = log10(s2) * (T)0.5;
    data[i] = {l, (T)0};
  }
}

// powertools.tabulate_sigmas interpolation step (powertools.py:155-163): np.interp in f64,
// linear in log10 k, zero outside the table.
template <typename T>
__global__ void stage_tabulate_kernel(cpx<T>* data, long long total, int n,
                                      const double* xp, const double* fp) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    double x = (double)data[i].x;
    double r;
    if (!(x >= xp[0]) || x > xp[n - 1]) {
      r = 0.0;   // includes -inf (DC) and NaN
    } else if (x == xp[n - 1]) {
      r = fp[n - 1];
    } else {
      int lo = 0, hi = n - 1;   // xp[lo] <= x < xp[hi]
      while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (xp[mid] <= x) lo = mid; else hi = mid;
      }
      double slope = __ddiv_rn(__dsub_rn(fp[lo + 1], fp[lo]), __dsub_rn(xp[lo + 1], xp[lo]));
      r = __dadd_rn(__dmul_rn(slope, __dsub_rn(x, xp[lo])), fp[lo]);
    }
    data[i].x = (T)r;
  }
}

// random.randomize (random.py:12-29): imag <- real, then every scalar is multiplied by an
// independent N(0,1) deviate.  Philox counter = complex element index.
template <typename T>
__global__ void stage_randomize_kernel(cpx<T>* data, long long total, uint32_t key0,
                                       uint32_t key1) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    philox_out o = philox4x32_10((uint32_t)i, (uint32_t)((uint64_t)i >> 32), 1u, 0u, key0, key1);
    cpx<T> nrm = box_muller<T>(o.w[0], o.w[1]);
    T s = data[i].x;
    data[i] = {s * nrm.x, s * nrm.y};
  }
}

// transform.symmetrize (transform.py:114-158), packed and unpacked.  Out of place (src ->
// dst) so that masters are read before any slave is written.
template <typename T>
__global__ void stage_symmetrize_kernel(const cpx<T>* src, cpx<T>* dst, int nx, int ny, int nz,
                                        int packed) {
  const int ncols = packed ? nz / 2 + 1 : nz;
  const long long total = (long long)nx * ny * ncols;
  const int hx = nx / 2, hy = ny / 2, hz = nz / 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int iz = (int)(i % ncols);
    long long r = i / ncols;
    int iy = (int)(r % ny), ix = (int)(r / ny);
    const bool ex = (ix == 0 || ix == hx), ey = (iy == 0 || iy == hy), ez = (iz == 0 || iz == hz);
    const bool xh = ix > hx, yh = iy > hy, zh = iz > hz;   // "hi" half (never true on edges)
    bool slave;
    if (ez) {
      slave = yh || (ey && xh);
    } else if (packed) {
      slave = false;
    } else if (!ex && !ey) {
      // interior: masters are (lo,lo,lo),(hi,lo,lo),(lo,hi,lo),(lo,lo,hi)
      slave = ((int)xh + (int)yh + (int)zh) >= 2;
    } else {
      // x or y on an edge plane, z interior: the z-high half mirrors the z-low half
      slave = zh;
    }
    cpx<T> v;
    if (slave) {
      int mx = (nx - ix) % nx, my = (ny - iy) % ny, mz = (nz - iz) % nz;
      cpx<T> m = src[((long long)mx * ny + my) * ncols + mz];
      v = {m.x, -m.y};
    } else {
      v = src[i];
    }
    if (ex && ey && ez) v.y = (T)0;
    if (ix == 0 && iy == 0 && iz == 0) v.x = (T)0;
    dst[i] = v;
  }
}

}  // namespace rf
