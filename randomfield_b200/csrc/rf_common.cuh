// rf_common.cuh -- shared device building blocks for the B200 field-synthesis path.
//
// Contents: complex type, compile-time loops, in-register radix-R DIF butterflies with
// constant twiddles, Philox4x32-10, Box-Muller, the sigma(k) table lookup, and the
// Hermitian pairing rule.  Everything here is header-only and used by both the fast
// tile kernels (rf_fast.cuh) and the stage-level kernels (rf_generic.cu).
//
// Reference behaviour restated (never copied): randomfield/powertools.py:27-61,125-164
// (k grid, sigma interpolation), randomfield/random.py:12-29 (sigma * normal deviates),
// randomfield/transform.py:140-158 (packed Hermitian fix-up).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <type_traits>

namespace rf {

// ----------------------------------------------------------------------------------------
// complex number, aligned so that float -> 8-byte and double -> 16-byte vector accesses
// ----------------------------------------------------------------------------------------
template <typename T>
struct alignas(2 * sizeof(T)) cpx {
  T x, y;
};

// Packed single precision (sm_100 add/mul/fma.f32x2 -> SASS FADD2 / FMUL2 / FFMA2): a cpx<float>
// is one 64-bit register pair, a complex add is ONE instruction and a complex multiply TWO.
// ptxas folds lane swaps, per-lane signs and scalar broadcasts into operand modifiers
// (R.F32x2.LO_HI.NP, R.F32), so the helpers below are written with plain float2 constructors.
// Each lane is an IEEE round-to-nearest operation: results do not depend on which kernel
// (fused or separate passes, one GPU or several) evaluates them.  RF_NO_F32X2 restores the
// scalar forms (A/B builds).
#if defined(__CUDA_ARCH__) && !defined(RF_NO_F32X2)
#define RF_PK 1
#else
#define RF_PK 0
#endif

__host__ __device__ __forceinline__ cpx<float> pk_add(float ax, float ay, float bx, float by) {
#if RF_PK
  const float2 r = __fadd2_rn(make_float2(ax, ay), make_float2(bx, by));
  return {r.x, r.y};
#else
  return {ax + bx, ay + by};
#endif
}
__host__ __device__ __forceinline__ cpx<float> pk_mul(float ax, float ay, float bx, float by) {
#if RF_PK
  const float2 r = __fmul2_rn(make_float2(ax, ay), make_float2(bx, by));
  return {r.x, r.y};
#else
  return {ax * bx, ay * by};
#endif
}
__host__ __device__ __forceinline__ cpx<float> pk_fma(float ax, float ay, float bx, float by, float cx, float cy) {
#if RF_PK
  const float2 r = __ffma2_rn(make_float2(ax, ay), make_float2(bx, by), make_float2(cx, cy));
  return {r.x, r.y};
#elif defined(__CUDA_ARCH__)
  return {__fmaf_rn(ax, bx, cx), __fmaf_rn(ay, by, cy)};
#else
  return {ax * bx + cx, ay * by + cy};
#endif
}
__host__ __device__ __forceinline__ cpx<double> pk_add(double ax, double ay, double bx, double by) {
  return {ax + bx, ay + by};
}
__host__ __device__ __forceinline__ cpx<double> pk_mul(double ax, double ay, double bx, double by) {
  return {ax * bx, ay * by};
}

template <typename T>
__host__ __device__ __forceinline__ cpx<T> operator+(cpx<T> a, cpx<T> b) {
  return pk_add(a.x, a.y, b.x, b.y);
}
template <typename T>
__host__ __device__ __forceinline__ cpx<T> operator-(cpx<T> a, cpx<T> b) {
  return pk_add(a.x, a.y, -b.x, -b.y);
}
// a * b = (a.x b.x - a.y b.y, a.y b.x + a.x b.y).  The float form pins the rounding sequence --
// t = (a.x b.x, a.y b.x) rounded, then one fused multiply-add per lane -- because the field
// must be bit-identical whichever kernels (fused or separate passes, one GPU or several)
// produce it; left to the compiler the choice of the fused product differs from kernel to kernel.
template <typename T>
__host__ __device__ __forceinline__ cpx<T> cmul(cpx<T> a, cpx<T> b) {
#ifdef __CUDA_ARCH__
  if constexpr (sizeof(T) == 4) {
    const cpx<float> t = pk_mul(a.x, a.y, b.x, b.x);
    return pk_fma(a.y, a.x, -b.y, b.y, t.x, t.y);
  }
#endif
  return {a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x};
}
template <typename T>
__host__ __device__ __forceinline__ cpx<T> cconj(cpx<T> a) {
  return {a.x, -a.y};
}
// a + i*b and a - i*b
template <typename T>
__host__ __device__ __forceinline__ cpx<T> add_i(cpx<T> a, cpx<T> b) {
  return pk_add(a.x, a.y, -b.y, b.x);
}
template <typename T>
__host__ __device__ __forceinline__ cpx<T> sub_i(cpx<T> a, cpx<T> b) {
  return pk_add(a.x, a.y, b.y, -b.x);
}

// c2r along the contiguous axis, Hermitian unpack of the pair (k, H-k) of a half-length line:
// a = X[k], c = X[H-k], w = exp(+2 pi i k / N); zk = E[k] + i O[k], zh = the same for H-k.
template <typename T>
__host__ __device__ __forceinline__ void c2r_unpack_pair(cpx<T> a, cpx<T> c, cpx<T> w, cpx<T>& zk, cpx<T>& zh) {
  const cpx<T> e = pk_add(a.x, a.y, c.x, -c.y);   // X[k] + conj(X[H-k])
  const cpx<T> d = pk_add(a.x, a.y, -c.x, c.y);   // X[k] - conj(X[H-k])
  const cpx<T> o = cmul(d, w);
  zk = pk_add(e.x, e.y, -o.y, o.x);
  zh = pk_add(e.x, -e.y, o.y, o.x);
}
// k = 0 and k = H/2 of the same unpack (X[0] = P0 + i PN holds two real values, w = i at H/2)
template <typename T>
__host__ __device__ __forceinline__ void c2r_unpack_ends(cpx<T> a, cpx<T> m, cpx<T>& z0, cpx<T>& zm) {
  z0 = pk_add(a.x, a.x, a.y, -a.y);
  zm = pk_mul(m.x, m.y, (T)2, (T)-2);
}
// running (sum, sum of squares) of a real field held as complex pairs: acc1 += x, acc2 += x*x per lane
__host__ __device__ __forceinline__ void stats_accum(cpx<float> x, cpx<float>& acc1, cpx<float>& acc2) {
  acc1 = pk_add(acc1.x, acc1.y, x.x, x.y);
  acc2 = pk_fma(x.x, x.y, x.x, x.y, acc2.x, acc2.y);
}

// ----------------------------------------------------------------------------------------
// compile-time loop: body receives std::integral_constant<int, I>
// ----------------------------------------------------------------------------------------
template <int B, int E, class F>
__host__ __device__ __forceinline__ void sfor(F&& f) {
  if constexpr (B < E) {
    f(std::integral_constant<int, B>{});
    sfor<B + 1, E>(f);
  }
}

// ----------------------------------------------------------------------------------------
// 32nd roots of unity as compile-time constants
// ----------------------------------------------------------------------------------------
__host__ __device__ constexpr double rf_c32(int k) {
  // cos(2*pi*k/32) for k in [0, 8]
  return k == 0   ? 1.0
         : k == 1 ? 0.98078528040323044913
         : k == 2 ? 0.92387953251128675613
         : k == 3 ? 0.83146961230254523708
         : k == 4 ? 0.70710678118654752440
         : k == 5 ? 0.55557023301960222474
         : k == 6 ? 0.38268343236508977173
         : k == 7 ? 0.19509032201612826785
                  : 0.0;
}
__host__ __device__ constexpr double rf_cos32(int k) {
  k = ((k % 32) + 32) % 32;
  if (k > 16) k = 32 - k;               // cos is even about 16
  return (k > 8) ? -rf_c32(16 - k) : rf_c32(k);
}
__host__ __device__ constexpr double rf_sin32(int k) { return rf_cos32(k - 8); }

// multiply by W32^IDX, where W = exp(+2*pi*i/32) for the inverse and exp(-...) forward.
template <int IDX, bool INV, typename T>
__device__ __forceinline__ cpx<T> mul_tw32(cpx<T> d) {
  constexpr int k = ((IDX % 32) + 32) % 32;
  if constexpr (k == 0) {
    return d;
  } else if constexpr (k == 16) {
    return {-d.x, -d.y};
  } else if constexpr (k == 8) {
    if constexpr (INV) return {-d.y, d.x};
    else return {d.y, -d.x};
  } else if constexpr (k == 24) {
    if constexpr (INV) return {d.y, -d.x};
    else return {-d.y, d.x};
  } else if constexpr (k == 4 || k == 12 || k == 20 || k == 28) {
    constexpr T s = (T)0.70710678118654752440;
    constexpr T cs = (k == 4 || k == 28) ? (T)1 : (T)-1;          // sign of cos
    constexpr T sn0 = (k == 4 || k == 12) ? (T)1 : (T)-1;         // sign of sin(+)
    constexpr T sn = INV ? sn0 : -sn0;
    // (cs + i sn) * (x + i y) * s = (cs x - sn y) + i (cs y + sn x)
    const cpx<T> u = pk_add(cs * d.x, cs * d.y, -sn * d.y, sn * d.x);
    return pk_mul(u.x, u.y, s, s);
  } else {
    constexpr T c = (T)rf_cos32(k);
    constexpr T s = (T)(INV ? rf_sin32(k) : -rf_sin32(k));
    if constexpr (sizeof(T) == 4) {
      // same rounding sequence as cmul(d, {c, s})
      const cpx<float> t = pk_mul(d.x, d.y, c, c);
      return pk_fma(d.y, d.x, -s, s, t.x, t.y);
    } else {
      return {d.x * c - d.y * s, d.x * s + d.y * c};
    }
  }
}

// ----------------------------------------------------------------------------------------
// in-register radix-R decimation-in-frequency DFT, R in {2,4,8,16,32}.
// Output k of the DFT ends up in v[brev<R>(k)].
// ----------------------------------------------------------------------------------------
template <int R>
__host__ __device__ constexpr int brev(int k) {
  int r = 0;
  for (int b = 1; b < R; b <<= 1) {
    r = (r << 1) | (k & 1);
    k >>= 1;
  }
  return r;
}

template <int R, int HALF, bool INV, typename T>
__device__ __forceinline__ void dif_stage(cpx<T> (&v)[R]) {
  sfor<0, R / (2 * HALF)>([&](auto gi) {
    constexpr int g = decltype(gi)::value * 2 * HALF;
    sfor<0, HALF>([&](auto ki) {
      constexpr int k = decltype(ki)::value;
      cpx<T> a = v[g + k], b = v[g + k + HALF];
      v[g + k] = a + b;
      v[g + k + HALF] = mul_tw32<(k * 16) / HALF, INV>(a - b);
    });
  });
  if constexpr (HALF > 1) dif_stage<R, HALF / 2, INV>(v);
}

template <int R, bool INV, typename T>
__device__ __forceinline__ void fft_dif(cpx<T> (&v)[R]) {
  static_assert(R == 1 || R == 2 || R == 4 || R == 8 || R == 16 || R == 32, "radix");
  if constexpr (R > 1) dif_stage<R, R / 2, INV>(v);
}

// ----------------------------------------------------------------------------------------
// Philox4x32-R (Salmon et al. 2011).  Counter = 128 bit, key = 64 bit.  R = RF_PHILOX_ROUNDS.
// ----------------------------------------------------------------------------------------
#ifndef RF_PHILOX_ROUNDS
#define RF_PHILOX_ROUNDS 10
#endif
struct philox_out {
  uint32_t w[4];
};

__host__ __device__ __forceinline__ philox_out philox4x32_10(uint32_t c0, uint32_t c1,
                                                             uint32_t c2, uint32_t c3,
                                                             uint32_t k0, uint32_t k1) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  constexpr uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < RF_PHILOX_ROUNDS; ++r) {
    uint64_t p0 = (uint64_t)M0 * c0;
    uint64_t p1 = (uint64_t)M1 * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  return {{c0, c1, c2, c3}};
}

// Box-Muller: two 32-bit words -> one complex standard-normal deviate (re, im each N(0,1)).
// u1 = (wa + 0.5) / 2^32 in (0,1) -- all 32 bits of the word: the radius reaches
// sqrt(2 * 33 * ln 2) = 6.76 (float32 rounds wa to 24 significant bits, which keeps the exponent,
// i.e. the resolution of the tail); angle = pi * (int32)wb / 2^31 in [-pi, pi).
template <typename T>
__device__ __forceinline__ cpx<T> box_muller(uint32_t wa, uint32_t wb);

// MUFU wrappers with flush-to-zero semantics: a bare MUFU, without the denormal pre/post
// scaling (FSETP + 2 predicated ops) the non-ftz forms carry.  Every argument on the hot path
// is a normal number or zero (2^-33 <= u1 <= 1, -2 ln u1 = 0 or >= 1.19e-7, k^2 >= 5.9e-7 up to 4096^3 at 2 Mpc/h),
// for which both forms return the same bits.
__device__ __forceinline__ float lg2_ftz(float x) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float sqrt_ftz(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <>
__device__ __forceinline__ cpx<float> box_muller<float>(uint32_t wa, uint32_t wb) {
  // u1 = (wa + 0.5) 2^-32 in (0, 1]: the scaling is exact, so log2(u1) keeps its resolution near
  // u1 = 1 (small radii); the top 128 values of wa round to u1 = 1, i.e. r = 0
  float u1 = fmaf(__uint2float_rn(wa), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
  float l2 = lg2_ftz(u1);                                   // MUFU.LG2
  float r2 = l2 * -1.3862943611198906f;                     // -2 ln2 * log2(u1) = -2 ln(u1)
  float r = sqrt_ftz(r2);                                   // MUFU.SQRT
  float ang = (float)(int32_t)wb * 1.4629180792671596e-9f;  // pi / 2^31
  float s, c;
  __sincosf(ang, &s, &c);
  return pk_mul(c, s, r, r);
}

template <>
__device__ __forceinline__ cpx<double> box_muller<double>(uint32_t wa, uint32_t wb) {
  double u1 = ((double)wa + 0.5) * (1.0 / 4294967296.0);
  double r = sqrt(-2.0 * log(u1));
  double s, c;
  sincospi((double)(int32_t)wb * (1.0 / 2147483648.0), &s, &c);
  return {r * c, r * s};
}

// ----------------------------------------------------------------------------------------
// sigma(k) table in device memory (built on the host by rf_set_power)
//
// knots[j]   : the user's table (k_j^2, log2 k_j^2, sigma_j, slope on segment j).
// cells[c]   : direct-indexed lookup.  Cell c covers the k^2 values whose float bit pattern,
//              shifted right by `shift`, equals lut_base + c (exponent + a few mantissa bits,
//              so cells are uniform inside each octave).  The host picks enough mantissa
//              bits that a cell holds at most one knot, which makes the evaluation
//              branch-free: sigma = c0 + c1*log2(k^2) with (c0, c1) of the segment below
//              (A) or above (B) the knot `t` inside the cell.  Cells that would hold more
//              than one knot, and the cell of the last knot, are flagged and take the
//              knot-walking path.
// ----------------------------------------------------------------------------------------
template <typename T>
struct alignas(4 * sizeof(T)) knot {
  T ksq;    // k_j^2 rounded to T (bin edges: bin j is ksq[j] <= k^2 < ksq[j+1])
  T l2;     // log2(k_j^2) rounded to T
  T sig;    // sigma_j (already divided by nx*ny*nz)
  T slope;  // d sigma / d log2(k^2) on segment j
};

template <typename T>
struct alignas(8 * sizeof(T)) sigcell {
  T c0a, c1a, c0b, c1b;   // sigma = c0 + c1 * log2(k^2) below / at-or-above t
  T t;                    // knot inside the cell (k^2), +inf if none
  T bin_a;                // canonical bin index below t, stored as a number (exact)
  T slow;                 // != 0: walk the knots instead
  T pad;
};

template <typename T>
struct SigmaTable {
  const knot<T>* knots;     // n entries
  const sigcell<T>* cells;  // lut_n entries; first and last are all-zero sentinels
  const void* hinge;        // float4 {c0, c1, lt, d} per cell (float path only), see sigma_hinge
  int n;
  int lut_n;
  int lut_base;             // cell id of cells[0]
  int shift;                // cell id = (high bits of k^2) >> shift
};

__device__ __forceinline__ int coarse_cell(float v, int shift) { return __float_as_int(v) >> shift; }
__device__ __forceinline__ int coarse_cell(double v, int shift) { return __double2hiint(v) >> shift; }
__device__ __forceinline__ float rf_log2(float v) { return __log2f(v); }
__device__ __forceinline__ double rf_log2(double v) { return log2(v); }
__device__ __forceinline__ float rf_tiny(float) { return 1.17549435e-38f; }
__device__ __forceinline__ double rf_tiny(double) { return 2.2250738585072014e-308; }

// knot-walking evaluation from a start bin (cells flagged slow).  Returns sigma; *bin
// receives the canonical bin (largest j with knots[j].ksq <= ksq, or -1).
template <typename T>
__device__ __noinline__ T sigma_walk(const SigmaTable<T>& tab, T ksq, int j, int* bin) {
  while (j + 1 < tab.n && __ldg(&tab.knots[j + 1].ksq) <= ksq) ++j;
  *bin = j;
  bool outside = (j < 0) || (j == tab.n - 1 && ksq > __ldg(&tab.knots[tab.n - 1].ksq));
  int seg = max(min(j, tab.n - 2), 0);
  const knot<T>* kp = tab.knots + seg;
  T l2 = __ldg(&kp->l2), sg = __ldg(&kp->sig), sl = __ldg(&kp->slope);
  T v = sg + sl * (rf_log2(ksq) - l2);
  return outside ? (T)0 : v;
}

template <typename T>
__device__ __forceinline__ T sigma_eval(const SigmaTable<T>& tab, T ksq, int* bin_out = nullptr) {
  // cells[0] / cells[lut_n-1] are the all-zero "below the table" / "above the table" cells
  const int c = min(max(coarse_cell(ksq, tab.shift) - tab.lut_base, 0), tab.lut_n - 1);
  const sigcell<T>* cp = tab.cells + c;
  T c0a, c1a, c0b, c1b, t, bin_a, slow;
  if constexpr (sizeof(T) == 4) {
    float4 lo = __ldg(reinterpret_cast<const float4*>(cp));
    float4 hi = __ldg(reinterpret_cast<const float4*>(cp) + 1);
    c0a = lo.x; c1a = lo.y; c0b = lo.z; c1b = lo.w; t = hi.x; bin_a = hi.y; slow = hi.z;
  } else {
    const double2* q = reinterpret_cast<const double2*>(cp);
    double2 a = __ldg(q), b = __ldg(q + 1), d = __ldg(q + 2), e = __ldg(q + 3);
    c0a = a.x; c1a = a.y; c0b = b.x; c1b = b.y; t = d.x; bin_a = d.y; slow = e.x;
  }
  if (slow != (T)0) {
    int bin;
    T v = sigma_walk(tab, ksq, (int)bin_a, &bin);
    if (bin_out) *bin_out = bin;
    return v;
  }
  const bool up = ksq >= t;
  if (bin_out) *bin_out = (int)bin_a + (up ? 1 : 0);
  const T c0 = up ? c0b : c0a, c1 = up ? c1b : c1a;
  return c0 + c1 * rf_log2(fmax(ksq, rf_tiny((T)0)));
}

// Branch-free single-load evaluation used by the fused generator:
//   sigma = c0 + c1*lg + d*max(lg - lt, 0),  lg = log2(k^2)
// i.e. the segment below the cell's knot plus a hinge of slope change d at the knot (log2 of
// the knot = lt; d = 0 when the cell holds no knot).  Valid for every cell that lies strictly
// between the first and the last knot's cells and holds at most one knot; the host only
// enables kernels that use it when all non-zero modes of the grid fall in such cells.
__device__ __forceinline__ float sigma_hinge(const SigmaTable<float>& tab, float ksq) {
  const int c = min(max(coarse_cell(ksq, tab.shift) - tab.lut_base, 0), tab.lut_n - 1);
  const float4 h = __ldg(reinterpret_cast<const float4*>(tab.hinge) + c);
  const float lg = lg2_ftz(ksq);
  return fmaf(h.w, fmaxf(lg - h.z, 0.f), fmaf(h.y, lg, h.x));
}
// same, with the cell table staged in shared memory and indexed without clamping: `hinge0`
// points at the (virtual) entry of cell id 0, and the host's hinge_ok check guarantees that the
// cell of every non-zero mode of the grid lies inside the table
__device__ __forceinline__ float sigma_hinge_smem(const float4* hinge0, int shift, float ksq) {
  const float4 h = hinge0[coarse_cell(ksq, shift)];
  const float lg = lg2_ftz(ksq);
  return fmaf(h.w, fmaxf(lg - h.z, 0.f), fmaf(h.y, lg, h.x));
}

template <typename T>
__device__ __forceinline__ int sigma_bin(const SigmaTable<T>& tab, T ksq) {
  int b;
  (void)sigma_eval(tab, ksq, &b);
  return b;
}

// ----------------------------------------------------------------------------------------
// grid geometry + mode generation
// ----------------------------------------------------------------------------------------
template <typename T>
struct ModeGeom {
  int nx, ny, nz;          // full grid
  T k0x2, k0y2, k0z2;      // (2*pi/(n*spacing))^2 per axis, rounded to T
  uint32_t key0, key1;     // Philox key (seed)
  const double* noise;     // injected white noise in reference order, or nullptr
  int potential;           // 1: scale modes by 1/k^2 (Newtonian potential source)
};

// signed mode number squared as T (exact for n <= 8192)
template <typename T>
__device__ __forceinline__ T mode_sq(int i, int n) {
  int j = (i <= n / 2) ? i : i - n;
  return (T)(j * j);
}

// canonical k^2: fma(jx^2, k0x2, fma(jy^2, k0y2, jz^2 * k0z2)) with explicit rounding
__device__ __forceinline__ float ksq_canon(float jx2, float jy2, float jz2, float a, float b, float c) {
  return __fmaf_rn(jx2, a, __fmaf_rn(jy2, b, __fmul_rn(jz2, c)));
}
__device__ __forceinline__ double ksq_canon(double jx2, double jy2, double jz2, double a, double b, double c) {
  return __fma_rn(jx2, a, __fma_rn(jy2, b, __dmul_rn(jz2, c)));
}

// Hermitian pairing on the planes iz in {0, nz/2} (public packed layout):
// an element is a slave iff iy > ny/2, or iy in {0, ny/2} and ix > nx/2; its master is
// ((nx-ix)%nx, (ny-iy)%ny, iz).   (transform.py:140-158, verified against it in tests.)
struct Pairing {
  int mx, my;
  bool conj;
  bool vertex;
};
__device__ __forceinline__ Pairing hermitian_pairing(int ix, int iy, int iz, int nx, int ny, int nz) {
  Pairing p{ix, iy, false, false};
  if (iz == 0 || iz == nz / 2) {
    bool yedge = (iy == 0 || iy == ny / 2);
    bool slave = (iy > ny / 2) || (yedge && ix > nx / 2);
    if (slave) {
      p.mx = (nx - ix) % nx;
      p.my = (ny - iy) % ny;
      p.conj = true;
    }
    p.vertex = yedge && (ix == 0 || ix == nx / 2);
  }
  return p;
}

// raw (unsymmetrised) mode: sigma(|k|) * complex normal deviate.
// Philox counter: q = ((ix mod nx/2) * ny + iy) * (nz/2+1) + iz; words (0,1) serve
// ix < nx/2 and words (2,3) serve ix >= nx/2.
template <typename T>
__device__ __forceinline__ cpx<T> raw_mode(const ModeGeom<T>& g, const SigmaTable<T>& tab,
                                           int ix, int iy, int iz) {
  T ksq = ksq_canon(mode_sq<T>(ix, g.nx), mode_sq<T>(iy, g.ny), mode_sq<T>(iz, g.nz),
                    g.k0x2, g.k0y2, g.k0z2);
  T sig = sigma_eval(tab, ksq);
  if (g.potential) sig = (ksq > (T)0) ? sig / ksq : (T)0;
  const int nzh = g.nz / 2 + 1;
  if (g.noise) {
    size_t idx = 2 * (((size_t)ix * g.ny + iy) * nzh + iz);
    double nr = __ldg(g.noise + idx), ni = __ldg(g.noise + idx + 1);
    return {(T)((double)sig * nr), (T)((double)sig * ni)};
  }
  const int hx = g.nx / 2;
  int px = (ix >= hx) ? ix - hx : ix;
  uint64_t q = ((uint64_t)px * g.ny + iy) * nzh + iz;
  philox_out o = philox4x32_10((uint32_t)q, (uint32_t)(q >> 32), 0u, 0u, g.key0, g.key1);
  cpx<T> n = (ix >= hx) ? box_muller<T>(o.w[2], o.w[3]) : box_muller<T>(o.w[0], o.w[1]);
  return pk_mul(n.x, n.y, sig, sig);
}

// symmetrised mode in the public (nx, ny, nz/2+1) layout
template <typename T>
__device__ __forceinline__ cpx<T> sym_mode(const ModeGeom<T>& g, const SigmaTable<T>& tab,
                                           int ix, int iy, int iz) {
  Pairing p = hermitian_pairing(ix, iy, iz, g.nx, g.ny, g.nz);
  cpx<T> v = raw_mode(g, tab, p.mx, p.my, iz);
  if (p.conj) v.y = -v.y;
  if (p.vertex) v.y = (T)0;
  if (ix == 0 && iy == 0 && iz == 0) v.x = (T)0;
  return v;
}

}  // namespace rf
