// rf_big.cuh -- strided-axis passes for axes of 2048 and 4096 points (sm_100a, float32).
//
// The 64 KB tile ring of rf_pipe.cuh does not stretch to these lengths: a [N][TZ] tile with
// 64-byte rows is 128 KB at N = 2048 (TZ = 8) and, with 32-byte rows, 128 KB at N = 4096
// (TZ = 4).  Here one CTA per SM owns ONE padded buffer and runs a three-step FFT in place:
//
//   step A  R1-point DFTs over stride-M samples (M = N/R1).  An item reads rows j + M*m and
//           writes rows k1*M + j: the same set of rows, so the step is in place per item.
//   step B  R2-point DFTs inside each k1 block (stride M2 = M/R2), in place per item too.
//   step C  R3-point DFTs over contiguous rows; output k1 + R1*(k2 + R2*k3) is a permutation,
//           so every thread reads its items, all threads synchronise, and the results go
//           back into the same memory in the dense [N][TZ] layout the TMA store reads.
//
// The threads of a CTA hold the whole tile in registers at each step (32 complex values per
// thread, at most 128 registers).  Intermediate layout: one spare row after every M rows, which makes
// all three steps bank-conflict free (lanes run over 4 or 8 consecutive rows of 64 or 32
// bytes in steps A/B and in the dense store; in step C the lanes' k1 blocks start M+1 rows
// apart).  Twiddles W_N^k come from a shared-memory copy of the table.
//
//   k1b_generate_xfft : modes generated in registers -> x FFT -> TMA store      (4 B/cell)
//   k2b_strided_fft   : TMA load -> y FFT -> TMA store                          (8 B/cell)
#pragma once
#include "rf_pipe.cuh"

namespace rf {
namespace big {

using pipe::K1PMaps;
using pipe::K1PParams;
using pipe::K2PParams;

// NT threads per CTA, CTAS resident CTAs per SM (NT * CTAS == 512: the whole register file at
// 128 registers per thread).  N = 2048 runs two independent 64 KB tiles per SM so that one
// CTA's loads and stores overlap the other's butterflies; at N = 4096 a tile of 32-byte rows
// is already 128 KB.
template <int N>
struct BPlan;
#ifndef RF_BIG2048_TZ
#define RF_BIG2048_TZ 8
#endif
#ifndef RF_BIG2048_NT
#define RF_BIG2048_NT (64 * RF_BIG2048_TZ)
#endif
template <>
struct BPlan<2048> {
  static constexpr int R1 = 16, R2 = 16, R3 = 8, TZ = RF_BIG2048_TZ, NT = RF_BIG2048_NT,
                       CTAS = 8 / RF_BIG2048_TZ;
};
template <>
struct BPlan<4096> {
  static constexpr int R1 = 16, R2 = 16, R3 = 16, TZ = 4, NT = 512, CTAS = 1;
};

#ifndef RF_K2B_AUXM_2048
#define RF_K2B_AUXM_2048 (RF_BIG2048_TZ == 8 ? 10 : 6)
#endif
#ifndef RF_K2B_AUXM_4096
#define RF_K2B_AUXM_4096 8
#endif
template <int N>
struct BGeom {
  using P = BPlan<N>;
  static constexpr int R1 = P::R1, R2 = P::R2, R3 = P::R3, TZ = P::TZ, kThreads = P::NT;
  static constexpr int M = N / R1, M2 = M / R2;
  static_assert(M2 == R3, "three-step factorisation");
  static constexpr int kItemsA = M * TZ, kItemsB = R1 * M2 * TZ, kItemsC = R1 * R2 * TZ;
  static constexpr int IA = kItemsA / kThreads, IB = kItemsB / kThreads, IC = kItemsC / kThreads;
  static_assert(IA * kThreads == kItemsA && IB * kThreads == kItemsB && IC * kThreads == kItemsC, "items");
  static constexpr int kDenseBytes = N * TZ * 8;
  static constexpr int kBufBytes = (N + R1) * TZ * 8;   // one spare row per k1 block
  static constexpr int kBoxRows = 256;
  // element r (< M) of block k1, column c, in the padded intermediate layout
  __device__ static __forceinline__ int pidx(int k1, int r, int c) { return (k1 * (M + 1) + r) * TZ + c; }
  // twiddle tables in shared memory, laid out so that the lanes of a warp (consecutive j, or
  // consecutive jp) read consecutive or odd-strided words:
  //   twA[k1 * M + j]          = W_N^(j * k1)        (step A)
  //   twB[jp * (R2 + 1) + k2]  = W_N^(R1 * jp * k2)  (step B)
  static constexpr int kTwA = N, kTwB = M2 * (R2 + 1);
  static constexpr int kTwBytes = (kTwA + kTwB) * 8;
  // k2b: row blocks [0, kAuxM) of the next tile are prefetched into an auxiliary buffer
  static constexpr int kAuxM = N == 2048 ? RF_K2B_AUXM_2048 : N == 4096 ? RF_K2B_AUXM_4096 : 6;
  static constexpr int kAuxBytes = kAuxM * M * TZ * 8;
};

template <int N>
__device__ __forceinline__ void fill_twiddles(cpx<float>* __restrict__ twA, const cpx<float>* __restrict__ tw,
                                              int tid) {
  using G = BGeom<N>;
  cpx<float>* twB = twA + G::kTwA;
  for (int q = tid; q < N; q += G::kThreads) {
    const int k1 = q / G::M, j = q % G::M;
    twA[q] = ldg_cpx(tw + j * k1);
  }
  for (int q = tid; q < G::M2 * G::R2; q += G::kThreads) {
    const int jp = q / G::R2, k2 = q % G::R2;
    twB[jp * (G::R2 + 1) + k2] = ldg_cpx(tw + G::R1 * jp * k2);
  }
}

// step A tail: twiddle by W_N^(j*k1) and store into the padded layout
template <int N>
__device__ __forceinline__ void step_a_store(cpx<float>* __restrict__ buf, const cpx<float>* __restrict__ tws,
                                             cpx<float> (&v)[BPlan<N>::R1], int j, int c) {
  using G = BGeom<N>;
  cpx<float>* wp = buf + G::pidx(0, j, c);
  sfor<0, G::R1>([&](auto ki) {
    constexpr int k1 = decltype(ki)::value;
    cpx<float> x = v[brev<G::R1>(k1)];
    if constexpr (k1 > 0) x = cmul(x, ld_cpx(tws + k1 * G::M + j));
    st_cpx(wp + (G::pidx(k1, 0, 0) - G::pidx(0, 0, 0)), x);
  });
}

// step B on a tile whose step-A output sits in the padded layout: in place per item
template <int N>
__device__ __forceinline__ void step_b(cpx<float>* __restrict__ buf, const cpx<float>* __restrict__ tws, int tid) {
  using G = BGeom<N>;
  constexpr int R2 = G::R2, TZ = G::TZ, M2 = G::M2, kThreads = G::kThreads;
#pragma unroll
  for (int q = 0; q < G::IB; ++q) {
    const int w = tid + q * kThreads;
    const int c = w % TZ, qq = w / TZ;
    const int jp = qq % M2, k1 = qq / M2;
    cpx<float>* wp = buf + G::pidx(k1, jp, c);
    cpx<float> v[R2];
    sfor<0, R2>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      v[m] = ld_cpx(wp + M2 * m * TZ);
    });
    fft_dif<R2, true>(v);
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      cpx<float> x = v[brev<R2>(k2)];
      if constexpr (k2 > 0) x = cmul(x, ld_cpx(tws + G::kTwA + jp * (R2 + 1) + k2));
      st_cpx(wp + M2 * k2 * TZ, x);
    });
  }
}
// step C, first half: every thread pulls its R3-point items out of the padded layout
template <int N>
__device__ __forceinline__ void step_c_read(const cpx<float>* __restrict__ buf, int tid,
                                            cpx<float> (&u)[BGeom<N>::IC][BPlan<N>::R3]) {
  using G = BGeom<N>;
  constexpr int R1 = G::R1, R3 = G::R3, TZ = G::TZ, M2 = G::M2, kThreads = G::kThreads;
  sfor<0, G::IC>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kThreads;
    const int c = w % TZ, qq = w / TZ;
    const int k1 = qq % R1, k2 = qq / R1;
    const cpx<float>* wp = buf + G::pidx(k1, k2 * M2, c);
    sfor<0, R3>([&](auto ji) {
      constexpr int jp = decltype(ji)::value;
      u[q][jp] = ld_cpx(wp + jp * TZ);
    });
  });
}
// steps B and C, leaving the dense [N][TZ] result at the start of the buffer (output
// k1 + R1*(k2 + R2*k3) is a permutation: read everything, synchronise, write).  Ends with
// every thread's writes issued; the caller fences and synchronises before the TMA store.
template <int N>
__device__ __forceinline__ void steps_bc(cpx<float>* __restrict__ buf, const cpx<float>* __restrict__ tws, int tid) {
  using G = BGeom<N>;
  constexpr int R1 = G::R1, R2 = G::R2, R3 = G::R3, TZ = G::TZ, kThreads = G::kThreads;
  step_b<N>(buf, tws, tid);
  __syncthreads();
  cpx<float> u[G::IC][R3];
  step_c_read<N>(buf, tid, u);
  __syncthreads();
  sfor<0, G::IC>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kThreads;
    const int c = w % TZ, qq = w / TZ;
    const int k1 = qq % R1, k2 = qq / R1;
    fft_dif<R3, true>(u[q]);
    cpx<float>* op = buf + (k1 + R1 * k2) * TZ + c;
    sfor<0, R3>([&](auto ki) {
      constexpr int k3 = decltype(ki)::value;
      st_cpx(op + R1 * R2 * k3 * TZ, u[q][brev<R3>(k3)]);
    });
  });
}

template <int N>
__host__ __device__ constexpr int kb_smem_bytes(int extra = 0) {
  return (BGeom<N>::kBufBytes + BGeom<N>::kTwBytes + 64 + extra + 127) / 128 * 128;
}
// k2b keeps the first AUXM of the R1 row blocks of the NEXT tile in a second buffer
template <int N>
constexpr int k2b_smem_bytes() {
  return kb_smem_bytes<N>() + BGeom<N>::kAuxBytes;
}

// ----------------------------------------------------------------------------------------
// k2b: inverse FFT along a strided axis of N = 2048 / 4096 rows.  Same tile walk, load map and
// parameters as pipe::k2p_strided_fft (box_rows rows per TMA box).
//
// A 128 KB tile leaves no room for a second one, so the copy/compute overlap is built from
// what is left: rows [0, AUXM*M) of the next tile stream into an auxiliary buffer while the
// current tile is transformed (step A reads them from there), and rows [AUXM*M, N) go into
// the main buffer as soon as step C has pulled the current tile into registers.  The results
// leave the registers with plain 8-byte stores (each warp writes 4 or 8 full 64/32-byte row
// segments), so nothing of the current tile has to stay in shared memory after step C's
// reads and the second half of the load overlaps the last butterflies.
// ----------------------------------------------------------------------------------------
template <int N>
__global__ void __launch_bounds__(BPlan<N>::NT, BPlan<N>::CTAS)
k2b_strided_fft(const __grid_constant__ CUtensorMap map_in, const K2PParams p) {
  using G = BGeom<N>;
  constexpr int R1 = G::R1, R2 = G::R2, R3 = G::R3, TZ = G::TZ, M = G::M, M2 = G::M2, kThreads = G::kThreads;
  constexpr int AUXM = G::kAuxM, kSplit = AUXM * M;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const buf = reinterpret_cast<cpx<float>*>(smem_raw);
  cpx<float>* const tws = reinterpret_cast<cpx<float>*>(smem_raw + G::kBufBytes);
  uint64_t* const full = reinterpret_cast<uint64_t*>(smem_raw + G::kBufBytes + G::kTwBytes);   // [0] aux, [1] main
  cpx<float>* const aux = reinterpret_cast<cpx<float>*>(smem_raw + kb_smem_bytes<N>());   // 128-byte aligned
  const int tid = threadIdx.x;
  if (tid == 0) {
    pipe::mbar_init(&full[0], 1);
    pipe::mbar_init(&full[1], 1);
    pipe::fence_barrier_init();
    pipe::tma_prefetch_desc(&map_in);
  }
  fill_twiddles<N>(tws, p.tw, tid);
  __syncthreads();
  const int first = blockIdx.x, step = gridDim.x;
  // rows [lo, hi) of tile t -> dst (dense position of row r: dst + r*TZ), completion on bar
  auto load_rows = [&](int t, int lo, int hi, cpx<float>* dst, uint64_t* bar) {
    int ck, o, z0;
    p.decode(t, ck, o, z0, TZ);
    pipe::mbar_expect_tx(bar, (uint32_t)((hi - lo) * TZ * 8));
    for (int r0 = lo; r0 < hi; r0 += p.box_rows) {
      const int src = r0 / p.rows_per_src, rl = r0 - src * p.rows_per_src;
      pipe::tma_load_4d(dst + r0 * TZ, &map_in, bar, 2 * z0, rl, o, ck * p.nsrc + src);
    }
  };
  if (tid == 0 && first < p.ntiles) {
    load_rows(first, 0, kSplit, aux, &full[0]);
    load_rows(first, kSplit, N, buf, &full[1]);
  }
  int i = 0;
  for (int t = first; t < p.ntiles; t += step, ++i) {
    pipe::mbar_wait(&full[0], i & 1);
    pipe::mbar_wait(&full[1], i & 1);
    // ---- step A: dense rows -> registers ----
    cpx<float> v[G::IA][R1];
    sfor<0, G::IA>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int w = tid + q * kThreads;
      sfor<0, R1>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[q][m] = ld_cpx((m < AUXM ? aux : buf) + w + M * m * TZ);
      });
    });
    __syncthreads();
    if (tid == 0 && t + step < p.ntiles) load_rows(t + step, 0, kSplit, aux, &full[0]);
    sfor<0, G::IA>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int w = tid + q * kThreads;
      fft_dif<R1, true>(v[q]);
      step_a_store<N>(buf, tws, v[q], w / TZ, w % TZ);
    });
    __syncthreads();
    step_b<N>(buf, tws, tid);
    __syncthreads();
    // ---- step C: registers <- padded layout; the main buffer is free after the barrier ----
    cpx<float> u[G::IC][R3];
    step_c_read<N>(buf, tid, u);
    __syncthreads();
    if (tid == 0 && t + step < p.ntiles) load_rows(t + step, kSplit, N, buf, &full[1]);
    int ck, oo, zz;
    p.decode(t, ck, oo, zz, TZ);
    cpx<float>* const outp = p.out + (size_t)oo * p.out_outer + zz + p.z_out_base + ck * p.ck_wc;
    sfor<0, G::IC>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int w = tid + q * kThreads;
      const int c = w % TZ, qq = w / TZ;
      const int k1 = qq % R1, k2 = qq / R1;
      fft_dif<R3, true>(u[q]);
      cpx<float>* op = outp + (size_t)(k1 + R1 * k2) * p.out_row + c;
      sfor<0, R3>([&](auto ki) {
        constexpr int k3 = decltype(ki)::value;
        st_cpx(op + (size_t)(R1 * R2 * k3) * p.out_row, u[q][brev<R3>(k3)]);
      });
    });
  }
  (void)M2;
}

// ----------------------------------------------------------------------------------------
// k1b: pass 1 for NX = 2048 / 4096.  Tile = [NX rows][TZ columns] of one local y index, modes
// generated in registers exactly as in pipe::k1q_generate_xfft (same Philox counters, same
// sigma cells), stores through the per-destination tensor maps of K1PMaps.
// ----------------------------------------------------------------------------------------
template <int NX, bool INJ, bool POT>
__global__ void __launch_bounds__(BPlan<NX>::NT, BPlan<NX>::CTAS)
k1b_generate_xfft(const __grid_constant__ K1PMaps maps, const K1PParams p) {
  using G = BGeom<NX>;
  constexpr int R1 = G::R1, TZ = G::TZ, M = G::M, kThreads = G::kThreads;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const buf = reinterpret_cast<cpx<float>*>(smem_raw);
  cpx<float>* const tws = reinterpret_cast<cpx<float>*>(smem_raw + G::kBufBytes);
  float4* const hinge_s = reinterpret_cast<float4*>(smem_raw + G::kBufBytes + G::kTwBytes + 64);
  const int tid = threadIdx.x;
  const int first = blockIdx.x, step = gridDim.x;
  if (tid == 0) pipe::tma_prefetch_desc(&maps.m[0]);
  fill_twiddles<NX>(tws, p.tw, tid);
  if constexpr (!INJ) {
    for (int q = tid; q < p.tab.lut_n; q += kThreads)
      hinge_s[q] = __ldg(reinterpret_cast<const float4*>(p.tab.hinge) + q);
  }
  __syncthreads();
  const float4* const hinge0 = hinge_s - p.tab.lut_base;
  const int shift = p.tab.shift;
  const ModeGeom<float>& g = p.geom;
  const int nzh = g.nz / 2 + 1;
  const uint64_t dq = (uint64_t)M * g.ny * nzh;   // Philox counter step between m and m+1

  for (int t = first; t < p.ntiles; t += step) {
    const int iyl = t / p.ztiles, z0 = (t % p.ztiles) * TZ;
    const int iy = p.y0 + iyl;
    cpx<float> v[G::IA][R1];
    sfor<0, G::IA>([&](auto qi) {
      constexpr int qn = decltype(qi)::value;
      const int w = tid + qn * kThreads;
      const int c = w % TZ, j = w / TZ;
      const int iz = p.z_base + z0 + c;
      if (iz == 0) {
        const cpx<float>* pp = p.plane + (size_t)j * p.ny_local + iyl;
        sfor<0, R1>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          v[qn][m] = ldg_cpx(pp + (size_t)(M * m) * p.ny_local);
        });
      } else if constexpr (INJ) {
        sfor<0, R1>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          v[qn][m] = raw_mode(g, p.tab, j + M * m, iy, iz);
        });
      } else {
        const float fj = (float)j;
        const float ksq_yz = ksq_canon(0.f, mode_sq<float>(iy, g.ny), mode_sq<float>(iz, g.nz), g.k0x2,
                                       g.k0y2, g.k0z2);
        const uint64_t q0 = ((uint64_t)j * g.ny + iy) * nzh + iz;
        sfor<0, R1 / 2>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          const float f1 = fj + (float)(M * m), f2 = f1 - (float)(NX / 2);
          const float ks1 = __fmaf_rn(f1 * f1, g.k0x2, ksq_yz);
          const float ks2 = __fmaf_rn(f2 * f2, g.k0x2, ksq_yz);
          float s1 = sigma_hinge_smem(hinge0, shift, ks1), s2 = sigma_hinge_smem(hinge0, shift, ks2);
          if constexpr (POT) {
            s1 = __fdividef(s1, ks1);
            s2 = __fdividef(s2, ks2);
          }
          const uint64_t q = q0 + (uint64_t)m * dq;
          philox_out o = philox4x32_10((uint32_t)q, (uint32_t)(q >> 32), 0u, 0u, g.key0, g.key1);
          cpx<float> n1 = box_muller<float>(o.w[0], o.w[1]);
          cpx<float> n2 = box_muller<float>(o.w[2], o.w[3]);
          v[qn][m] = pk_mul(n1.x, n1.y, s1, s1);
          v[qn][m + R1 / 2] = pk_mul(n2.x, n2.y, s2, s2);
        });
      }
      fft_dif<R1, true>(v[qn]);
    });
    // the previous tile's TMA store must have finished reading the buffer
    if (tid == 0) pipe::bulk_wait_read0();
    __syncthreads();
    sfor<0, G::IA>([&](auto qi) {
      constexpr int qn = decltype(qi)::value;
      const int w = tid + qn * kThreads;
      step_a_store<NX>(buf, tws, v[qn], w / TZ, w % TZ);
    });
    __syncthreads();
    steps_bc<NX>(buf, tws, tid);
    pipe::fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
      const int box = p.nxl < G::kBoxRows ? p.nxl : G::kBoxRows;
      for (int r0 = 0; r0 < NX; r0 += box) {
        const int d = r0 / p.nxl;
        pipe::tma_store_3d(&maps.m[d], buf + r0 * TZ, 2 * z0, iyl, r0 - d * p.nxl);
      }
      pipe::bulk_commit();
    }
  }
  if (tid == 0) pipe::bulk_wait0();
}

}  // namespace big
}  // namespace rf
