// rf_fuse12.cuh -- pass 1 (generate + x FFT) and pass 2 (y FFT) in ONE persistent kernel.
//
// Pass 1 is issue-bound (Philox, Box-Muller, sigma(k), butterflies: HBM 30 % busy) and pass 2
// is HBM-bound (issue slots 40 % busy).  Run back to back they cost the sum; here they share
// every SM and cost little more than the longer of the two.  One CTA of 512 threads per SM:
//
//   warps 0-7  (group X): the k1q pass-1 loop -- modes of a [NX][TZ] tile generated in
//              registers, x FFT through one padded buffer, TMA store.  Tiles are walked in
//              chunks of zc z-tiles: all (y, z-tile) of chunk 0, then chunk 1, ...
//   warps 8-15 (group Y): the pass-2 loop on tiles [NY][TZ] of one x, z-tile major as well.
//              A tile of z-tile zt is loaded once all ny pass-1 tiles of zt have been stored
//              (one counter per z-tile: release-increment by group X after the store has
//              completed, acquire-poll by the thread that issues the loads).  TMA load into a
//              dense buffer, step A into the padded buffer, step B in place, TMA store; the
//              next tile's load is issued as soon as step A has read the dense buffer.
//
// Group X never waits for group Y and every CTA is resident (grid <= SM count, one CTA per
// SM), so the z-tile counters always complete: no deadlock.  The two groups synchronise among
// themselves with named barriers 1 and 2.  Single GPU, nx == ny <= 1024, Philox path (the
// injected-noise path and slab-decomposed runs use the separate kernels).
#pragma once
#include "rf_big.cuh"

namespace rf {
namespace fuse12 {

using pipe::kConsumers;
using pipe::SPlan;
using pipe::LayoutStridedPad;

constexpr int kThreads = 2 * kConsumers;

struct XSync {
  __device__ __forceinline__ void operator()() const { asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory"); }
};
struct YSync {
  __device__ __forceinline__ void operator()() const { asm volatile("bar.sync 2, %0;" ::"n"(kConsumers) : "memory"); }
};

// mbarrier wait that lets the hardware park the warp for up to ~1 us per probe: group Y often
// waits for group X, and a tight try_wait loop would compete with it for issue slots
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "RF_WAITR_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
      "@p bra RF_DONER_%=;\n"
      "bra RF_WAITR_%=;\n"
      "RF_DONER_%=:\n"
      "}\n" ::"r"(pipe::smem_u32(bar)),
      "r"(parity), "r"(1000u)
      : "memory");
}

struct K12Params {
  pipe::K1PParams k1;       // ztiles, ntiles as in k1p; tiles are walked z-tile major here
  const cpx<float>* tw_y;   // exp(+2 pi i k / ny)
  int nplanes;              // x planes of pass 2 (== nx)
  int hinge_first;          // first staged cell of the sigma table (index into tab.hinge)
  int hinge_cells;          // number of staged cells
  cpx<float>* out;          // pass-2 output: out[x * out_outer + y * out_row + z]
  long long out_outer, out_row;
  int dbg;                  // timing experiments: 1 = group Y idle, 2 = group X idle (counters preset)
  int zc;                   // z-tiles per chunk (power of two, divides ztiles)
  unsigned int* zt_done;    // [ztiles / zc] zeroed before launch: pass-1 tiles stored per chunk
};

// shared-memory carve-up (bytes): dense_y | work_y | work_x | tw_y | tw_x | barriers | sigma cells
template <int N>
struct Smem {
  using P = SPlan<N>;
  static constexpr int kDenseY = 0;
  static constexpr int kWorkY = P::kDenseBytes;
  static constexpr int kWorkX = kWorkY + (big::BGeom<N>::kBufBytes + 127) / 128 * 128;
  static constexpr int kTwY = kWorkX + P::kWorkBytes;
  static constexpr int kTwX = kTwY + (big::BGeom<N>::kTwBytes + 15) / 16 * 16;
  static constexpr int kBars = kTwX + N * 8;
  static constexpr int kHinge = kBars + 64;
  static_assert(kWorkY % 128 == 0 && kWorkX % 128 == 0, "TMA buffers need 128-byte alignment");
};
template <int N>
constexpr int k12_smem_bytes(int hinge_cells) {
  return Smem<N>::kHinge + hinge_cells * 16;
}

template <int N, bool POT>
__global__ void __launch_bounds__(kThreads, 1)
k12_generate_xyfft(const __grid_constant__ pipe::K1PMaps maps, const __grid_constant__ CUtensorMap map_in,
                   const K12Params p) {
  using P = SPlan<N>;
  using L = LayoutStridedPad<P::TZ, P::M>;
  using S = Smem<N>;
  constexpr int R1 = P::R1, R2 = P::R2, M = P::M, TZ = P::TZ;
  static_assert(M * TZ == kConsumers, "one step-A item per thread");
  static_assert(R1 % R2 == 0, "step B items per thread");
  constexpr int IB = R1 / R2;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const dense_y = reinterpret_cast<cpx<float>*>(smem_raw + S::kDenseY);
  cpx<float>* const work_y = reinterpret_cast<cpx<float>*>(smem_raw + S::kWorkY);
  cpx<float>* const work_x = reinterpret_cast<cpx<float>*>(smem_raw + S::kWorkX);
  cpx<float>* const tws_y = reinterpret_cast<cpx<float>*>(smem_raw + S::kTwY);
  cpx<float>* const tws_x = reinterpret_cast<cpx<float>*>(smem_raw + S::kTwX);
  uint64_t* const full = reinterpret_cast<uint64_t*>(smem_raw + S::kBars);
  float4* const hinge_s = reinterpret_cast<float4*>(smem_raw + S::kHinge);
  const int tid = threadIdx.x;
  const int first = blockIdx.x, step = gridDim.x;
  const pipe::K1PParams& k1 = p.k1;

  if (tid == 0) {
    pipe::mbar_init(full, 1);
    pipe::fence_barrier_init();
    pipe::tma_prefetch_desc(&maps.m[0]);
    pipe::tma_prefetch_desc(&map_in);
  }
  // twiddle tables, [k1 * M + j] = W_N^(j * k1): lanes (consecutive j) read consecutive words
  for (int q = tid; q < N; q += kThreads) {
    tws_x[q] = ldg_cpx(k1.tw + (q % M) * (q / M));
    tws_y[q] = ldg_cpx(p.tw_y + (q % big::BGeom<N>::M) * (q / big::BGeom<N>::M));
  }
  for (int q = tid; q < big::BGeom<N>::M2 * big::BGeom<N>::R2; q += kThreads) {
    const int jp = q / big::BGeom<N>::R2, k2 = q % big::BGeom<N>::R2;
    tws_y[big::BGeom<N>::kTwA + jp * (big::BGeom<N>::R2 + 1) + k2] = ldg_cpx(p.tw_y + big::BGeom<N>::R1 * jp * k2);
  }
  for (int q = tid; q < p.hinge_cells; q += kThreads)
    hinge_s[q] = __ldg(reinterpret_cast<const float4*>(k1.tab.hinge) + p.hinge_first + q);
  __syncthreads();

  if (tid < kConsumers) {
    // =================== group X: pass 1, chunk major ===================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 168;");
    const float4* const hinge0 = hinge_s - (k1.tab.lut_base + p.hinge_first);
    const int shift = k1.tab.shift;
    const ModeGeom<float>& g = k1.geom;
    const int c = tid % TZ, j = tid / TZ;
    const int nzh = g.nz / 2 + 1;
    const float fj = (float)j;
    const uint64_t dq = (uint64_t)M * g.ny * nzh;
    cpx<float>* const wa = work_x + L::idx2(0, j, c);
    int prev_zt = -1, pend0 = -1, pend1 = -1;   // chunks of the last three stores, oldest first
    auto signal = [&](int zt) {
      if (zt >= 0) {
        asm volatile("fence.proxy.async;" ::: "memory");
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.zt_done + zt) : "memory");
      }
    };
    for (int t = first; t < ((p.dbg & 2) ? 0 : k1.ntiles); t += step) {
      // chunk-major; inside a chunk z-tiles run fastest, so the CTAs of one wave store runs of
      // zc * 64 contiguous bytes per (x, y) instead of isolated 64-byte pieces
      const int per_chunk = k1.ny_local * p.zc;
      const int chunk = t / per_chunk, r = t - chunk * per_chunk;
      const int iyl = r / p.zc, zt = chunk * p.zc + (r - iyl * p.zc);
      const int z0 = zt * TZ;
      const int iy = k1.y0 + iyl;
      const int iz = k1.z_base + z0 + c;
      cpx<float> v[R1];
      if (iz == 0) {
        const cpx<float>* pp = k1.plane + (size_t)j * k1.ny_local + iyl;
        sfor<0, R1>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          v[m] = ldg_cpx(pp + (size_t)(M * m) * k1.ny_local);
        });
      } else {
        const float ksq_yz = ksq_canon(0.f, mode_sq<float>(iy, g.ny), mode_sq<float>(iz, g.nz), g.k0x2,
                                       g.k0y2, g.k0z2);
        const uint64_t q0 = ((uint64_t)j * g.ny + iy) * nzh + iz;
        sfor<0, R1 / 2>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          const float f1 = fj + (float)(M * m), f2 = f1 - (float)(N / 2);
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
          v[m] = pk_mul(n1.x, n1.y, s1, s1);
          v[m + R1 / 2] = pk_mul(n2.x, n2.y, s2, s2);
        });
      }
      fft_dif<R1, true>(v);
      // work_x must have been read out by the previous tile's store before it is rewritten
      if (tid == 0) pipe::bulk_wait_read0();
      XSync()();
      sfor<0, R1>([&](auto ki) {
        constexpr int kk = decltype(ki)::value;
        cpx<float> x = v[brev<R1>(kk)];
        if constexpr (kk > 0) x = cmul(x, ld_cpx(tws_x + kk * M + j));
        st_cpx(wa + (L::idx2(kk, 0, 0) - L::idx2(0, 0, 0)), x);
      });
      XSync()();
      cpx<float> u[IB][R2];
      sfor<0, IB>([&](auto qi) {
        constexpr int q = decltype(qi)::value;
        const int w = tid + q * kConsumers;
        const cpx<float>* wp = work_x + L::idx2(w / TZ, 0, w % TZ);
        sfor<0, R2>([&](auto ji) {
          constexpr int jj = decltype(ji)::value;
          u[q][jj] = ld_cpx(wp + (L::idx2(0, jj, 0) - L::idx2(0, 0, 0)));
        });
      });
      XSync()();
      sfor<0, IB>([&](auto qi) {
        constexpr int q = decltype(qi)::value;
        const int w = tid + q * kConsumers;
        const int kk = w / TZ, cc = w % TZ;
        fft_dif<R2, true>(u[q]);
        sfor<0, R2>([&](auto ki) {
          constexpr int k2 = decltype(ki)::value;
          st_cpx(work_x + (kk + R1 * k2) * TZ + cc, u[q][brev<R2>(k2)]);
        });
      });
      pipe::fence_proxy_async();
      XSync()();
      if (tid == 0) {
        for (int r0 = 0; r0 < N; r0 += P::kBoxRows) pipe::tma_store_3d(&maps.m[0], work_x + r0 * TZ, 2 * z0, iyl, r0);
        pipe::bulk_commit();
        // a store is only reported to group Y two tiles later: waiting for the write of the tile
        // just issued (HBM latency under load) would stall the whole group
        asm volatile("cp.async.bulk.wait_group 2;" ::: "memory");
        signal(pend0);
        pend0 = pend1;
        pend1 = prev_zt;
        prev_zt = chunk;
      }
    }
    if (tid == 0) {
      pipe::bulk_wait0();
      signal(pend0);
      signal(pend1);
      signal(prev_zt);
    }
    return;
  }

  // =================== group Y: pass 2, chunk major ===================
  // Three-step radix-16/8/8 engine of rf_big.cuh (in place per item, at most 16 values per
  // thread in registers), results stored from registers: this group runs in 88 registers so
  // that group X can have 168.
  {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
    using G = big::BGeom<N>;
    constexpr int GR1 = G::R1, GR3 = G::R3, GM = G::M, GM2 = G::M2;
    static_assert(G::TZ == TZ && G::kThreads == kConsumers, "pass-2 plan");
    const int yt = tid - kConsumers;
    const int ytiles = (p.dbg & 1) ? 0 : k1.ztiles * p.nplanes;
    const int per_chunk = p.nplanes * p.zc;
    // load tile u = (chunk, x plane o, z-tile) once pass 1 has stored every (y, z-tile) of the chunk
    auto load_tile = [&](int u) {
      const int chunk = u / per_chunk, r = u - chunk * per_chunk;
      const int o = r / p.zc, zt = chunk * p.zc + (r - o * p.zc);
      const unsigned int need = (unsigned int)(k1.ny_local * p.zc);
      unsigned int have;
      for (;;) {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(have) : "l"(p.zt_done + chunk) : "memory");
        if (have >= need) break;
        __nanosleep(200);
      }
      asm volatile("fence.proxy.async;" ::: "memory");
      pipe::mbar_expect_tx(full, (uint32_t)P::kDenseBytes);
      for (int r0 = 0; r0 < N; r0 += P::kBoxRows)
        pipe::tma_load_4d(dense_y + r0 * TZ, &map_in, full, 2 * zt * TZ, r0, o, 0);
    };
    if (yt == 0 && first < ytiles) load_tile(first);
    int i = 0;
    for (int u = first; u < ytiles; u += step, ++i) {
      mbar_wait_relaxed(full, i & 1);
      // ---- step A: dense tile -> padded work buffer ----
#pragma unroll
      for (int q = 0; q < G::IA; ++q) {
        const int w = yt + q * kConsumers;
        cpx<float> v[GR1];
        sfor<0, GR1>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          v[m] = ld_cpx(dense_y + w + GM * m * TZ);
        });
        fft_dif<GR1, true>(v);
        big::step_a_store<N>(work_y, tws_y, v, w / TZ, w % TZ);
      }
      YSync()();
      if (yt == 0 && u + step < ytiles) load_tile(u + step);   // dense_y is free
      big::step_b<N>(work_y, tws_y, yt);
      YSync()();
      // ---- step C: R3-point DFTs, results straight to global memory ----
      {
        const int chunk = u / per_chunk, r = u - chunk * per_chunk;
        const int o = r / p.zc, zt = chunk * p.zc + (r - o * p.zc);
        cpx<float>* const outp = p.out + (size_t)o * p.out_outer + zt * TZ;
#pragma unroll
        for (int q = 0; q < G::IC; ++q) {
          const int w = yt + q * kConsumers;
          const int cc = w % TZ, qq = w / TZ;
          const int kk1 = qq % GR1, kk2 = qq / GR1;
          const cpx<float>* wp = work_y + G::pidx(kk1, kk2 * GM2, cc);
          cpx<float> v[GR3];
          sfor<0, GR3>([&](auto ji) {
            constexpr int jp = decltype(ji)::value;
            v[jp] = ld_cpx(wp + jp * TZ);
          });
          fft_dif<GR3, true>(v);
          cpx<float>* op = outp + (size_t)(kk1 + GR1 * kk2) * p.out_row + cc;
          sfor<0, GR3>([&](auto ki) {
            constexpr int k3 = decltype(ki)::value;
            st_cpx(op + (size_t)(GR1 * G::R2 * k3) * p.out_row, v[brev<GR3>(k3)]);
          });
        }
      }
      YSync()();   // work_y may be overwritten by the next tile's step A
    }
  }
}

}  // namespace fuse12
}  // namespace rf
