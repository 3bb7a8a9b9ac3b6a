// NOT BUILT -- kept as the record of a measured experiment: slower than
// k23q_fused_yz (2.60 ms against 2.36 ms at 1024^3).  To try it again: copy next to rf_k23q.cuh, include it
// from rf_launch.h and add the launcher of git show HEAD~:randomfield_b200/csrc/rf_fused_inst.cu.
//
// rf_k23h.cuh -- the fused y + z pass on HALF tiles with THREE independent FFT groups.
//
// k23q_fused_yz (rf_k23q.cuh) moves a 64 KB tile through shared memory six times (TMA in, load A,
// store A', load B, store B', TMA out: 3072 cycles of the 128 B/clk port) and spends about 2200
// FMA-pipe cycles on its butterflies -- and because its eight FFT warps walk through the phases of
// a tile together, the two add up, and with two warps per scheduler every dependent-issue and
// shared-memory latency is exposed (ncu, round 2: each FFT warp issues one instruction every six
// cycles; shared-memory port 60 %, FMA pipe 25 % busy).  Here the unit of work is HALF a tile,
//   y pass: [NY][TZ/2 columns] (columns are independent), z pass: TL/2 lines (lines are independent),
// 32 KB in a ring of SIX half-tile slots, and twelve FFT warps form three groups of four that each
// take every third half tile and run free of each other: three warps per scheduler, in different
// phases, so one group's shared-memory sweeps run under another's butterflies; three slots are
// being transformed while three are in transit.  The step twiddles are read from the shared-memory
// tables right where they are used (no 64-register twiddle arrays), which is what lets 448
// threads fit the register file.  Same arithmetic on the same values in the same order as
// k2p / k3q / k23q: the field is bit-identical.
//   bar 1       : all FFT threads (final statistics)
//   bar 2 + g   : the 128 threads of group g
// Loader / storer warps, tickets (now per half tile) and plane counters are k23q's.
#pragma once
#include "rf_k23q.cuh"

namespace rf {
namespace pipe {

constexpr int kGroupH = kConsumers / 2;   // 128 threads per FFT group

constexpr int kGroupsH = 3;
constexpr int kSlotsH = 6;
constexpr int kFftThreadsH = kGroupsH * kGroupH;      // 384
constexpr int kThreadsH = kFftThreadsH + 64;          // + loader and storer warps

template <int NY, int NZ>
struct K23H {
  using P2 = SPlan<NY>;
  using P3 = LPlan<NZ>;
  static constexpr int TZh = P2::TZ / 2;                 // columns of a y-pass half tile
  static constexpr int TEh = P3::TE / 2;                 // lines of a z-pass half tile
  static constexpr int kHalf2 = (NY + P2::R1) * TZh * 8;   // padded y-pass half tile
  static constexpr int kHalf3 = TEh * P3::WP * 8;          // padded z-pass half tile
  static constexpr int kHalf = (((kHalf2 > kHalf3 ? kHalf2 : kHalf3) + 127) / 128) * 128;
  static constexpr int kDense2 = NY * TZh * 8, kDense3 = TEh * P3::H * 8;   // bytes moved per half tile
  static constexpr int kSmem = kSlotsH * kHalf + 4 * kBarBytes + (NY + NZ) * 8;
  static constexpr bool kOk = K3Fuse<NZ>::value && P2::TZ >= 8 && (P2::M * TZh == kGroupH) &&
                              (P3::M / 2) * TEh == kGroupH && kSmem <= 227 * 1024;
};

struct GroupSync {
  int id;
  __device__ __forceinline__ void operator()() const { named_sync(id, kGroupH); }
};
// one y-pass half tile, in place in `B` (dense [N][TZh] at the start of a kHalf buffer); t in [0, 128)
template <int N>
__device__ __forceinline__ void k2h_tile_body(cpx<float>* __restrict__ B, int t, const cpx<float>* tws,
                                              const GroupSync gsync) {
  using P = SPlan<N>;
  constexpr int R1 = P::R1, R2 = P::R2, M = P::M, TZh = P::TZ / 2, IB = R1 / R2;
  using L = LayoutStridedPad<TZh, M>;
  static_assert(M * TZh == kGroupH && R1 % R2 == 0 && R1 * R2 == N, "plan");
  const int c = t % TZh, j = t / TZh;
  cpx<float> v[R1];
  {
    const cpx<float>* Dp = B + j * TZh + c;
    sfor<0, R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      v[m] = ld_cpx(Dp + M * m * TZh);
    });
  }
  gsync();
  fft_dif<R1, true>(v);
  {
    cpx<float>* wa = B + L::idx2(0, j, c);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x = v[brev<R1>(k1)];
      if constexpr (k1 > 0) x = cmul(x, ld_cpx(tws + k1 * M + j));
      st_cpx(wa + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x);
    });
  }
  gsync();
  cpx<float> u[IB][R2];
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = t + q * kGroupH;
    const cpx<float>* wp = B + L::idx2(w / TZh, 0, w % TZh);
    sfor<0, R2>([&](auto ji) {
      constexpr int jj = decltype(ji)::value;
      u[q][jj] = ld_cpx(wp + (L::idx2(0, jj, 0) - L::idx2(0, 0, 0)));
    });
  });
  gsync();
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    fft_dif<R2, true>(u[q]);
  });
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = t + q * kGroupH;
    const int k1 = w / TZh, cc = w % TZh;
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      st_cpx(B + (k1 + R1 * k2) * TZh + cc, u[q][brev<R2>(k2)]);
    });
  });
}

// one z-pass half tile (TE/2 lines), in place in `B` (dense [TEh][H] at the start of a kHalf buffer);
// the fused-unpack form of k3q_tile_body (K3Fuse) for 128 threads; t in [0, 128)
template <int NZ>
__device__ __forceinline__ void k3h_tile_body(cpx<float>* __restrict__ B, int t, const cpx<float>* twHt,
                                              const cpx<float>* twN, const float* zs, cpx<float>& a1,
                                              cpx<float>& a2, const GroupSync gsync) {
  using P = LPlan<NZ>;
  constexpr int H = P::H, HE = P::HE, R1 = P::R1, R2 = HE / R1, M = P::M, TEh = P::TE / 2;
  using L = LayoutLinesPad<P::WP, P::M>;
  constexpr int PH = M / 2, KS = K3Fuse<NZ>::kStep;
  constexpr int IB = R1 * TEh / kGroupH;
  static_assert(K3Fuse<NZ>::value && PH * TEh == kGroupH && IB * kGroupH == R1 * TEh && IB * R2 == 32, "plan");
  // ---- step A: dense -> registers (+ Hermitian unpack) | barrier | registers -> padded ----
  const int pp = t % PH, l = t / PH;
  const int j0 = pp, j1 = pp == 0 ? PH : M - pp;
  const cpx<float>* Lp = B + l * HE;
  cpx<float> v0[R1], v1[R1];
  sfor<0, R1>([&](auto mi) {
    constexpr int m = decltype(mi)::value;
    v0[m] = ld_cpx(Lp + j0 + M * m);
    v1[m] = ld_cpx(Lp + j1 + M * m);
  });
  const cpx<float> wj = ld_cpx(twN + (pp == 0 ? PH : pp));
  gsync();
  if (pp != 0) {
    sfor<0, R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      const cpx<float> a = v0[m], c = v1[R1 - 1 - m];
      const cpx<float> e = pk_add(a.x, a.y, c.x, -c.y);
      const cpx<float> d = pk_add(a.x, a.y, -c.x, c.y);
      const cpx<float> o = mul_tw32<KS * m, true>(cmul(d, wj));
      v0[m] = pk_add(e.x, e.y, -o.y, o.x);
      v1[R1 - 1 - m] = pk_add(e.x, -e.y, o.y, o.x);
    });
  } else {
    {
      cpx<float> z0, zm;
      c2r_unpack_ends(v0[0], v0[R1 / 2], z0, zm);
      v0[0] = z0;
      v0[R1 / 2] = zm;
    }
    sfor<1, R1 / 2>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      const cpx<float> a = v0[m], c = v0[R1 - m];
      const cpx<float> e = pk_add(a.x, a.y, c.x, -c.y);
      const cpx<float> d = pk_add(a.x, a.y, -c.x, c.y);
      const cpx<float> o = mul_tw32<KS * m, true>(d);
      v0[m] = pk_add(e.x, e.y, -o.y, o.x);
      v0[R1 - m] = pk_add(e.x, -e.y, o.y, o.x);
    });
    sfor<0, R1 / 2>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      const cpx<float> a = v1[m], c = v1[R1 - 1 - m];
      const cpx<float> e = pk_add(a.x, a.y, c.x, -c.y);
      const cpx<float> d = pk_add(a.x, a.y, -c.x, c.y);
      const cpx<float> o = mul_tw32<KS * m, true>(cmul(d, wj));
      v1[m] = pk_add(e.x, e.y, -o.y, o.x);
      v1[R1 - 1 - m] = pk_add(e.x, -e.y, o.y, o.x);
    });
  }
  fft_dif<R1, true>(v0);
  fft_dif<R1, true>(v1);
  {
    cpx<float>* wp0 = B + L::idx2(0, j0, l);
    cpx<float>* wp1 = B + L::idx2(0, j1, l);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x0 = v0[brev<R1>(k1)], x1 = v1[brev<R1>(k1)];
      if constexpr (k1 > 0) {
        x0 = cmul(x0, ld_cpx(twHt + k1 * M + j0));
        x1 = cmul(x1, ld_cpx(twHt + k1 * M + j1));
      }
      st_cpx(wp0 + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x0);
      st_cpx(wp1 + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x1);
    });
  }
  gsync();
  // ---- step B: padded -> registers | barrier | registers -> dense (+ statistics, z scale) ----
  cpx<float> u[IB][R2];
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = t + q * kGroupH;
    int k1, ll;
    L::split(w, R1, k1, ll);
    const cpx<float>* wp = B + L::idx2(k1, 0, ll);
    sfor<0, R2>([&](auto ji) {
      constexpr int j = decltype(ji)::value;
      u[q][j] = ld_cpx(wp + (L::idx2(0, j, 0) - L::idx2(0, 0, 0)));
    });
  });
  gsync();
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    fft_dif<R2, true>(u[q]);
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      stats_accum(u[q][brev<R2>(k2)], a1, a2);
    });
  });
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = t + q * kGroupH;
    int k1, ll;
    L::split(w, R1, k1, ll);
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      cpx<float> x = u[q][brev<R2>(k2)];
      const int k = k1 + R1 * k2;
      if (zs) {
        float2 z = __ldg(reinterpret_cast<const float2*>(zs) + k);
        x = pk_mul(x.x, x.y, z.x, z.y);
      }
      st_cpx(B + ll * H + k, x);
    });
  });
}

// map_in / map_out: the tensor maps of k23q with the box cut to TZ/2 columns.  p.t2 / p.t3 count HALF
// tiles per plane (twice k23q's), and the plane counters count half-tile stores.
template <int NY, int NZ>
__global__ void __launch_bounds__(kThreadsH, 1)
k23h_fused_yz(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_out,
              const K23Params p) {
  using P2 = SPlan<NY>;
  using P3 = LPlan<NZ>;
  using G = K23H<NY, NZ>;
  static_assert(G::kOk, "k23h: unsupported (NY, NZ)");
  constexpr int kSlotElems = G::kHalf / 8;
  constexpr int TZh = G::TZh;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const buf0 = reinterpret_cast<cpx<float>*>(smem_raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + kSlotsH * G::kHalf);
  uint64_t* full = bars;                  // [6] half tile landed (or end marker published)
  uint64_t* done = bars + kSlotsH;        // [6] the slot's FFT group finished the half tile
  uint64_t* empty = bars + 2 * kSlotsH;   // [6] the half tile's store has been read out of the slot
  volatile int* desc = reinterpret_cast<volatile int*>(bars + 3 * kSlotsH);   // [6][4]: type, plane, idx
  cpx<float>* const tw2s = reinterpret_cast<cpx<float>*>(smem_raw + kSlotsH * G::kHalf + 4 * kBarBytes);
  cpx<float>* const twHs = tw2s + NY;
  cpx<float>* const twNs = twHs + NZ / 2;
  static_assert(3 * kSlotsH * 8 + kSlotsH * 16 <= 4 * kBarBytes, "barrier area");
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int b = 0; b < kSlotsH; ++b) {
      mbar_init(&full[b], 1);
      mbar_init(&done[b], kGroupH);
      mbar_init(&empty[b], 1);
    }
    fence_barrier_init();
  }
  // step-twiddle tables in [k1][j] order (see k23q)
  for (int q = tid; q < NY; q += kThreadsH) tw2s[q] = ldg_cpx(p.k2.tw + (q % P2::M) * (q / P2::M));
  for (int q = tid; q < NZ / 2; q += kThreadsH) {
    twHs[q] = ldg_cpx(p.k3.twH + (q % P3::M) * (q / P3::M));
    twNs[q] = ldg_cpx(p.k3.twN + q);
  }
  __syncthreads();
  char* const gbase = reinterpret_cast<char*>(p.k3.buf);

  if (tid >= kFftThreadsH) {
    if (tid == kFftThreadsH) {
      // ===== loader: one ticket = one half tile =====
      tma_prefetch_desc(&map_in);
      unsigned int ticket = atomicAdd(p.ticket, 1u);
      int b = 0, ph = 0;
      for (int n = 0;; ++n) {
        const unsigned int ticket_next = atomicAdd(p.ticket, 1u);
        if (n >= kSlotsH) mbar_wait(&empty[b], ph ^ 1);   // (ph flips every kSlotsH half tiles)
        int type, plane, idx;
        k23_decode(p, ticket, type, plane, idx);
        ticket = ticket_next;
        desc[4 * b + 0] = type; desc[4 * b + 1] = plane; desc[4 * b + 2] = idx;
        cpx<float>* dst = buf0 + b * kSlotElems;
        if (type == kTileEnd) {
          // every group has to see the end: publish it in this slot and in the next two
          for (int e = 0; e < kGroupsH; ++e) {
            if (e > 0) {
              if (++b == kSlotsH) { b = 0; ph ^= 1; }
              if (n + e >= kSlotsH) mbar_wait(&empty[b], ph ^ 1);
              desc[4 * b + 0] = kTileEnd;
            }
            mbar_arrive(&full[b]);
          }
          break;
        }
        if (type == kTileK2) {
          const int z0 = idx * TZh;
          mbar_expect_tx(&full[b], (uint32_t)G::kDense2);
          for (int r0 = 0; r0 < NY; r0 += p.k2.box_rows) {
            const int src = r0 / p.k2.rows_per_src, rl = r0 - src * p.k2.rows_per_src;
            tma_load_4d(dst + r0 * TZh, &map_in, &full[b], 2 * z0, rl, plane, src);
          }
        } else {
          for (;;) {
            unsigned int have;
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(have) : "l"(p.plane_done + plane) : "memory");
            if (have >= (unsigned)p.t2) break;
            __nanosleep(200);
          }
          asm volatile("fence.proxy.async;" ::: "memory");
          const size_t off = ((size_t)plane * p.t3 + idx) * (size_t)G::kDense3;
          mbar_expect_tx(&full[b], (uint32_t)G::kDense3);
          bulk_g2s(dst, gbase + off, (uint32_t)G::kDense3, &full[b]);
        }
        if (++b == kSlotsH) { b = 0; ph ^= 1; }
      }
    } else if (tid == kFftThreadsH + 32) {
      // ===== storer =====
      // Up to kRd stores may still be reading their slots when the next one is issued: with one
      // half tile every ~1.3 us per SM, waiting for each read-out in turn (k23q's scheme) makes
      // this thread's serial chain the bottleneck.  A slot goes back to the loader kRd stores later.
      tma_prefetch_desc(&map_out);
      constexpr int kSigLag = 2 * RF_K23Q_SIGLAG;
      constexpr int kRd = 2;
      int pend[kSigLag];
      int rd[kRd];
#pragma unroll
      for (int q = 0; q < kSigLag; ++q) pend[q] = -1;
#pragma unroll
      for (int q = 0; q < kRd; ++q) rd[q] = -1;
      auto signal = [&](int plane) {
        if (plane >= 0) {
          asm volatile("fence.proxy.async;" ::: "memory");
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.plane_done + plane) : "memory");
        }
      };
      // before sitting idle: hand back every slot and signal every finished plane
      auto flush = [&]() {
        bulk_wait0();
#pragma unroll
        for (int q = 0; q < kRd; ++q) {
          if (rd[q] >= 0) mbar_arrive(&empty[rd[q]]);
          rd[q] = -1;
        }
#pragma unroll
        for (int q = 0; q < kSigLag; ++q) { signal(pend[q]); pend[q] = -1; }
      };
      int b = 0, ph = 0;
      for (;;) {
        if (!mbar_test(&full[b], ph)) {
          flush();
          mbar_wait(&full[b], ph);
        }
        const int type = desc[4 * b + 0], plane = desc[4 * b + 1], idx = desc[4 * b + 2];
        if (type == kTileEnd) break;
        if (!mbar_test(&done[b], ph)) {
          flush();
          mbar_wait(&done[b], ph);
        }
        const cpx<float>* src = buf0 + b * kSlotElems;
        if (type == kTileK2) {
          const int z0 = idx * TZh;
          for (int r0 = 0; r0 < NY; r0 += P2::kBoxRows) tma_store_3d(&map_out, src + r0 * TZh, 2 * z0, r0, plane);
        } else {
          const size_t off = ((size_t)plane * p.t3 + idx) * (size_t)G::kDense3;
          bulk_s2g(gbase + off, src, (uint32_t)G::kDense3);
        }
        bulk_commit();
        // the store issued kRd half tiles ago has been read out of its slot
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kRd) : "memory");
        if (rd[0] >= 0) mbar_arrive(&empty[rd[0]]);
#pragma unroll
        for (int q = 0; q + 1 < kRd; ++q) rd[q] = rd[q + 1];
        rd[kRd - 1] = b;
        asm volatile("cp.async.bulk.wait_group %0;" ::"n"(kSigLag) : "memory");
        signal(pend[0]);
#pragma unroll
        for (int q = 0; q + 1 < kSigLag; ++q) pend[q] = pend[q + 1];
        pend[kSigLag - 1] = type == kTileK2 ? plane : -1;
        if (++b == kSlotsH) { b = 0; ph ^= 1; }
      }
      flush();
    }
    return;
  }

  // ===== FFT warps: three groups of four; group g takes half tiles g, g + 3, ... of this CTA =====
  const int grp = tid / kGroupH, t = tid % kGroupH;
  const GroupSync gsync{2 + grp};
  double s1 = 0.0, s2 = 0.0;
  const float* zs = p.k3.zscale;
  int b = grp, ph = 0;      // slots grp, grp + 3, grp, ...: the phase flips every second visit
  for (;;) {
    cpx<float>* const B = buf0 + b * kSlotElems;
    mbar_wait(&full[b], ph);
    const int type = desc[4 * b + 0];
    if (type == kTileEnd) break;
    if (type == kTileK2) {
      k2h_tile_body<NY>(B, t, tw2s, gsync);
    } else {
      cpx<float> a1 = {0.f, 0.f}, a2 = {0.f, 0.f};
      k3h_tile_body<NZ>(B, t, twHs, twNs, zs, a1, a2, gsync);
      s1 += (double)a1.x + (double)a1.y;
      s2 += (double)a2.x + (double)a2.y;
    }
    fence_proxy_async();
    mbar_arrive(&done[b]);
    b += kGroupsH;
    if (b >= kSlotsH) { b -= kSlotsH; ph ^= 1; }
  }

  if (p.k3.partials) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    __shared__ double red[2][kFftThreadsH / 32];
    const int wid = tid >> 5, lane = tid & 31;
    if (lane == 0) { red[0][wid] = s1; red[1][wid] = s2; }
    named_sync(1, kFftThreadsH);
    if (tid == 0) {
      double t1 = 0, t2 = 0;
      for (int q = 0; q < kFftThreadsH / 32; ++q) { t1 += red[0][q]; t2 += red[1][q]; }
      p.k3.partials[2 * (size_t)blockIdx.x] = t1;
      p.k3.partials[2 * (size_t)blockIdx.x + 1] = t2;
    }
  }
}

}  // namespace pipe
}  // namespace rf
