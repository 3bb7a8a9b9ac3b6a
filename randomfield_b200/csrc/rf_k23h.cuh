// rf_k23h.cuh -- the fused y + z pass with TWO FFT groups on half tiles ("ping-pong").
//
// k23q_fused_yz (rf_k23q.cuh) moves a 64 KB tile through shared memory six times (TMA in, load A,
// store A', load B, store B', TMA out: 3072 cycles of the 128 B/clk port) and spends about 2200
// FMA-pipe cycles on its butterflies -- and because its eight FFT warps walk through the phases of
// a tile together, the two add up (ncu, round 2: a tile takes the sum, 5300 cycles, with the
// shared-memory port 60 % and the FMA pipe 25 % busy).  Here the FFT warps are two groups of four.
// Every tile is cut in two independent halves, one per group, each in its own half buffer:
//   y-pass tile [NY][TZ columns]  ->  two dense sub-tiles [NY][TZ/2]   (columns are independent)
//   z-pass tile [TL lines][NZ/2]  ->  two sets of TL/2 lines            (lines are independent)
// and the groups hand a token back and forth through named barriers: a group runs butterflies only
// while it holds the token, and does its shared-memory phases without it.  One group's loads and
// stores therefore run under the other group's butterflies.  Same arithmetic on the same values in
// the same order as k2p / k3q / k23q: the field is bit-identical.
//   bar 1       : all 256 FFT threads (final statistics)
//   bar 2 + g   : the 128 threads of group g
//   bar 4 + g   : group g may run butterflies (bar.sync by group g, bar.arrive by the other group)
// Loader / storer warps, tickets, plane counters and the three-deep buffer ring are k23q's.
#pragma once
#include "rf_k23q.cuh"

namespace rf {
namespace pipe {

constexpr int kGroupH = kConsumers / 2;   // 128 threads per FFT group

template <int NY, int NZ>
struct K23H {
  using P2 = SPlan<NY>;
  using P3 = LPlan<NZ>;
  static constexpr int TZh = P2::TZ / 2;                 // columns of a y-pass half tile
  static constexpr int TEh = P3::TE / 2;                 // lines of a z-pass half tile
  static constexpr int kHalf2 = (NY + P2::R1) * TZh * 8;   // padded y-pass half tile
  static constexpr int kHalf3 = TEh * P3::WP * 8;          // padded z-pass half tile
  static constexpr int kHalf = (((kHalf2 > kHalf3 ? kHalf2 : kHalf3) + 127) / 128) * 128;
  static constexpr int kBuf = 2 * kHalf;
  static constexpr int kSmem = 3 * kBuf + 2 * kBarBytes + (NY + NZ) * 8;
  static constexpr bool kOk = K3Fuse<NZ>::value && P2::TZ >= 8 && (P2::M * TZh == kGroupH) &&
                              (P3::M / 2) * TEh == kGroupH && kSmem <= 227 * 1024;
};

struct GroupSync {
  int id;
  __device__ __forceinline__ void operator()() const { named_sync(id, kGroupH); }
};
// the butterfly token: strict alternation between the two groups
struct FmaToken {
  int mine, other;
  __device__ __forceinline__ void acquire() const { named_sync(mine, kConsumers); }
  __device__ __forceinline__ void release() const { named_arrive(other, kConsumers); }
};

// one y-pass half tile, in place in `B` (dense [N][TZh] at the start of a kHalf buffer); t in [0, 128)
template <int N>
__device__ __forceinline__ void k2h_tile_body(cpx<float>* __restrict__ B, int t, const cpx<float>* tws,
                                              const GroupSync gsync, const FmaToken tok) {
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
  cpx<float> twr[R1];
  sfor<0, R1>([&](auto ki) {
    constexpr int k1 = decltype(ki)::value;
    twr[k1] = ld_cpx(tws + k1 * M + j);
  });
  gsync();
  tok.acquire();
  fft_dif<R1, true>(v);
  sfor<1, R1>([&](auto ki) {
    constexpr int k1 = decltype(ki)::value;
    v[brev<R1>(k1)] = cmul(v[brev<R1>(k1)], twr[k1]);
  });
  tok.release();
  {
    cpx<float>* wa = B + L::idx2(0, j, c);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      st_cpx(wa + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), v[brev<R1>(k1)]);
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
  tok.acquire();
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    fft_dif<R2, true>(u[q]);
  });
  tok.release();
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
                                              cpx<float>& a2, const GroupSync gsync, const FmaToken tok) {
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
  cpx<float> twr[R1], twr2[R1];
  sfor<0, R1>([&](auto ki) {
    constexpr int k1 = decltype(ki)::value;
    twr[k1] = ld_cpx(twHt + k1 * M + j0);
    twr2[k1] = ld_cpx(twHt + k1 * M + j1);
  });
  const cpx<float> wj = ld_cpx(twN + (pp == 0 ? PH : pp));
  gsync();
  tok.acquire();
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
  sfor<1, R1>([&](auto ki) {
    constexpr int k1 = decltype(ki)::value;
    v0[brev<R1>(k1)] = cmul(v0[brev<R1>(k1)], twr[k1]);
    v1[brev<R1>(k1)] = cmul(v1[brev<R1>(k1)], twr2[k1]);
  });
  tok.release();
  {
    cpx<float>* wp0 = B + L::idx2(0, j0, l);
    cpx<float>* wp1 = B + L::idx2(0, j1, l);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      st_cpx(wp0 + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), v0[brev<R1>(k1)]);
      st_cpx(wp1 + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), v1[brev<R1>(k1)]);
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
  tok.acquire();
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    fft_dif<R2, true>(u[q]);
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      stats_accum(u[q][brev<R2>(k2)], a1, a2);
    });
  });
  tok.release();
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

// map_in / map_out: the tensor maps of k23q with the box cut to TZ/2 columns
template <int NY, int NZ>
__global__ void __launch_bounds__(kThreadsQ, 1)
k23h_fused_yz(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_out,
              const K23Params p) {
  using P2 = SPlan<NY>;
  using P3 = LPlan<NZ>;
  using G = K23H<NY, NZ>;
  static_assert(G::kOk, "k23h: unsupported (NY, NZ)");
  constexpr int kBufElems = G::kBuf / 8, kHalfElems = G::kHalf / 8;
  constexpr int TZh = G::TZh;
  constexpr uint32_t kHalf3Bytes = (uint32_t)G::TEh * P3::H * 8;   // dense z-pass half tile
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const buf0 = reinterpret_cast<cpx<float>*>(smem_raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + 3 * G::kBuf);
  uint64_t* full = bars;        // [3] tile landed (or end marker published)
  uint64_t* done = bars + 3;    // [3] both FFT groups finished the tile
  uint64_t* empty = bars + 6;   // [3] the tile's store has been read out of the buffer
  volatile int* desc = reinterpret_cast<volatile int*>(bars + 9);   // [3][4]: type, plane, idx
  cpx<float>* const tw2s = reinterpret_cast<cpx<float>*>(smem_raw + 3 * G::kBuf + 2 * kBarBytes);
  cpx<float>* const twHs = tw2s + NY;
  cpx<float>* const twNs = twHs + NZ / 2;
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int b = 0; b < 3; ++b) {
      mbar_init(&full[b], 1);
      mbar_init(&done[b], kConsumers);
      mbar_init(&empty[b], 1);
    }
    fence_barrier_init();
  }
  // step-twiddle tables in [k1][j] order (see k23q)
  for (int q = tid; q < NY; q += kThreadsQ) tw2s[q] = ldg_cpx(p.k2.tw + (q % P2::M) * (q / P2::M));
  for (int q = tid; q < NZ / 2; q += kThreadsQ) {
    twHs[q] = ldg_cpx(p.k3.twH + (q % P3::M) * (q / P3::M));
    twNs[q] = ldg_cpx(p.k3.twN + q);
  }
  __syncthreads();
  char* const gbase = reinterpret_cast<char*>(p.k3.buf);

  if (tid >= kConsumers) {
    if (tid == kConsumers) {
      // ===== loader (as k23q; every tile arrives as two half tiles) =====
      tma_prefetch_desc(&map_in);
      unsigned int ticket = atomicAdd(p.ticket, 1u);
      int b = 0, ph = 0;
      for (int n = 0;; ++n) {
        const unsigned int ticket_next = atomicAdd(p.ticket, 1u);
        if (n >= 3) mbar_wait(&empty[b], ph ^ 1);
        int type, plane, idx;
        k23_decode(p, ticket, type, plane, idx);
        ticket = ticket_next;
        desc[4 * b + 0] = type; desc[4 * b + 1] = plane; desc[4 * b + 2] = idx;
        cpx<float>* dst = buf0 + b * kBufElems;
        if (type == kTileEnd) {
          mbar_arrive(&full[b]);
          break;
        }
        if (type == kTileK2) {
          const int z0 = idx * P2::TZ;
          mbar_expect_tx(&full[b], (uint32_t)P2::kDenseBytes);
          for (int g = 0; g < 2; ++g)
            for (int r0 = 0; r0 < NY; r0 += p.k2.box_rows) {
              const int src = r0 / p.k2.rows_per_src, rl = r0 - src * p.k2.rows_per_src;
              tma_load_4d(dst + g * kHalfElems + r0 * TZh, &map_in, &full[b], 2 * (z0 + g * TZh), rl, plane, src);
            }
        } else {
          for (;;) {
            unsigned int have;
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(have) : "l"(p.plane_done + plane) : "memory");
            if (have >= (unsigned)p.t2) break;
            __nanosleep(200);
          }
          asm volatile("fence.proxy.async;" ::: "memory");
          const size_t off = ((size_t)plane * p.t3 + idx) * (size_t)P3::kDenseBytes;
          mbar_expect_tx(&full[b], (uint32_t)P3::kDenseBytes);
          for (int g = 0; g < 2; ++g)
            bulk_g2s(dst + g * kHalfElems, gbase + off + (size_t)g * kHalf3Bytes, kHalf3Bytes, &full[b]);
        }
        if (++b == 3) { b = 0; ph ^= 1; }
      }
    } else if (tid == kConsumers + 32) {
      // ===== storer (as k23q) =====
      tma_prefetch_desc(&map_out);
      constexpr int kSigLag = RF_K23Q_SIGLAG;
      int pend[kSigLag];
#pragma unroll
      for (int q = 0; q < kSigLag; ++q) pend[q] = -1;
      auto signal = [&](int plane) {
        if (plane >= 0) {
          asm volatile("fence.proxy.async;" ::: "memory");
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.plane_done + plane) : "memory");
        }
      };
      auto flush = [&]() {
        bulk_wait0();
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
        const cpx<float>* src = buf0 + b * kBufElems;
        if (type == kTileK2) {
          const int z0 = idx * P2::TZ;
          constexpr int kBox = P2::kBoxRows;
          for (int g = 0; g < 2; ++g)
            for (int r0 = 0; r0 < NY; r0 += kBox)
              tma_store_3d(&map_out, src + g * kHalfElems + r0 * TZh, 2 * (z0 + g * TZh), r0, plane);
        } else {
          const size_t off = ((size_t)plane * p.t3 + idx) * (size_t)P3::kDenseBytes;
          for (int g = 0; g < 2; ++g)
            bulk_s2g(gbase + off + (size_t)g * kHalf3Bytes, src + g * kHalfElems, kHalf3Bytes);
        }
        bulk_commit();
        bulk_wait_read0();
        mbar_arrive(&empty[b]);
        asm volatile("cp.async.bulk.wait_group %0;" ::"n"(kSigLag) : "memory");
        signal(pend[0]);
#pragma unroll
        for (int q = 0; q + 1 < kSigLag; ++q) pend[q] = pend[q + 1];
        pend[kSigLag - 1] = type == kTileK2 ? plane : -1;
        if (++b == 3) { b = 0; ph ^= 1; }
      }
      flush();
    }
    return;
  }

  // ===== FFT warps: two groups of four =====
  const int grp = tid / kGroupH, t = tid % kGroupH;
  const GroupSync gsync{2 + grp};
  const FmaToken tok{4 + grp, 4 + (1 - grp)};
  if (grp == 1) named_arrive(4, kConsumers);      // group 0 holds the token first
  double s1 = 0.0, s2 = 0.0;
  const float* zs = p.k3.zscale;
  int b = 0, ph = 0;
  for (;;) {
    cpx<float>* const B = buf0 + b * kBufElems + grp * kHalfElems;
    mbar_wait(&full[b], ph);
    const int type = desc[4 * b + 0];
    if (type == kTileEnd) break;
    if (type == kTileK2) {
      k2h_tile_body<NY>(B, t, tw2s, gsync, tok);
    } else {
      cpx<float> a1 = {0.f, 0.f}, a2 = {0.f, 0.f};
      k3h_tile_body<NZ>(B, t, twHs, twNs, zs, a1, a2, gsync, tok);
      s1 += (double)a1.x + (double)a1.y;
      s2 += (double)a2.x + (double)a2.y;
    }
    fence_proxy_async();
    mbar_arrive(&done[b]);
    if (++b == 3) { b = 0; ph ^= 1; }
  }
  // the last release of group 1 has no taker: consume it so that no barrier is left half-arrived
  if (grp == 0) named_sync(4, kConsumers);

  if (p.k3.partials) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    __shared__ double red[2][kConsumers / 32];
    const int wid = tid >> 5, lane = tid & 31;
    if (lane == 0) { red[0][wid] = s1; red[1][wid] = s2; }
    ConsumerSync()();
    if (tid == 0) {
      double t1 = 0, t2 = 0;
      for (int q = 0; q < kConsumers / 32; ++q) { t1 += red[0][q]; t2 += red[1][q]; }
      p.k3.partials[2 * (size_t)blockIdx.x] = t1;
      p.k3.partials[2 * (size_t)blockIdx.x + 1] = t2;
    }
  }
}

}  // namespace pipe
}  // namespace rf
