// rf_k3q.cuh -- pass 3 (c2r along the contiguous axis) with a THREE-deep tile ring.
//
// k3p keeps two dense 64 KB buffers plus a 66 KB padded work buffer: while one tile is
// transformed the other buffer has to drain its store AND receive the next load, so the FFT
// warps wait for HBM latency about 40 % of the time (5.5 TB/s).  Here the transform of a tile
// is in place in its own buffer -- step A: every thread reads its radix-R1 items, barrier,
// writes the padded layout; step B: reads its radix-R2 items, barrier, writes the dense
// layout -- so the work buffer goes away and three 66 KB buffers fit: one loading, one being
// transformed, one draining, at all times.  Same arithmetic, in the same order, as k3p.
#pragma once
#include "rf_pipe.cuh"

namespace rf {
namespace pipe {

template <int NZ>
constexpr int k3q_smem_bytes() {
  return 3 * LPlan<NZ>::kWorkBytes + kBarBytes;
}

// Fused Hermitian unpack (NZ = 256, 512, 1024): the pair (k, H-k) of the unpack lives in the step-A
// items j and M - j of the same line (k = j + M m  <->  H - k = (M - j) + M (R1 - 1 - m)), so a thread
// that owns BOTH items unpacks in registers while it loads step A: the separate unpack sweep (one
// more read and one more write of the tile through shared memory, and its barrier) goes away.
// Thread -> (p, line), p = tid % (M/2): items (p, M - p) for p >= 1 and the two self-paired items
// (0, M/2) for p = 0.  The unpack twiddle exp(2 pi i k / NZ) = W_NZ^j * W_(2 R1)^m is one per-thread
// value times a compile-time 32nd root of unity.
template <int NZ>
struct K3Fuse {
  using P = LPlan<NZ>;
  static constexpr bool value = P::SPLIT == 1 && (P::R1 == 16 || P::R1 == 8) && P::M >= 4 &&
                                (P::M * P::TE == 2 * kConsumers) && NZ >= 256;
  static constexpr int kStep = 32 / (2 * P::R1);   // W_(2 R1)^m = W_32^(kStep m)
};

// twiddles of this thread for k3q_tile_body: step-A factors of its item(s) and the unpack factor(s).
// tabH holds exp(2 pi i k / H) -- or, TRANSPOSED, the step table W_H^(j k1) at [k1 * M + j] (SPLIT == 1) --
// and tabN exp(2 pi i k / NZ) (global memory or shared-memory copies).
template <int NZ, bool TRANSPOSED = false, class LD>
__device__ __forceinline__ void k3q_load_twiddles(cpx<float> (&twr)[LPlan<NZ>::R1], cpx<float> (&twr2)[LPlan<NZ>::R1],
                                                  cpx<float> (&twu)[K3Consts<NZ>::kPairsPerThread],
                                                  const cpx<float>* tabH, const cpx<float>* tabN, int tid, LD&& ld) {
  using P = LPlan<NZ>;
  constexpr int H = P::H, M = P::M, R1 = P::R1;
  if constexpr (K3Fuse<NZ>::value) {
    const int pp = tid % (M / 2);
    const int j0 = pp, j1 = pp == 0 ? M / 2 : M - pp;
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      twr[k1] = ld(tabH + (TRANSPOSED ? k1 * M + j0 : P::SPLIT * j0 * k1));
      twr2[k1] = ld(tabH + (TRANSPOSED ? k1 * M + j1 : P::SPLIT * j1 * k1));
    });
    twu[0] = ld(tabN + (pp == 0 ? M / 2 : pp));
  } else if constexpr (P::SPLIT == 2 && (H / 4) % kConsumers == 0) {
    // quad sweep: twu[2q] = W_NZ^k, twu[2q+1] = W_NZ^(H/2-k), k = tid + 256 q < H/4
    sfor<0, K3Consts<NZ>::kPairsPerThread / 2>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int k = tid + q * kConsumers;
      twu[2 * q] = ld(tabN + k);
      twu[2 * q + 1] = ld(tabN + (H / 2 - k) % (H / 2) + (k == 0 ? H / 4 : 0));   // k = 0: the (H/4, 3H/4) pair
    });
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      twr[k1] = ld(tabH + (TRANSPOSED ? k1 * M + tid % M : P::SPLIT * (tid % M) * k1));
      twr2[k1] = twr[k1];
    });
  } else {
    sfor<0, K3Consts<NZ>::kPairsPerThread>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      twu[q] = ld(tabN + (tid + q * kConsumers) % (H / 2));
    });
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      twr[k1] = ld(tabH + (TRANSPOSED ? k1 * M + tid % M : P::SPLIT * (tid % M) * k1));
      twr2[k1] = twr[k1];
    });
  }
}

// one tile, in place in `B` (dense [TL][H] at the start of a kWorkBytes buffer)
template <int NZ>
__device__ __forceinline__ void k3q_tile_body(cpx<float>* __restrict__ B, int tid,
                                              const cpx<float> (&twr)[LPlan<NZ>::R1],
                                              const cpx<float> (&twr2)[LPlan<NZ>::R1],
                                              const cpx<float> (&twu)[K3Consts<NZ>::kPairsPerThread],
                                              const float* zs, cpx<float>& a1, cpx<float>& a2, const cpx<float>* twH) {
  using P = LPlan<NZ>;
  constexpr int H = P::H, HE = P::HE, R1 = P::R1, R2 = HE / R1, M = P::M;
  constexpr int kPairsPerThread = K3Consts<NZ>::kPairsPerThread;
  using L = LayoutLinesPad<P::WP, P::M>;
  constexpr int IA = M * P::TE / kConsumers, IB = R1 * P::TE / kConsumers;
  static_assert(IA * kConsumers == M * P::TE && IB * kConsumers == R1 * P::TE, "items per thread");
  static_assert(IA * R1 == 32 && IB * R2 == 32, "32 values per thread and step");
  if constexpr (K3Fuse<NZ>::value) {
    // ---- fused: dense -> registers (+ Hermitian unpack) | barrier | registers -> padded ----
    static_assert(IA == 2, "two items per thread");
    constexpr int PH = M / 2, KS = K3Fuse<NZ>::kStep;
    const int pp = tid % PH, l = tid / PH;
    const int j0 = pp, j1 = pp == 0 ? PH : M - pp;
    const cpx<float>* Lp = B + l * HE;
    cpx<float> v0[R1], v1[R1];
    sfor<0, R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      v0[m] = ld_cpx(Lp + j0 + M * m);
      v1[m] = ld_cpx(Lp + j1 + M * m);
    });
    ConsumerSync()();
    // twu[0] = W_NZ^j0 (p >= 1) / W_NZ^(M/2) (p = 0): the caller loads it for this mapping
    const cpx<float> wj = twu[0];
    if (pp != 0) {
      // k = j0 + M m (item 0)  <->  H - k = j1 + M (R1 - 1 - m) (item 1)
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
      // item 0 (j = 0): k = M m <-> M (R1 - m); k = 0 and k = H/2 (m = R1/2) are the real ends
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
      // item 1 (j = M/2): k = M/2 + M m <-> M/2 + M (R1 - 1 - m)
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
    cpx<float>* wp0 = B + L::idx2(0, j0, l);
    cpx<float>* wp1 = B + L::idx2(0, j1, l);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x0 = v0[brev<R1>(k1)], x1 = v1[brev<R1>(k1)];
      if constexpr (k1 > 0) {
        x0 = cmul(x0, twr[k1]);
        x1 = cmul(x1, twr2[k1]);
      }
      st_cpx(wp0 + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x0);
      st_cpx(wp1 + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x1);
    });
    ConsumerSync()();
  } else {
  if constexpr (P::SPLIT == 2 && (H / 4) % kConsumers == 0) {
    // ---- Hermitian unpack AND the radix-2 split across the line halves in ONE sweep ----
    // The unpack pairs (k, H-k), (H/2-k, H/2+k) and the split butterflies (k, k+H/2), (H/2-k, H-k) close
    // over the four elements {k, H/2-k, H/2+k, H-k}: one read and one write of the tile instead of two
    // of each, one barrier instead of two.  Same operations on the same values as the two sweeps.
    constexpr int QPT = (H / 4) / kConsumers;     // quads per thread and line
    static_assert(2 * QPT == kPairsPerThread, "twiddle registers");
#pragma unroll 1
    for (int l = 0; l < P::TL; ++l) {
      cpx<float>* Lp = B + l * H;
      sfor<0, QPT>([&](auto qi) {
        constexpr int q = decltype(qi)::value;
        const int k = tid + q * kConsumers;       // 0 <= k < H/4
        if (q == 0 && k == 0) {
          // k = 0: the real ends (0, H/2) and their butterfly (W^0 = 1); and the self-paired (H/4, 3H/4)
          cpx<float> z0, zm;
          c2r_unpack_ends(Lp[0], Lp[H / 2], z0, zm);
          Lp[0] = z0 + zm;
          Lp[H / 2] = cmul(z0 - zm, ldg_cpx(twH));
          cpx<float> za, zb;
          c2r_unpack_pair(Lp[H / 4], Lp[3 * H / 4], twu[2 * q + 1], za, zb);   // twN[H/4 - 0 ... ] see loader
          Lp[H / 4] = za + zb;
          Lp[3 * H / 4] = cmul(za - zb, ldg_cpx(twH + H / 4));
        } else {
          cpx<float> z1, z4, z2, z3;
          c2r_unpack_pair(Lp[k], Lp[H - k], twu[2 * q], z1, z4);                       // Z[k], Z[H-k]
          c2r_unpack_pair(Lp[H / 2 - k], Lp[H / 2 + k], twu[2 * q + 1], z2, z3);       // Z[H/2-k], Z[H/2+k]
          Lp[k] = z1 + z3;
          Lp[H / 2 + k] = cmul(z1 - z3, ldg_cpx(twH + k));
          Lp[H / 2 - k] = z2 + z4;
          Lp[H - k] = cmul(z2 - z4, ldg_cpx(twH + H / 2 - k));
        }
      });
    }
    ConsumerSync()();
  } else {
  // ---- Hermitian unpack, in place (as k3_tile_body) ----
  if constexpr (H / 2 >= kConsumers) {
#pragma unroll 2
    for (int l = 0; l < P::TL; ++l) {
      cpx<float>* Lp = B + l * H;
      sfor<0, kPairsPerThread>([&](auto qi) {
        constexpr int q = decltype(qi)::value;
        const int k = tid + q * kConsumers;
        if (q == 0 && k == 0) {
          cpx<float> a = Lp[0], m = Lp[H / 2];
          c2r_unpack_ends(a, m, Lp[0], Lp[H / 2]);
        } else {
          cpx<float> a = Lp[k], c = Lp[H - k];
          c2r_unpack_pair(a, c, twu[q], Lp[k], Lp[H - k]);
        }
      });
    }
  } else {
    constexpr int kLinesPerSweep = kConsumers / (H / 2);
    const int k = tid % (H / 2);
    for (int l = tid / (H / 2); l < P::TL; l += kLinesPerSweep) {
      cpx<float>* Lp = B + l * H;
      if (k == 0) {
        cpx<float> a = Lp[0], m = Lp[H / 2];
        c2r_unpack_ends(a, m, Lp[0], Lp[H / 2]);
      } else {
        cpx<float> a = Lp[k], c = Lp[H - k];
        c2r_unpack_pair(a, c, twu[0], Lp[k], Lp[H - k]);
      }
    }
  }
  ConsumerSync()();
  if constexpr (P::SPLIT == 2) {
    for (int w = tid; w < P::TL * (H / 2); w += kConsumers) {
      const int l = w / (H / 2), k = w % (H / 2);
      cpx<float>* Lp = B + l * H + k;
      cpx<float> a = Lp[0], c = Lp[H / 2];
      cpx<float> tw = ldg_cpx(twH + k);
      Lp[0] = a + c;
      Lp[H / 2] = cmul(a - c, tw);
    }
    ConsumerSync()();
  }
  }   // (separate unpack / split sweeps)
  // ---- step A: dense -> registers | barrier | registers -> padded ----
  cpx<float> v[IA][R1];
  sfor<0, IA>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kConsumers;
    int j, l;
    L::split(w, M, j, l);
    const cpx<float>* Lp = B + l * HE + j;
    sfor<0, R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      v[q][m] = ld_cpx(Lp + M * m);
    });
  });
  ConsumerSync()();
  sfor<0, IA>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kConsumers;
    int j, l;
    L::split(w, M, j, l);
    fft_dif<R1, true>(v[q]);
    cpx<float>* wp = B + L::idx2(0, j, l);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x = v[q][brev<R1>(k1)];
      if constexpr (k1 > 0) x = cmul(x, twr[k1]);
      st_cpx(wp + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x);
    });
  });
  ConsumerSync()();
  }   // (!K3Fuse)
  // ---- step B: padded -> registers | barrier | registers -> dense (+ statistics, z scale) ----
  cpx<float> u[IB][R2];
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kConsumers;
    int k1, l;
    L::split(w, R1, k1, l);
    const cpx<float>* wp = B + L::idx2(k1, 0, l);
    sfor<0, R2>([&](auto ji) {
      constexpr int j = decltype(ji)::value;
      u[q][j] = ld_cpx(wp + (L::idx2(0, j, 0) - L::idx2(0, 0, 0)));
    });
  });
  ConsumerSync()();
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kConsumers;
    int k1, l;
    L::split(w, R1, k1, l);
    fft_dif<R2, true>(u[q]);
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      cpx<float> x = u[q][brev<R2>(k2)];
      const int k = k1 + R1 * k2;
      stats_accum(x, a1, a2);
      const int pos = P::SPLIT == 1 ? k : P::SPLIT * k + (l % P::SPLIT);
      const int line = l / P::SPLIT;
      if (zs) {
        float2 z = __ldg(reinterpret_cast<const float2*>(zs) + pos);
        x = pk_mul(x.x, x.y, z.x, z.y);
      }
      st_cpx(B + line * H + pos, x);
    });
  });
}

template <int NZ>
__global__ void __launch_bounds__(kThreads, 1) k3q_c2r_lines(const K3PParams p) {
  using P = LPlan<NZ>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int kBufElems = P::kWorkBytes / 8;
  static_assert(P::kWorkBytes % 128 == 0, "bulk copies need 16-byte aligned buffers");
  cpx<float>* const buf0 = reinterpret_cast<cpx<float>*>(smem_raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + 3 * P::kWorkBytes);
  uint64_t* full = bars;       // [3] tile landed
  uint64_t* done = bars + 3;   // [3] FFT warps finished the tile
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int b = 0; b < 3; ++b) {
      mbar_init(&full[b], 1);
      mbar_init(&done[b], kConsumers);
    }
    fence_barrier_init();
  }
  __syncthreads();
  const int first = blockIdx.x, step = gridDim.x;
  char* gbase = reinterpret_cast<char*>(p.buf) + (size_t)p.tile0 * P::kDenseBytes;
  constexpr size_t kTile = P::kDenseBytes;

  if (tid >= kConsumers) {
    if (tid == kConsumers) {
      int n = 0;
      for (int t = first; t < p.ntiles && n < 3; t += step, ++n) {
        mbar_expect_tx(&full[n], (uint32_t)kTile);
        bulk_g2s(buf0 + n * kBufElems, gbase + (size_t)t * kTile, (uint32_t)kTile, &full[n]);
      }
      int i = 0, b = 0, ph = 0;
      for (int t = first; t < p.ntiles; t += step, ++i) {
        // while the FFT warps work on tile i: wait for the store of tile i-1 to be read out
        // of its buffer and refill that buffer with tile i+2 (two tile times ahead of its use)
        if (i >= 1) {
          const int pb = b == 0 ? 2 : b - 1;
          const int t3 = t + 2 * step;
          if (t3 < p.ntiles) {
            bulk_wait_read0();
            mbar_expect_tx(&full[pb], (uint32_t)kTile);
            bulk_g2s(buf0 + pb * kBufElems, gbase + (size_t)t3 * kTile, (uint32_t)kTile, &full[pb]);
          }
        }
        mbar_wait(&done[b], ph);
        bulk_s2g(gbase + (size_t)t * kTile, buf0 + b * kBufElems, (uint32_t)kTile);
        bulk_commit();
        if (++b == 3) { b = 0; ph ^= 1; }
      }
      bulk_wait0();
    }
    return;
  }

  double s1 = 0.0, s2 = 0.0;
  const float* zs = p.zscale;
  cpx<float> twr[P::R1], twr2[P::R1];
  cpx<float> twu[K3Consts<NZ>::kPairsPerThread];
  k3q_load_twiddles<NZ, false>(twr, twr2, twu, p.twH, p.twN, tid, [](const cpx<float>* q) { return ldg_cpx(q); });
  int b = 0, ph = 0;
  for (int t = first; t < p.ntiles; t += step) {
    cpx<float>* const B = buf0 + b * kBufElems;
    mbar_wait(&full[b], ph);
    cpx<float> a1 = {0.f, 0.f}, a2 = {0.f, 0.f};
    k3q_tile_body<NZ>(B, tid, twr, twr2, twu, zs, a1, a2, p.twH);
    s1 += (double)a1.x + (double)a1.y;
    s2 += (double)a2.x + (double)a2.y;
    fence_proxy_async();
    mbar_arrive(&done[b]);
    if (++b == 3) { b = 0; ph ^= 1; }
  }
  if (p.partials) {
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
      p.partials[2 * (size_t)blockIdx.x] = t1;
      p.partials[2 * (size_t)blockIdx.x + 1] = t2;
    }
  }
}

}  // namespace pipe
}  // namespace rf
