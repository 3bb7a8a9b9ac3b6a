// rf_k3f.cuh -- r2c along the contiguous axis: the first pass of the FORWARD transform the P(k)
// estimator needs (rf_r2c_power; the reference plans it as plan_r2c, generate.py:79-80,
// transform.py:278-301, and never executes it).
//
// The estimator only needs |F(k)|^2, and F = conj(G) with G the transform of the real field with
// the INVERSE sign, so the whole forward transform runs on the inverse-sign engine of the hot
// path: this kernel (z), then k2p_strided_fft along y and along x.  A real line f[0..NZ) is read
// as H = NZ/2 complex values z[n] = f[2n] + i f[2n+1]; Z = the H-point transform of z (steps A and
// B of k3q, same radix plan); then the pairs (k, H-k) are combined:
//     E = (Z[k] + conj Z[H-k]) / 2,  O = (Z[k] - conj Z[H-k]) / 2i,  w = exp(+2 pi i k / NZ)
//     G[k] = E + w O,   G[H-k] = conj(E - w O);   G[0] = Re Z[0] + Im Z[0],  G[H] = Re Z[0] - Im Z[0]
// G[0] and G[H] are real and share column 0 (G[0] + i G[H]) -- the packed-Nyquist layout of the
// hot path, read in the other direction.  In place on [nlines][NZ/2] complex, three-deep tile ring
// (the skeleton of k3q_c2r_lines).
#pragma once
#include "rf_k3q.cuh"

namespace rf {
namespace pipe {

template <int NZ>
__device__ __forceinline__ void k3f_tile_body(cpx<float>* __restrict__ B, int tid,
                                              const cpx<float> (&twr)[LPlan<NZ>::R1], const cpx<float>* twH,
                                              const cpx<float>* twN) {
  using P = LPlan<NZ>;
  constexpr int H = P::H, HE = P::HE, R1 = P::R1, R2 = HE / R1, M = P::M;
  using L = LayoutLinesPad<P::WP, P::M>;
  constexpr int IA = M * P::TE / kConsumers, IB = R1 * P::TE / kConsumers;
  static_assert(IA * kConsumers == M * P::TE && IB * kConsumers == R1 * P::TE, "items per thread");
  if constexpr (P::SPLIT == 2) {
    // radix-2 DIF across the line halves: the engine then runs on 2 TL half lines
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
  // ---- step B: padded -> registers | barrier | registers -> dense ----
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
      const int k = k1 + R1 * k2;
      const int pos = P::SPLIT == 1 ? k : P::SPLIT * k + (l % P::SPLIT);
      st_cpx(B + (l / P::SPLIT) * H + pos, u[q][brev<R2>(k2)]);
    });
  });
  ConsumerSync()();
  // ---- combine the pairs (k, H-k): Z -> G ----
  for (int w = tid; w < P::TL * (H / 2); w += kConsumers) {
    const int l = w / (H / 2), k = w % (H / 2);
    cpx<float>* Lp = B + l * H;
    if (k == 0) {
      const cpx<float> a = Lp[0];
      Lp[0] = {a.x + a.y, a.x - a.y};        // G[0] + i G[H]; Z[H/2] is G[H/2] already
    } else {
      const cpx<float> zk = Lp[k], zh = Lp[H - k];
      const cpx<float> s = pk_add(zk.x, zk.y, zh.x, -zh.y);    // Z[k] + conj Z[H-k]
      const cpx<float> d = pk_add(zk.x, zk.y, -zh.x, zh.y);    // Z[k] - conj Z[H-k]
      const cpx<float> e = pk_mul(s.x, s.y, 0.5f, 0.5f);
      const cpx<float> o = pk_mul(d.y, d.x, 0.5f, -0.5f);      // d / 2i
      const cpx<float> t = cmul(o, ldg_cpx(twN + k));
      Lp[k] = e + t;
      const cpx<float> g = e - t;
      Lp[H - k] = {g.x, -g.y};
    }
  }
}

template <int NZ>
__global__ void __launch_bounds__(kThreads, 1) k3f_r2c_lines(const K3PParams p) {
  using P = LPlan<NZ>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int kBufElems = P::kWorkBytes / 8;
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

  cpx<float> twr[P::R1];
  load_step_twiddles<P::R1, P::SPLIT>(twr, p.twH, tid % P::M);
  int b = 0, ph = 0;
  for (int t = first; t < p.ntiles; t += step) {
    cpx<float>* const B = buf0 + b * kBufElems;
    mbar_wait(&full[b], ph);
    k3f_tile_body<NZ>(B, tid, twr, p.twH, p.twN);
    fence_proxy_async();
    mbar_arrive(&done[b]);
    if (++b == 3) { b = 0; ph ^= 1; }
  }
}

}  // namespace pipe
}  // namespace rf
