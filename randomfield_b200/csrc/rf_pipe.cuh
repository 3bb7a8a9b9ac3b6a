// rf_pipe.cuh -- persistent, TMA-pipelined versions of the three hot-path kernels (sm_100a).
//
// One CTA per SM walks a strided list of tiles.  Warp 8 (one elected lane) is the copy
// engine driver: it streams tiles between HBM and a two-deep ring of dense shared-memory
// buffers with bulk-async copies (cp.async.bulk / cp.async.bulk.tensor) and mbarriers; warps
// 0-7 do the FFT of the tile that has landed while the next one is in flight and the
// previous one drains.  The FFT itself is the two-step in-register radix engine of
// rf_fast.cuh, run out of place between the dense buffer and a padded (bank-conflict-free)
// work buffer, so the dense buffer can be handed to the copy engine as is.
//
//   k1p_generate_xfft : modes generated in registers -> x FFT -> TMA store      (4 B/cell)
//   k2p_strided_fft   : TMA load -> y FFT -> TMA store                          (8 B/cell)
//   k3p_c2r_lines     : bulk load -> Hermitian unpack + z FFT -> bulk store     (8 B/cell)
//
// float32 only; sizes outside the plan table run on the rf_fast.cuh kernels.
#pragma once
#include <cuda.h>

#include "rf_fast.cuh"

namespace rf {
namespace pipe {

constexpr int kConsumers = 256;              // 8 FFT warps
constexpr int kThreads = kConsumers + 32;    // + 1 copy warp
constexpr int kTileBytes = 64 * 1024;        // dense tile

// ----------------------------------------------------------------------------------------
// PTX wrappers
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "RF_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra RF_DONE_%=;\n"
      "bra RF_WAIT_%=;\n"
      "RF_DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (bulk copies)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// contiguous global -> shared, completion counted on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// contiguous shared -> global, tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
               "r"(smem_u32(src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
      "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, const void* src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(m),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
// barrier over the consumer warps only (named barrier 1)
struct ConsumerSync {
  __device__ __forceinline__ void operator()() const {
    asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
  }
};

// ----------------------------------------------------------------------------------------
// plans and padded work-buffer layouts
// ----------------------------------------------------------------------------------------
// strided axis (k1p, k2p): dense tile [N rows][TZ cols]; work tile has one spare row after
// every M = N/R1 rows so that the R1 sub-transforms of step B start in different banks
template <int N>
struct SPlan;
#define RF_SPLAN(NN, A, B, TZV)                              \
  template <>                                                \
  struct SPlan<NN> {                                         \
    static constexpr int R1 = A, R2 = B, TZ = TZV, M = NN / A; \
    static constexpr int kDenseBytes = NN * TZV * 8;         \
    static constexpr int kWorkBytes = (NN + A) * TZV * 8;    \
    static constexpr int kBoxRows = NN < 256 ? NN : 256;     \
  };
// (M * TZ == kConsumers: every FFT thread owns one (j, column) of step A for the whole run,
// so its R1-1 step-A twiddles live in registers)
RF_SPLAN(64, 8, 8, 32)
RF_SPLAN(128, 16, 8, 32)
RF_SPLAN(256, 16, 16, 16)
RF_SPLAN(512, 32, 16, 16)
RF_SPLAN(1024, 32, 32, 8)
#undef RF_SPLAN

template <int TZ, int M>
struct LayoutStridedPad {
  // element (k1*M + j) of column c
  __device__ static __forceinline__ int idx2(int k1, int j, int c) { return (k1 * (M + 1) + j) * TZ + c; }
  __device__ static __forceinline__ void split(int w, int count, int& idx, int& c) {
    c = w % TZ;
    idx = w / TZ;
  }
};

// contiguous axis (k3p): dense tile [TL lines][H]; work line has one spare element after
// every M = H/R1 elements
template <int NZ>
struct LPlan;
// H = complex line length (NZ/2).  The FFT engine runs on HE = H / SPLIT points: for NZ = 4096
// a radix-2 DIF stage in shared memory splits each line into two independent half-lines
// (even / odd outputs), so the two-step radix-32 engine still applies.
#define RF_LPLAN2(NN, A, B)                                  \
  template <>                                                \
  struct LPlan<NN> {                                         \
    static constexpr int H = NN / 2, R1 = A, R2 = B;         \
    static constexpr int HE = A * B, SPLIT = H / HE, M = HE / A; \
    static constexpr int TL = 8192 / H;                      \
    static constexpr int TE = TL * SPLIT;                    \
    static constexpr int WP = HE + A;                        \
    static constexpr int kDenseBytes = TL * H * 8;           \
    static constexpr int kWorkBytes = TE * WP * 8;           \
  };
RF_LPLAN2(64, 8, 4)
RF_LPLAN2(128, 8, 8)
RF_LPLAN2(256, 16, 8)
RF_LPLAN2(512, 16, 16)
RF_LPLAN2(1024, 16, 32)
RF_LPLAN2(2048, 32, 32)
RF_LPLAN2(4096, 32, 32)
#undef RF_LPLAN2

template <int WP, int M>
struct LayoutLinesPad {
  // element (k1*M + j) of line l
  __device__ static __forceinline__ int idx2(int k1, int j, int l) { return l * WP + k1 * (M + 1) + j; }
  __device__ static __forceinline__ void split(int w, int count, int& idx, int& l) {
    idx = w % count;
    l = w / count;
  }
};

// step A: R1-point DFTs over the stride-M samples supplied by loadA, times the step twiddles
// twr[k1] = W_N^(j*k1) (j is fixed per thread), into the padded work buffer
template <int N, int R1, class L, class LoadA>
__device__ __forceinline__ void step_a(cpx<float>* __restrict__ work, int tid, int items_c,
                                       const cpx<float> (&twr)[R1], LoadA&& loadA) {
  constexpr int M = N / R1;
  for (int w = tid; w < M * items_c; w += kConsumers) {
    int j, c;
    L::split(w, M, j, c);
    cpx<float> v[R1];
    loadA(v, j, c);
    fft_dif<R1, true>(v);
    cpx<float>* wp = work + L::idx2(0, j, c);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x = v[brev<R1>(k1)];
      if constexpr (k1 > 0) x = cmul(x, twr[k1]);
      st_cpx(wp + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x);
    });
  }
}
// step B: R2-point DFTs (R2 == M) out of the work buffer; output element k1 + R1*k2
template <int N, int R1, class L, class StoreF>
__device__ __forceinline__ void step_b(const cpx<float>* __restrict__ work, int tid, int items_c,
                                       StoreF&& storeF) {
  constexpr int R2 = N / R1;
  for (int w = tid; w < R1 * items_c; w += kConsumers) {
    int k1, c;
    L::split(w, R1, k1, c);
    cpx<float> v[R2];
    const cpx<float>* wp = work + L::idx2(k1, 0, c);
    sfor<0, R2>([&](auto ji) {
      constexpr int j = decltype(ji)::value;
      v[j] = ld_cpx(wp + (L::idx2(0, j, 0) - L::idx2(0, 0, 0)));
    });
    fft_dif<R2, true>(v);
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      storeF(k1 + R1 * k2, c, v[brev<R2>(k2)]);
    });
  }
}
// the step-A twiddles of this thread: W_N^(j*k1), k1 < R1
template <int R1, int STRIDE = 1>
__device__ __forceinline__ void load_step_twiddles(cpx<float> (&twr)[R1], const cpx<float>* tw, int j) {
  sfor<0, R1>([&](auto ki) {
    constexpr int k1 = decltype(ki)::value;
    twr[k1] = ldg_cpx(tw + j * k1 * STRIDE);
  });
}

// same from a shared-memory copy of the table
template <int R1>
__device__ __forceinline__ void load_step_twiddles_smem(cpx<float> (&twr)[R1], const cpx<float>* tws, int j) {
  sfor<0, R1>([&](auto ki) {
    constexpr int k1 = decltype(ki)::value;
    twr[k1] = ld_cpx(tws + j * k1);
  });
}

constexpr int kBarBytes = 64;   // full[2], done[2] mbarriers

// FFT-warp work on one k3 tile: Hermitian unpack, step A, step B (+ statistics, z scale).
// D: dense tile [TL][H]; twr: step-A twiddles of this thread; twu: its unpack twiddle(s).
template <int NZ>
struct K3Consts {
  using P = LPlan<NZ>;
  static constexpr int H = P::H;
  static constexpr int kPairsPerThread = (H / 2 + kConsumers - 1) / kConsumers;   // per line
};
template <int NZ>
__device__ __forceinline__ void k3_load_twiddles(cpx<float> (&twr)[LPlan<NZ>::R1],
                                                 cpx<float> (&twu)[K3Consts<NZ>::kPairsPerThread],
                                                 const cpx<float>* twH, const cpx<float>* twN, int tid) {
  constexpr int H = LPlan<NZ>::H;
  sfor<0, K3Consts<NZ>::kPairsPerThread>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    twu[q] = ldg_cpx(twN + (tid + q * kConsumers) % (H / 2));
  });
  // table twH holds W_H^k; the engine needs W_HE^(j*k1) = W_H^(SPLIT*j*k1)
  load_step_twiddles<LPlan<NZ>::R1, LPlan<NZ>::SPLIT>(twr, twH, tid % LPlan<NZ>::M);
}
template <int NZ>
__device__ __forceinline__ void k3_tile_body(cpx<float>* __restrict__ D, cpx<float>* __restrict__ work, int tid,
                                             const cpx<float> (&twr)[LPlan<NZ>::R1],
                                             const cpx<float> (&twu)[K3Consts<NZ>::kPairsPerThread],
                                             const float* zs, cpx<float>& a1, cpx<float>& a2,
                                             const cpx<float>* twH = nullptr) {
  using P = LPlan<NZ>;
  constexpr int H = P::H;
  constexpr int HE = P::HE;
  constexpr int kPairsPerThread = K3Consts<NZ>::kPairsPerThread;
  using L = LayoutLinesPad<P::WP, P::M>;
  static_assert(kConsumers % (H / 2) == 0 || (H / 2) % kConsumers == 0, "unpack mapping");
    // Hermitian unpack in place: Z[k] = E[k] + i*O[k] for the pair (k, H-k)
    if constexpr (H / 2 >= kConsumers) {
      // thread owns pairs k = tid + q*kConsumers of every line
#pragma unroll 2
      for (int l = 0; l < P::TL; ++l) {
        cpx<float>* Lp = D + l * H;
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
      // several lines per sweep: thread owns pair k = tid % (H/2) of lines tid/(H/2) + ...
      constexpr int kLinesPerSweep = kConsumers / (H / 2);
      const int k = tid % (H / 2);
      for (int l = tid / (H / 2); l < P::TL; l += kLinesPerSweep) {
        cpx<float>* Lp = D + l * H;
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
      // radix-2 DIF across the two halves of every line, in place:
      //   first half <- Z[k] + Z[k+H/2] (even outputs), second half <- (Z[k] - Z[k+H/2]) W_H^k (odd)
      for (int w = tid; w < P::TL * (H / 2); w += kConsumers) {
        const int l = w / (H / 2), k = w % (H / 2);
        cpx<float>* Lp = D + l * H + k;
        cpx<float> a = Lp[0], c = Lp[H / 2];
        cpx<float> tw = ldg_cpx(twH + k);
        Lp[0] = a + c;
        Lp[H / 2] = cmul(a - c, tw);
      }
      ConsumerSync()();
    }
    // the tile is now TE = TL*SPLIT independent lines of HE points, contiguous in D
    step_a<HE, P::R1, L>(work, tid, P::TE, twr, [&](cpx<float>(&v)[P::R1], int j, int l) {
      const cpx<float>* Lp = D + l * HE + j;
      sfor<0, P::R1>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[m] = ld_cpx(Lp + P::M * m);
      });
    });
    ConsumerSync()();
    step_b<HE, P::R1, L>(work, tid, P::TE, [&](int k, int l, cpx<float> x) {
      stats_accum(x, a1, a2);
      // engine line l = SPLIT*line + parity holds outputs parity, parity + SPLIT, ...
      const int pos = P::SPLIT == 1 ? k : P::SPLIT * k + (l % P::SPLIT);
      const int line = l / P::SPLIT;
      if (zs) {
        float2 z = __ldg(reinterpret_cast<const float2*>(zs) + pos);
        x = pk_mul(x.x, x.y, z.x, z.y);
      }
      st_cpx(D + line * H + pos, x);
    });
}

// ----------------------------------------------------------------------------------------
// k3p: c2r along the contiguous axis, packed-Nyquist input, in place (see k3_c2r_lines)
// ----------------------------------------------------------------------------------------
struct K3PParams {
  float* buf;              // in: complex [nlines][nz/2]; out: real [nlines][nz] (same bytes)
  int ntiles;              // tiles to process (nlines / TL for the whole slab)
  int tile0;               // first tile (x-chunked runs)
  const cpx<float>* twH;   // exp(+2 pi i k / (nz/2))
  const cpx<float>* twN;   // exp(+2 pi i k / nz), k < nz/2
  const float* zscale;     // optional per-z factor (nz entries) or nullptr
  double* partials;        // optional [gridDim.x][2] (sum, sum of squares) or nullptr
};

template <int NZ>
constexpr int k3p_smem_bytes() {
  return 2 * LPlan<NZ>::kDenseBytes + LPlan<NZ>::kWorkBytes + kBarBytes;
}

template <int NZ>
__global__ void __launch_bounds__(kThreads, 1) k3p_c2r_lines(const K3PParams p) {
  using P = LPlan<NZ>;
  constexpr int H = P::H;
  using L = LayoutLinesPad<P::WP, P::M>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const dense0 = reinterpret_cast<cpx<float>*>(smem_raw);
  constexpr int kDenseElems = P::kDenseBytes / 8;
  cpx<float>* const work = reinterpret_cast<cpx<float>*>(smem_raw + 2 * P::kDenseBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + 2 * P::kDenseBytes + P::kWorkBytes);
  uint64_t* full = bars;       // [2] tile landed in dense[b]
  uint64_t* done = bars + 2;   // [2] consumers finished writing dense[b]
  const int tid = threadIdx.x;

  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_init(&done[0], kConsumers);
    mbar_init(&done[1], kConsumers);
    fence_barrier_init();
  }
  __syncthreads();

  const int first = blockIdx.x, step = gridDim.x;
  char* gbase = reinterpret_cast<char*>(p.buf) + (size_t)p.tile0 * P::kDenseBytes;
  constexpr size_t kTile = P::kDenseBytes;

  if (tid >= kConsumers) {
    // ===== copy warp: one lane drives the bulk copies =====
    if (tid == kConsumers) {
      int n = 0;
      for (int t = first; t < p.ntiles && n < 2; t += step, ++n) {
        mbar_expect_tx(&full[n], (uint32_t)kTile);
        bulk_g2s(dense0 + n * kDenseElems, gbase + (size_t)t * kTile, (uint32_t)kTile, &full[n]);
      }
      int i = 0;
      for (int t = first; t < p.ntiles; t += step, ++i) {
        const int b = i & 1;
        mbar_wait(&done[b], (i >> 1) & 1);
        bulk_s2g(gbase + (size_t)t * kTile, dense0 + b * kDenseElems, (uint32_t)kTile);
        bulk_commit();
        const int t2 = t + 2 * step;
        if (t2 < p.ntiles) {
          bulk_wait_read0();
          mbar_expect_tx(&full[b], (uint32_t)kTile);
          bulk_g2s(dense0 + b * kDenseElems, gbase + (size_t)t2 * kTile, (uint32_t)kTile, &full[b]);
        }
      }
      bulk_wait0();
    }
    return;
  }

  // ===== FFT warps =====
  double s1 = 0.0, s2 = 0.0;
  const float* zs = p.zscale;
  cpx<float> twr[P::R1];
  cpx<float> twu[K3Consts<NZ>::kPairsPerThread];
  k3_load_twiddles<NZ>(twr, twu, p.twH, p.twN, tid);
  int i = 0;
  for (int t = first; t < p.ntiles; t += step, ++i) {
    const int b = i & 1;
    cpx<float>* const D = dense0 + b * kDenseElems;
    mbar_wait(&full[b], (i >> 1) & 1);

    cpx<float> a1 = {0.f, 0.f}, a2 = {0.f, 0.f};
    k3_tile_body<NZ>(D, work, tid, twr, twu, zs, a1, a2, p.twH);
    s1 += (double)a1.x + (double)a1.y;
    s2 += (double)a2.x + (double)a2.y;
    fence_proxy_async();
    mbar_arrive(&done[b]);
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

// FFT-warp work on one k2 tile (dense [N][TZ], in place through the padded work buffer)
template <int N>
__device__ __forceinline__ void k2_tile_body(cpx<float>* __restrict__ D, cpx<float>* __restrict__ work, int tid,
                                             const cpx<float> (&twr)[SPlan<N>::R1]) {
  using P = SPlan<N>;
  using L = LayoutStridedPad<P::TZ, P::M>;
  step_a<N, P::R1, L>(work, tid, P::TZ, twr, [&](cpx<float>(&v)[P::R1], int jj, int cc) {
    const cpx<float>* Dp = D + jj * P::TZ + cc;
    sfor<0, P::R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      v[m] = ld_cpx(Dp + P::M * m * P::TZ);
    });
  });
  ConsumerSync()();
  step_b<N, P::R1, L>(work, tid, P::TZ, [&](int k, int cc, cpx<float> x) { st_cpx(D + k * P::TZ + cc, x); });
}

// ----------------------------------------------------------------------------------------
// k2p: inverse FFT along a strided axis.  Tile = [N rows][TZ columns] of one outer index.
//   load map  (4-D): (2*ncols floats, rows per source block, outer, source block)
//   store map (3-D): (2*ncols floats, N rows, outer)
// Single GPU: both maps describe the field buffer (in place; a tile reads and writes the
// same cells).  Slab-decomposed: the load map describes the all-to-all receive buffer
// [src][nx/G][ny/G][ncols], whose source blocks each hold ny/G rows of every line.
// ----------------------------------------------------------------------------------------
struct K2PParams {
  const cpx<float>* tw;    // exp(+2 pi i k / N)
  int ntiles;              // tiles to process (outer * ztiles for the whole slab)
  int tile0;               // first tile (x-chunked runs); a multiple of ztiles
  int ztiles;              // ncols / TZ
  int rows_per_src;        // rows of a line held by one source block (N when single GPU)
  int box_rows;            // rows per TMA box = min(256, rows_per_src)
  int z_out_base;          // column offset of this launch in the output (chunked runs)
  // one launch over several chunks of the chunk-major receive buffer (split schedule): tile t
  // belongs to chunk t / ck_tiles, whose source blocks follow the previous chunk's in the load
  // map's last dimension and whose columns start ck_wc further right in the output
  int ck_tiles = 0x7fffffff, ck_wc = 0, nsrc = 1;
  // tile t of the launch -> chunk, outer index, first column (within the chunk)
  __device__ __forceinline__ void decode(int t, int& ck, int& o, int& z0, int tz) const {
    ck = t / ck_tiles;
    const int r = t - ck * ck_tiles + tile0;
    o = r / ztiles;
    z0 = (r - o * ztiles) * tz;
  }
  // output addressing for kernels that store from registers (rf_big.cuh): element
  // out[outer * out_outer + row * out_row + column]
  cpx<float>* out;
  long long out_outer, out_row;
};

template <int N>
constexpr int k2p_smem_bytes() {
  return 2 * SPlan<N>::kDenseBytes + SPlan<N>::kWorkBytes + kBarBytes;
}

template <int N>
__global__ void __launch_bounds__(kThreads, 1)
k2p_strided_fft(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_out,
                const K2PParams p) {
  using P = SPlan<N>;
  using L = LayoutStridedPad<P::TZ, P::M>;
  static_assert(P::M * P::TZ == kConsumers, "one step-A item per thread");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const dense0 = reinterpret_cast<cpx<float>*>(smem_raw);
  constexpr int kDenseElems = P::kDenseBytes / 8;
  cpx<float>* const work = reinterpret_cast<cpx<float>*>(smem_raw + 2 * P::kDenseBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + 2 * P::kDenseBytes + P::kWorkBytes);
  uint64_t* full = bars;
  uint64_t* done = bars + 2;
  const int tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_init(&done[0], kConsumers);
    mbar_init(&done[1], kConsumers);
    fence_barrier_init();
  }
  __syncthreads();
  const int first = blockIdx.x, step = gridDim.x;

  if (tid >= kConsumers) {
    if (tid == kConsumers) {
      tma_prefetch_desc(&map_in);
      tma_prefetch_desc(&map_out);
      auto load_tile = [&](int t, int b) {
        int ck, o, z0;
        p.decode(t, ck, o, z0, P::TZ);
        mbar_expect_tx(&full[b], (uint32_t)P::kDenseBytes);
        for (int r0 = 0; r0 < N; r0 += p.box_rows) {
          const int src = r0 / p.rows_per_src, rl = r0 - src * p.rows_per_src;
          tma_load_4d(dense0 + b * kDenseElems + r0 * P::TZ, &map_in, &full[b], 2 * z0, rl, o,
                      ck * p.nsrc + src);
        }
      };
      int n = 0;
      for (int t = first; t < p.ntiles && n < 2; t += step, ++n) load_tile(t, n);
      int i = 0;
      for (int t = first; t < p.ntiles; t += step, ++i) {
        const int b = i & 1;
        mbar_wait(&done[b], (i >> 1) & 1);
        int ck, o, z0;
        p.decode(t, ck, o, z0, P::TZ);
        for (int r0 = 0; r0 < N; r0 += P::kBoxRows)
          tma_store_3d(&map_out, dense0 + b * kDenseElems + r0 * P::TZ,
                       2 * (z0 + p.z_out_base + ck * p.ck_wc), r0, o);
        bulk_commit();
        const int t2 = t + 2 * step;
        if (t2 < p.ntiles) {
          bulk_wait_read0();
          load_tile(t2, b);
        }
      }
      bulk_wait0();
    }
    return;
  }

  const int c = tid % P::TZ, j = tid / P::TZ;
  cpx<float> twr[P::R1];
  load_step_twiddles<P::R1>(twr, p.tw, j);
  int i = 0;
  for (int t = first; t < p.ntiles; t += step, ++i) {
    const int b = i & 1;
    cpx<float>* const D = dense0 + b * kDenseElems;
    mbar_wait(&full[b], (i >> 1) & 1);
    k2_tile_body<N>(D, work, tid, twr);
    fence_proxy_async();
    mbar_arrive(&done[b]);
    // the next tile's step A overwrites the work buffer: everyone must be out of step B
    ConsumerSync()();
  }
  (void)c;
}

// ----------------------------------------------------------------------------------------
// k1p: modes generated in registers (Philox + Box-Muller + sigma(k) + Hermitian pairing),
// inverse FFT along x, TMA store.  Tile = [NX rows][TZ columns] of one local y index.
//   store map (3-D): (2*ncols floats, ny_local, NX)
// ----------------------------------------------------------------------------------------
struct K1PParams {
  ModeGeom<float> geom;
  SigmaTable<float> tab;
  const cpx<float>* tw;    // exp(+2 pi i k / nx)
  const cpx<float>* plane; // [nx][ny_local] packed iz = 0 / nz/2 column (gen_packed_plane_kernel)
  int ny_local, y0;        // y-slab owned by this rank; y0 = first global iy
  int ntiles, ztiles;      // ny_local * ztiles, (nz/2) / TZ
  int nxl;                 // rows per destination rank (== NX when not slab-decomposed)
  int z_base;              // first global packed column of this launch (chunked runs); tile and
                           // store coordinates are relative to it
};

// one store map per destination rank: (2*ncols floats, ny_local, nx/G) over this rank's source
// block in that rank's receive buffer (peer memory over NVLink; the own block is local).
// Without slab decomposition maps[0] is the field buffer itself.
struct alignas(64) K1PMaps {
  CUtensorMap m[8];
};

template <int NX>
constexpr int k1p_smem_bytes() {
  return 2 * SPlan<NX>::kDenseBytes + SPlan<NX>::kWorkBytes + kBarBytes;
}

// No copy warp here: the kernel is issue-bound, there is nothing to load, and 8 warps (2 per
// scheduler) may use up to 255 registers where 9 may use 168.  Thread 0 issues the tile's TMA
// stores after step B and, two tiles later, checks that they have drained the buffer.
template <int NX, bool INJ, bool POT>
__global__ void __launch_bounds__(kConsumers, 1)
k1p_generate_xfft(const __grid_constant__ K1PMaps maps, const K1PParams p) {
  using P = SPlan<NX>;
  using L = LayoutStridedPad<P::TZ, P::M>;
  static_assert(P::M * P::TZ == kConsumers, "one step-A item per thread");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const dense0 = reinterpret_cast<cpx<float>*>(smem_raw);
  constexpr int kDenseElems = P::kDenseBytes / 8;
  cpx<float>* const work = reinterpret_cast<cpx<float>*>(smem_raw + 2 * P::kDenseBytes);
  const int tid = threadIdx.x;
  const int first = blockIdx.x, step = gridDim.x;
  if (tid == 0) tma_prefetch_desc(&maps.m[0]);

  const ModeGeom<float>& g = p.geom;
  const int c = tid % P::TZ, j = tid / P::TZ;
  const int nzh = g.nz / 2 + 1;
  cpx<float> twr[P::R1];
  load_step_twiddles<P::R1>(twr, p.tw, j);
  const float fj = (float)j;
  const uint64_t dq = (uint64_t)P::M * g.ny * nzh;   // Philox counter step between m and m+1
  int i = 0;
  for (int t = first; t < p.ntiles; t += step, ++i) {
    const int b = i & 1;
    cpx<float>* const D = dense0 + b * kDenseElems;
    const int iyl = t / p.ztiles, z0 = (t % p.ztiles) * P::TZ;
    const int iy = p.y0 + iyl;

    step_a<NX, P::R1, L>(work, tid, P::TZ, twr, [&](cpx<float>(&v)[P::R1], int jj, int cc) {
      const int iz = p.z_base + z0 + cc;
      if (iz == 0) {
        // packed column (iz=0 plane) + i*(iz=nz/2 plane): precomputed by gen_packed_plane_kernel
        const cpx<float>* pp = p.plane + (size_t)jj * p.ny_local + iyl;
        sfor<0, P::R1>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          v[m] = ldg_cpx(pp + (size_t)(P::M * m) * p.ny_local);
        });
      } else if constexpr (INJ) {
        sfor<0, P::R1>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          v[m] = raw_mode(g, p.tab, jj + P::M * m, iy, iz);
        });
      } else {
        // k^2 = fma(jx^2, k0x2, fma(jy^2, k0y2, jz^2*k0z2)): the inner part is per tile
        const float ksq_yz = ksq_canon(0.f, mode_sq<float>(iy, g.ny), mode_sq<float>(iz, g.nz), g.k0x2,
                                       g.k0y2, g.k0z2);
        const uint64_t q0 = ((uint64_t)jj * g.ny + iy) * nzh + iz;
        sfor<0, P::R1 / 2>([&](auto mi) {
          constexpr int m = decltype(mi)::value;
          // ix = j + M*m < nx/2 and its partner ix + nx/2 (signed mode ix - nx/2) share a counter
          const float f1 = fj + (float)(P::M * m), f2 = f1 - (float)(NX / 2);
          const float ks1 = __fmaf_rn(f1 * f1, g.k0x2, ksq_yz);
          const float ks2 = __fmaf_rn(f2 * f2, g.k0x2, ksq_yz);
          float s1 = sigma_hinge(p.tab, ks1), s2 = sigma_hinge(p.tab, ks2);
          if constexpr (POT) {
            s1 = __fdividef(s1, ks1);   // k^2 > 0 off the iz = 0 plane
            s2 = __fdividef(s2, ks2);
          }
          const uint64_t q = q0 + (uint64_t)m * dq;
          philox_out o = philox4x32_10((uint32_t)q, (uint32_t)(q >> 32), 0u, 0u, g.key0, g.key1);
          cpx<float> n1 = box_muller<float>(o.w[0], o.w[1]);
          cpx<float> n2 = box_muller<float>(o.w[2], o.w[3]);
          v[m] = pk_mul(n1.x, n1.y, s1, s1);
          v[m + P::R1 / 2] = pk_mul(n2.x, n2.y, s2, s2);
        });
      }
    });
    // buffer b was last stored from two tiles ago: at most the previous tile's store may
    // still be reading shared memory
    if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    __syncthreads();
    step_b<NX, P::R1, L>(work, tid, P::TZ, [&](int k, int cc, cpx<float> x) { st_cpx(D + k * P::TZ + cc, x); });
    fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
      // boxes never straddle two destinations (box rows = min(256, nxl), both powers of two)
      const int box = p.nxl < P::kBoxRows ? p.nxl : P::kBoxRows;
      for (int r0 = 0; r0 < NX; r0 += box) {
        const int d = r0 / p.nxl;
        tma_store_3d(&maps.m[d], D + r0 * P::TZ, 2 * z0, iyl, r0 - d * p.nxl);
      }
      bulk_commit();
    }
  }
  if (tid == 0) bulk_wait0();
  (void)c;
}

// ----------------------------------------------------------------------------------------
// k1q: the same pass as k1p laid out for TWO resident CTAs per SM (16 warps instead of 8).
// k1p is issue-bound with two warps per scheduler and long dependent chains (Philox rounds,
// MUFU); it reaches 60 % issue utilisation.  Here one CTA needs a single padded buffer
// (66 KB) and at most 128 registers per thread:
//   * step A writes the padded work buffer straight from registers (as in k1p);
//   * step B is IN PLACE: every thread reads its R2-point sub-transform(s) out of the work
//     buffer, all threads synchronise, and the results go back into the same memory in the
//     dense [NX][TZ] layout the TMA store wants -- no separate store ring;
//   * the step twiddles W_NX^(j*k1) come from a shared-memory copy of the table instead of
//     62 registers; the sigma(k) cell table is staged in shared memory as well;
//   * the buffer is only written again after the next tile's modes have been generated in
//     registers (most of the tile time), so the TMA store read-out is hidden.
// While one CTA of an SM generates (ALU/XU heavy) the other usually transforms (FMA heavy).
// ----------------------------------------------------------------------------------------
template <int NX>
constexpr int k1q_smem_bytes(int hinge_cells) {
  return SPlan<NX>::kWorkBytes + NX * 8 + hinge_cells * 16;
}

template <int NX, bool INJ, bool POT>
__global__ void __launch_bounds__(kConsumers, 2)
k1q_generate_xfft(const __grid_constant__ K1PMaps maps, const K1PParams p) {
  using P = SPlan<NX>;
  using L = LayoutStridedPad<P::TZ, P::M>;
  constexpr int R1 = P::R1, R2 = P::R2, M = P::M, TZ = P::TZ;
  static_assert(M * TZ == kConsumers, "one step-A item per thread");
  static_assert(R1 % R2 == 0, "step B items per thread");
  constexpr int IB = R1 / R2;          // step-B items per thread (R1*TZ items / 256 threads)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const work = reinterpret_cast<cpx<float>*>(smem_raw);
  cpx<float>* const tws = reinterpret_cast<cpx<float>*>(smem_raw + P::kWorkBytes);
  float4* const hinge_s = reinterpret_cast<float4*>(smem_raw + P::kWorkBytes + NX * 8);
  const int tid = threadIdx.x;
  const int first = blockIdx.x, step = gridDim.x;
  if (tid == 0) tma_prefetch_desc(&maps.m[0]);
  // tws[k1 * M + j] = W_NX^(j * k1): the lanes of a warp (consecutive j) read consecutive words
  for (int q = tid; q < NX; q += kConsumers) tws[q] = ldg_cpx(p.tw + (q % M) * (q / M));
  if constexpr (!INJ) {
    for (int q = tid; q < p.tab.lut_n; q += kConsumers)
      hinge_s[q] = __ldg(reinterpret_cast<const float4*>(p.tab.hinge) + q);
  }
  __syncthreads();
  const float4* const hinge0 = hinge_s - p.tab.lut_base;   // entry of cell id 0 (never dereferenced)
  const int shift = p.tab.shift;

  const ModeGeom<float>& g = p.geom;
  const int c = tid % TZ, j = tid / TZ;
  const int nzh = g.nz / 2 + 1;
  const float fj = (float)j;
  const uint64_t dq = (uint64_t)M * g.ny * nzh;   // Philox counter step between m and m+1
  cpx<float>* const wa = work + L::idx2(0, j, c);
  for (int t = first; t < p.ntiles; t += step) {
    const int iyl = t / p.ztiles, z0 = (t % p.ztiles) * TZ;
    const int iy = p.y0 + iyl;
    const int iz = p.z_base + z0 + c;
    cpx<float> v[R1];
    if (iz == 0) {
      // packed column (iz=0 plane) + i*(iz=nz/2 plane): precomputed by gen_packed_plane_kernel
      const cpx<float>* pp = p.plane + (size_t)j * p.ny_local + iyl;
      sfor<0, R1>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[m] = ldg_cpx(pp + (size_t)(M * m) * p.ny_local);
      });
    } else if constexpr (INJ) {
      sfor<0, R1>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[m] = raw_mode(g, p.tab, j + M * m, iy, iz);
      });
    } else {
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
          s1 = __fdividef(s1, ks1);   // k^2 > 0 off the iz = 0 plane
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
    // the previous tile's TMA store must have finished reading the buffer
    if (tid == 0) bulk_wait_read0();
    __syncthreads();
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x = v[brev<R1>(k1)];
      if constexpr (k1 > 0) x = cmul(x, ld_cpx(tws + k1 * M + j));
      st_cpx(wa + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x);
    });
    __syncthreads();
    // step B in place: item w = tid + q*256 -> (k1, column) = (w / TZ, w % TZ)
    cpx<float> u[IB][R2];
    sfor<0, IB>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int w = tid + q * kConsumers;
      const cpx<float>* wp = work + L::idx2(w / TZ, 0, w % TZ);
      sfor<0, R2>([&](auto ji) {
        constexpr int jj = decltype(ji)::value;
        u[q][jj] = ld_cpx(wp + (L::idx2(0, jj, 0) - L::idx2(0, 0, 0)));
      });
    });
    __syncthreads();
    sfor<0, IB>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int w = tid + q * kConsumers;
      const int k1 = w / TZ, cc = w % TZ;
      fft_dif<R2, true>(u[q]);
      sfor<0, R2>([&](auto ki) {
        constexpr int k2 = decltype(ki)::value;
        st_cpx(work + (k1 + R1 * k2) * TZ + cc, u[q][brev<R2>(k2)]);
      });
    });
    fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
      // boxes never straddle two destinations (box rows = min(256, nxl), both powers of two)
      const int box = p.nxl < P::kBoxRows ? p.nxl : P::kBoxRows;
      for (int r0 = 0; r0 < NX; r0 += box) {
        const int d = r0 / p.nxl;
        tma_store_3d(&maps.m[d], work + r0 * TZ, 2 * z0, iyl, r0 - d * p.nxl);
      }
      bulk_commit();
    }
  }
  if (tid == 0) bulk_wait0();
}

// ----------------------------------------------------------------------------------------
// k1r: k1q's tile body run by TWO 256-thread groups of one 512-thread CTA that take turns
// ("ping-pong").  A tile has two phases that want different pipes: GENERATE (Philox = IMAD.WIDE,
// measured at 1.6 warp instructions per clock and SM on B200; Box-Muller and sigma(k) = MUFU at
// 16 lanes per clock) and TRANSFORM (radix-32 butterflies on the FP32 pipe, four passes through
// shared memory).  Two independent CTAs per SM (k1q) start together and, contending
// symmetrically, stay in lockstep: both generate, then both transform, and every pipe idles half
// of the time (ncu: no pipe above 35 %, issue slots 56 % busy, yet 2 CTAs per SM are only 11 %
// faster than one).  Here the two groups hand a token back and forth through named barriers:
// a group generates only while it holds the token and passes it on before it transforms, so one
// group's GENERATE always overlaps the other's TRANSFORM.
//   bar 1 + g : the 256 threads of group g (the __syncthreads of k1q)
//   bar 3 + g : group g may generate (bar.sync by group g, bar.arrive by the other group)
// ----------------------------------------------------------------------------------------
constexpr int kGroupsR = 2;
constexpr int kThreadsR = kGroupsR * kConsumers;

template <int NX>
constexpr int k1r_smem_bytes(int hinge_cells) {
  return kGroupsR * SPlan<NX>::kWorkBytes + NX * 8 + hinge_cells * 16;
}

__device__ __forceinline__ void named_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}

template <int NX, bool INJ, bool POT>
__global__ void __launch_bounds__(kThreadsR, 1)
k1r_generate_xfft(const __grid_constant__ K1PMaps maps, const K1PParams p) {
  using P = SPlan<NX>;
  using L = LayoutStridedPad<P::TZ, P::M>;
  constexpr int R1 = P::R1, R2 = P::R2, M = P::M, TZ = P::TZ;
  static_assert(M * TZ == kConsumers, "one step-A item per thread");
  static_assert(R1 % R2 == 0, "step B items per thread");
  constexpr int IB = R1 / R2;          // step-B items per thread (R1*TZ items / 256 threads)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int grp = threadIdx.x / kConsumers, tid = threadIdx.x % kConsumers;
  cpx<float>* const work = reinterpret_cast<cpx<float>*>(smem_raw + grp * P::kWorkBytes);
  cpx<float>* const tws = reinterpret_cast<cpx<float>*>(smem_raw + kGroupsR * P::kWorkBytes);
  float4* const hinge_s = reinterpret_cast<float4*>(smem_raw + kGroupsR * P::kWorkBytes + NX * 8);
  if (threadIdx.x == 0) tma_prefetch_desc(&maps.m[0]);
  // tws[k1 * M + j] = W_NX^(j * k1): the lanes of a warp (consecutive j) read consecutive words
  for (int q = threadIdx.x; q < NX; q += kThreadsR) tws[q] = ldg_cpx(p.tw + (q % M) * (q / M));
  if constexpr (!INJ) {
    for (int q = threadIdx.x; q < p.tab.lut_n; q += kThreadsR)
      hinge_s[q] = __ldg(reinterpret_cast<const float4*>(p.tab.hinge) + q);
  }
  __syncthreads();
  const float4* const hinge0 = hinge_s - p.tab.lut_base;   // entry of cell id 0 (never dereferenced)
  const int shift = p.tab.shift;
  const int bar_grp = 1 + grp, bar_mine = 3 + grp, bar_other = 3 + (1 - grp);
  // group 0 holds the token first
  if (grp == 1) named_arrive(3, kThreadsR);

  const ModeGeom<float>& g = p.geom;
  const int c = tid % TZ, j = tid / TZ;
  const int nzh = g.nz / 2 + 1;
  const float fj = (float)j;
  const uint64_t dq = (uint64_t)M * g.ny * nzh;   // Philox counter step between m and m+1
  cpx<float>* const wa = work + L::idx2(0, j, c);
  // tiles of this CTA: blockIdx.x + k * gridDim.x; group g takes k = g, g + 2, ...
  const int step = gridDim.x;
  for (int t = blockIdx.x + grp * step; t < p.ntiles; t += kGroupsR * step) {
    const int iyl = t / p.ztiles, z0 = (t % p.ztiles) * TZ;
    const int iy = p.y0 + iyl;
    const int iz = p.z_base + z0 + c;
    cpx<float> v[R1];
    named_sync(bar_mine, kThreadsR);              // ---- GENERATE: this group holds the token ----
    if (iz == 0) {
      // packed column (iz=0 plane) + i*(iz=nz/2 plane): precomputed by gen_packed_plane_kernel
      const cpx<float>* pp = p.plane + (size_t)j * p.ny_local + iyl;
      sfor<0, R1>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[m] = ldg_cpx(pp + (size_t)(M * m) * p.ny_local);
      });
    } else if constexpr (INJ) {
      sfor<0, R1>([&](auto mi) {
        constexpr int m = decltype(mi)::value;
        v[m] = raw_mode(g, p.tab, j + M * m, iy, iz);
      });
    } else {
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
          s1 = __fdividef(s1, ks1);   // k^2 > 0 off the iz = 0 plane
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
    // ---- TRANSFORM: pass the token on if the other group has a tile left ----
    if (t + step < p.ntiles) named_arrive(bar_other, kThreadsR);
    fft_dif<R1, true>(v);
    // the previous tile's TMA store must have finished reading this group's buffer
    if (tid == 0) bulk_wait_read0();
    named_sync(bar_grp, kConsumers);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x = v[brev<R1>(k1)];
      if constexpr (k1 > 0) x = cmul(x, ld_cpx(tws + k1 * M + j));
      st_cpx(wa + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x);
    });
    named_sync(bar_grp, kConsumers);
    // step B in place: item w = tid + q*256 -> (k1, column) = (w / TZ, w % TZ)
    cpx<float> u[IB][R2];
    sfor<0, IB>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int w = tid + q * kConsumers;
      const cpx<float>* wp = work + L::idx2(w / TZ, 0, w % TZ);
      sfor<0, R2>([&](auto ji) {
        constexpr int jj = decltype(ji)::value;
        u[q][jj] = ld_cpx(wp + (L::idx2(0, jj, 0) - L::idx2(0, 0, 0)));
      });
    });
    named_sync(bar_grp, kConsumers);
    sfor<0, IB>([&](auto qi) {
      constexpr int q = decltype(qi)::value;
      const int w = tid + q * kConsumers;
      const int k1 = w / TZ, cc = w % TZ;
      fft_dif<R2, true>(u[q]);
      sfor<0, R2>([&](auto ki) {
        constexpr int k2 = decltype(ki)::value;
        st_cpx(work + (k1 + R1 * k2) * TZ + cc, u[q][brev<R2>(k2)]);
      });
    });
    fence_proxy_async();
    named_sync(bar_grp, kConsumers);
    if (tid == 0) {
      // boxes never straddle two destinations (box rows = min(256, nxl), both powers of two)
      const int box = p.nxl < P::kBoxRows ? p.nxl : P::kBoxRows;
      for (int r0 = 0; r0 < NX; r0 += box) {
        const int d = r0 / p.nxl;
        tma_store_3d(&maps.m[d], work + r0 * TZ, 2 * z0, iyl, r0 - d * p.nxl);
      }
      bulk_commit();
    }
  }
  if (tid == 0) bulk_wait0();
}

// ----------------------------------------------------------------------------------------
// Tile schedule of the fused y + z pass (k23q_fused_yz, rf_k23q.cuh): the y pass and the z (c2r)
// pass in ONE persistent kernel, so that the y-pass output of an x-plane is still in L2 when the
// z pass reads it (the plane's tiles are visited a few planes later, and the 126 MB L2 is
// write-back): DRAM sees 8 B/cell for both passes instead of 16.  Tiles are handed out by a
// global ticket in a fixed order
//     K2(plane 0) ... K2(plane lag-1), K2(plane lag), K3(plane 0), K2(plane lag+1), K3(plane 1), ...
// A K3 tile waits until the ztiles K2 tiles of its plane have been stored (plane_done counter).
// It only depends on smaller tickets and every CTA works through its tickets in increasing
// order, so the smallest unfinished ticket can always proceed.
// ----------------------------------------------------------------------------------------
struct K23Params {
  K2PParams k2;               // ntiles / tile0 unused
  K3PParams k3;               // ntiles / tile0 unused
  int nplanes;                // local x planes
  int t2, t3;                 // K2 / K3 tiles per plane
  int lag;                    // planes between the y pass and the z pass (<= nplanes)
  unsigned int* ticket;       // zeroed before launch
  unsigned int* plane_done;   // [nplanes] zeroed before launch
};

enum { kTileK2 = 0, kTileK3 = 1, kTileEnd = 2 };

__device__ __forceinline__ void k23_decode(const K23Params& p, unsigned int ticket, int& type, int& plane,
                                           int& idx) {
  const unsigned int head = (unsigned)p.lag * p.t2, per = (unsigned)(p.t2 + p.t3);
  const unsigned int mid = (unsigned)(p.nplanes - p.lag) * per;
  const unsigned int total = (unsigned)p.nplanes * per;
  if (ticket >= total) { type = kTileEnd; plane = idx = 0; return; }
  if (ticket < head) { type = kTileK2; plane = ticket / p.t2; idx = ticket % p.t2; return; }
  unsigned int u = ticket - head;
  if (u < mid) {
    const int s = p.lag + u / per, r = u % per;
    if (r < p.t2) { type = kTileK2; plane = s; idx = r; }
    else { type = kTileK3; plane = s - p.lag; idx = r - p.t2; }
    return;
  }
  u -= mid;
  type = kTileK3; plane = p.nplanes - p.lag + u / p.t3; idx = u % p.t3;
}

// @@PIPE_END@@
}  // namespace pipe
}  // namespace rf
