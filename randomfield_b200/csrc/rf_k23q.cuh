// rf_k23q.cuh -- the y pass and the z (c2r) pass in ONE persistent kernel, three-deep tile ring.
//
// Tile order of K23Params (rf_pipe.cuh): the z pass of an x-plane runs a few planes after its y
// pass and reads the y-pass output from the write-back L2, so DRAM sees 8 B/cell for both passes
// instead of 16.  A first version (round 1, removed) lost to the separate kernels because its ring
// was two dense buffers plus a shared work buffer: store read-out, dependency probe and the next
// load were serial on one buffer while the other was being transformed.  Here both tile bodies
// run IN PLACE in their own buffer (k2q_tile_body below, k3q_tile_body of rf_k3q.cuh: read all,
// barrier, write all), the work buffer goes away and three 66 KB buffers fit: one loading, one
// being transformed, one draining.  Same arithmetic in the same order as k2p / k3q.
#pragma once
#include "rf_k3q.cuh"

#ifndef RF_K23Q_SIGLAG
#define RF_K23Q_SIGLAG 4
#endif

namespace rf {
namespace pipe {

// one y-pass tile, in place in `B` (dense [N rows][TZ columns] at the start of a kWorkBytes buffer)
template <int N>
__device__ __forceinline__ void k2q_tile_body(cpx<float>* __restrict__ B, int tid,
                                              const cpx<float> (&twr)[SPlan<N>::R1]) {
  using P = SPlan<N>;
  using L = LayoutStridedPad<P::TZ, P::M>;
  constexpr int R1 = P::R1, R2 = P::R2, M = P::M, TZ = P::TZ, IB = R1 / R2;
  static_assert(M * TZ == kConsumers && R1 % R2 == 0 && R1 * R2 == N, "plan");
  const int c = tid % TZ, j = tid / TZ;
  cpx<float> v[R1];
  {
    const cpx<float>* Dp = B + j * TZ + c;
    sfor<0, R1>([&](auto mi) {
      constexpr int m = decltype(mi)::value;
      v[m] = ld_cpx(Dp + M * m * TZ);
    });
  }
  ConsumerSync()();
  fft_dif<R1, true>(v);
  {
    cpx<float>* wa = B + L::idx2(0, j, c);
    sfor<0, R1>([&](auto ki) {
      constexpr int k1 = decltype(ki)::value;
      cpx<float> x = v[brev<R1>(k1)];
      if constexpr (k1 > 0) x = cmul(x, twr[k1]);
      st_cpx(wa + (L::idx2(k1, 0, 0) - L::idx2(0, 0, 0)), x);
    });
  }
  ConsumerSync()();
  cpx<float> u[IB][R2];
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kConsumers;
    const cpx<float>* wp = B + L::idx2(w / TZ, 0, w % TZ);
    sfor<0, R2>([&](auto ji) {
      constexpr int jj = decltype(ji)::value;
      u[q][jj] = ld_cpx(wp + (L::idx2(0, jj, 0) - L::idx2(0, 0, 0)));
    });
  });
  ConsumerSync()();
  sfor<0, IB>([&](auto qi) {
    constexpr int q = decltype(qi)::value;
    const int w = tid + q * kConsumers;
    const int k1 = w / TZ, cc = w % TZ;
    fft_dif<R2, true>(u[q]);
    sfor<0, R2>([&](auto ki) {
      constexpr int k2 = decltype(ki)::value;
      st_cpx(B + (k1 + R1 * k2) * TZ + cc, u[q][brev<R2>(k2)]);
    });
  });
}

template <int NY, int NZ>
__host__ __device__ constexpr int k23q_buf_bytes() {
  constexpr int a = SPlan<NY>::kWorkBytes, b = LPlan<NZ>::kWorkBytes;
  return ((a > b ? a : b) + 127) / 128 * 128;
}
template <int NY, int NZ>
constexpr int k23q_smem_bytes() {
  // tile ring, barriers + tile descriptors, twiddle tables (NY + 2*(NZ/2) entries)
  return 3 * k23q_buf_bytes<NY, NZ>() + 2 * kBarBytes + (NY + NZ) * 8;
}

// one non-blocking look at an mbarrier phase
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// Two producer warps: the LOADER (decode the next ticket, probe the plane counter, issue the TMA
// load) and the STORER (TMA store, wait for its read-out, hand the buffer back, signal finished
// planes).  In the first version one thread did all of that in sequence, and that chain of long-latency
// steps (read-out wait, acquire probe, release signal: about 3 us) was longer than a tile's
// butterflies.
constexpr int kThreadsQ = kConsumers + 64;

template <int NY, int NZ>
__global__ void __launch_bounds__(kThreadsQ, 1)
k23q_fused_yz(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_out,
              const K23Params p) {
  using P2 = SPlan<NY>;
  using P3 = LPlan<NZ>;
  static_assert(P3::SPLIT == 1, "k23q: NZ <= 2048");
  constexpr int kBuf = k23q_buf_bytes<NY, NZ>();
  constexpr int kBufElems = kBuf / 8;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  cpx<float>* const buf0 = reinterpret_cast<cpx<float>*>(smem_raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + 3 * kBuf);
  uint64_t* full = bars;        // [3] tile landed (or end marker published)
  uint64_t* done = bars + 3;    // [3] FFT warps finished the tile
  uint64_t* empty = bars + 6;   // [3] the tile's store has been read out of the buffer
  volatile int* desc = reinterpret_cast<volatile int*>(bars + 9);   // [3][4]: type, plane, idx
  // (the acquire loads / release reductions of the producer warps invalidate L1, so the per-tile
  // twiddle reloads of the FFT warps come from shared-memory copies of the tables)
  cpx<float>* const tw2s = reinterpret_cast<cpx<float>*>(smem_raw + 3 * kBuf + 2 * kBarBytes);
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
  // step-twiddle tables in [k1][j] order, W^(j k1) at [k1 * M + j]: the lanes of a warp (consecutive
  // j) read consecutive words -- indexed as W[j * k1] the per-tile reloads cost 15-20 % of all
  // shared-memory wavefronts in bank conflicts (ncu, round 2)
  for (int q = tid; q < NY; q += kThreadsQ) tw2s[q] = ldg_cpx(p.k2.tw + (q % P2::M) * (q / P2::M));
  for (int q = tid; q < NZ / 2; q += kThreadsQ) {
    twHs[q] = ldg_cpx(p.k3.twH + (q % P3::M) * (q / P3::M));
    twNs[q] = ldg_cpx(p.k3.twN + q);
  }
  __syncthreads();
  char* const gbase = reinterpret_cast<char*>(p.k3.buf);

  if (tid >= kConsumers) {
    if (tid == kConsumers) {
      // ===== loader =====
      tma_prefetch_desc(&map_in);
      // Tickets come from a global counter, fetched one tile ahead (the atomic's round trip is
      // off the critical path).  Unlike a static blockIdx-strided schedule this needs no
      // co-residency of all CTAs: a ticket is only ever held by a running CTA, a z-pass tile
      // depends on smaller tickets only, and every CTA works through its tickets in increasing
      // order -- so the smallest unfinished ticket can always proceed.
      unsigned int ticket = atomicAdd(p.ticket, 1u);
      int b = 0, ph = 0;
      for (int n = 0;; ++n) {
        const unsigned int ticket_next = atomicAdd(p.ticket, 1u);
        if (n >= 3) mbar_wait(&empty[b], ph ^ 1);   // (ph flips every 3 tiles: phase of tile n - 3)
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
          for (int r0 = 0; r0 < NY; r0 += p.k2.box_rows) {
            const int src = r0 / p.k2.rows_per_src, rl = r0 - src * p.k2.rows_per_src;
            tma_load_4d(dst + r0 * P2::TZ, &map_in, &full[b], 2 * z0, rl, plane, src);
          }
        } else {
          // the plane's y pass must be complete.  Blocking here is safe: every earlier tile of this
          // CTA is already loaded and is transformed, stored and signalled without the loader, and
          // a z-pass tile only depends on smaller tickets.
          for (;;) {
            unsigned int have;
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(have) : "l"(p.plane_done + plane) : "memory");
            if (have >= (unsigned)p.t2) break;
            __nanosleep(200);
          }
          asm volatile("fence.proxy.async;" ::: "memory");
          const size_t off = ((size_t)plane * p.t3 + idx) * (size_t)P3::kDenseBytes;
          mbar_expect_tx(&full[b], (uint32_t)P3::kDenseBytes);
          bulk_g2s(dst, gbase + off, (uint32_t)P3::kDenseBytes, &full[b]);
        }
        if (++b == 3) { b = 0; ph ^= 1; }
      }
    } else if (tid == kConsumers + 32) {
      // ===== storer =====
      tma_prefetch_desc(&map_out);
      // planes of the last kSigLag y-pass stores, not yet signalled: a store's completion is only
      // waited for kSigLag tiles later -- or at once when the storer would otherwise sit idle
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
      // before sitting idle: complete and signal what is pending (other CTAs may be waiting for
      // exactly these planes, and this CTA's next tile may in turn be waiting for theirs)
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
        if (type == kTileK2) {
          const int z0 = idx * P2::TZ;
          for (int r0 = 0; r0 < NY; r0 += P2::kBoxRows)
            tma_store_3d(&map_out, buf0 + b * kBufElems + r0 * P2::TZ, 2 * z0, r0, plane);
        } else {
          const size_t off = ((size_t)plane * p.t3 + idx) * (size_t)P3::kDenseBytes;
          bulk_s2g(gbase + off, buf0 + b * kBufElems, (uint32_t)P3::kDenseBytes);
        }
        bulk_commit();
        bulk_wait_read0();
        mbar_arrive(&empty[b]);
        // the store committed kSigLag tiles ago is complete once at most kSigLag groups are pending
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

  // ===== FFT warps =====
  double s1 = 0.0, s2 = 0.0;
  const float* zs = p.k3.zscale;
  int b = 0, ph = 0;
  for (;;) {
    cpx<float>* const B = buf0 + b * kBufElems;
    mbar_wait(&full[b], ph);
    const int type = desc[4 * b + 0];
    if (type == kTileEnd) break;
    if (type == kTileK2) {
      cpx<float> twr[P2::R1];
      sfor<0, P2::R1>([&](auto ki) {
        constexpr int k1 = decltype(ki)::value;
        twr[k1] = ld_cpx(tw2s + k1 * P2::M + tid / P2::TZ);
      });
      k2q_tile_body<NY>(B, tid, twr);
    } else {
      cpx<float> twr[P3::R1], twr2[P3::R1];
      cpx<float> twu[K3Consts<NZ>::kPairsPerThread];
      k3q_load_twiddles<NZ, true>(twr, twr2, twu, twHs, twNs, tid, [](const cpx<float>* q) { return ld_cpx(q); });
      cpx<float> a1 = {0.f, 0.f}, a2 = {0.f, 0.f};
      k3q_tile_body<NZ>(B, tid, twr, twr2, twu, zs, a1, a2, p.k3.twH);
      s1 += (double)a1.x + (double)a1.y;
      s2 += (double)a2.x + (double)a2.y;
    }
    fence_proxy_async();
    mbar_arrive(&done[b]);
    if (++b == 3) { b = 0; ph ^= 1; }
  }

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
