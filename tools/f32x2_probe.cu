// f32x2_probe.cu -- issue rate of scalar vs packed FP32 on sm_100a (B200).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_probe f32x2_probe.cu && ./f32x2_probe
// Each thread runs NCH independent dependency chains of one instruction kind; the kernel is
// launched with `warps` warps per SM sub-partition, and the rate is reported in warp
// instructions per clock per SM (4 schedulers: 4.0 = one instruction per scheduler per clock).
#include <cstdio>
#include <cuda_runtime.h>

template <int KIND, int NCH>
__global__ void probe(float* out, int iters, float a, float b) {
  float2 v[NCH];
#pragma unroll
  for (int i = 0; i < NCH; ++i) v[i] = make_float2(threadIdx.x + i, blockIdx.x - i);
  const float2 A = make_float2(a, a * 1.0001f), B = make_float2(b, b * 0.9999f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      if (KIND == 0) { v[i].x = v[i].x + A.x; v[i].y = v[i].y + A.y; }                       // 2 FADD
      else if (KIND == 1) v[i] = __fadd2_rn(v[i], A);                                         // 1 FADD2
      else if (KIND == 2) { v[i].x = __fmaf_rn(v[i].x, A.x, B.x); v[i].y = __fmaf_rn(v[i].y, A.y, B.y); }  // 2 FFMA
      else if (KIND == 3) v[i] = __ffma2_rn(v[i], A, B);                                      // 1 FFMA2
      else if (KIND == 4) { v[i].x = v[i].x * A.x; v[i].y = v[i].y * A.y; }                  // 2 FMUL
      else if (KIND == 5) v[i] = __fmul2_rn(v[i], A);                                         // 1 FMUL2
      else if (KIND == 6) {                                                                    // 2 IMAD.WIDE.U32 (Philox round)
        unsigned long long p = (unsigned long long)0xD2511F53u * __float_as_uint(v[i].x);
        unsigned long long q = (unsigned long long)0xCD9E8D57u * __float_as_uint(v[i].y);
        v[i].x = __uint_as_float((unsigned)(q >> 32) ^ (unsigned)p);
        v[i].y = __uint_as_float((unsigned)(p >> 32) ^ (unsigned)q);
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < NCH; ++i) s += v[i].x + v[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int KIND>
static void run(const char* name, int per_pair) {
  int dev = 0, sms = 0, khz = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  float* out;
  cudaMalloc(&out, sizeof(float) * sms * 1024);
  constexpr int NCH = 8;
  const int iters = 20000;
  for (int warps : {4, 8, 16, 32}) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<KIND, NCH><<<sms, warps * 32>>>(out, 100, 1.0001f, 0.5f);
    cudaEventRecord(e0);
    probe<KIND, NCH><<<sms, warps * 32>>>(out, iters, 1.0001f, 0.5f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double inst = (double)warps * iters * NCH * per_pair;   // warp instructions per SM
    const double clk = ms * 1e-3 * khz * 1e3;
    printf("%-10s warps/SM %2d: %.3f ms, %.2f warp-inst/clk/SM (nominal clock), %.1f pair-ops/clk/SM\n", name, warps, ms,
           inst / clk, (double)warps * iters * NCH * 32 / clk);
  }
  cudaFree(out);
}

int main() {
  run<0>("FADD", 2);
  run<1>("FADD2", 1);
  run<2>("FFMA", 2);
  run<3>("FFMA2", 1);
  run<4>("FMUL", 2);
  run<5>("FMUL2", 1);
  run<6>("IMAD.WIDE", 2);
  return 0;
}
