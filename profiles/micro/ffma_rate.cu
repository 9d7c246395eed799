// How fast is packed fp32 FMA (fma.rn.f32x2, SASS FFMA2) on B200 against scalar FFMA, and do the two
// mix?  The weight-stream forward and the conv kernels are FP32-pipe-bound; ncu shows
// `math_pipe_throttle` as their top stall at 50-58 % "FMA pipe active", which raises the question what
// 100 % means for FFMA2.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/micro/ffma_rate profiles/micro/ffma_rate.cu
//
// Every thread runs ITER rounds over NACC independent accumulator chains (enough ILP to hide latency);
// modes: 0 = FFMA only, 1 = FFMA2 only, 2 = alternating FFMA2 / FFMA (2:1 in flops).  Reports FMA
// (multiply-adds, not instructions) per clock per SM for 1, 2, 4, 8, 16 warps per scheduler-quad.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int NACC = 16;
constexpr int ITER = 4096;

template <int MODE>
__global__ void rate_kernel(float* out, float seed, long long* cycles) {
  float2 a2[NACC];
  float a1[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { a2[i] = make_float2(seed + i, seed - i); a1[i] = seed * i; }
  const float2 m2 = make_float2(1.0001f, 0.9999f), c2 = make_float2(seed, -seed);
  const float m1 = 1.0002f, c1 = seed * 0.5f;
  const long long t0 = clock64();
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      if (MODE == 0) { a1[i] = __fmaf_rn(a1[i], m1, c1); }
      if (MODE == 1) { a2[i] = __ffma2_rn(a2[i], m2, c2); }
      if (MODE == 2) { a2[i] = __ffma2_rn(a2[i], m2, c2); a1[i] = __fmaf_rn(a1[i], m1, c1); }
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += a2[i].x + a2[i].y + a1[i];
  if (s == 12345.678f) out[0] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
}

template <int MODE>
static double run(int threads, float* out, long long* cyc) {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  rate_kernel<MODE><<<sms, threads>>>(out, 1.5f, cyc);
  CK(cudaDeviceSynchronize());
  rate_kernel<MODE><<<sms, threads>>>(out, 1.5f, cyc);
  CK(cudaDeviceSynchronize());
  long long c = 0;
  CK(cudaMemcpy(&c, cyc, sizeof c, cudaMemcpyDeviceToHost));
  const double fma_per_thread = (double)ITER * NACC * (MODE == 0 ? 1 : MODE == 1 ? 2 : 3);
  return fma_per_thread * threads / (double)c;            // FMA per clock per SM
}

int main() {
  float* out;
  long long* cyc;
  CK(cudaMalloc(&out, 4));
  CK(cudaMalloc(&cyc, 8));
  printf("FMA (multiply-adds) per clock per SM; 128 = one FMA per lane per clock on 4 schedulers x 32 lanes\n");
  printf("%8s %12s %12s %16s\n", "threads", "FFMA", "FFMA2", "FFMA2+FFMA 1:1");
  for (int threads : {128, 256, 512, 1024}) {
    printf("%8d %12.1f %12.1f %16.1f\n", threads, run<0>(threads, out, cyc), run<1>(threads, out, cyc), run<2>(threads, out, cyc));
  }
  return 0;
}
