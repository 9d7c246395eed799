// Does B200's L2 keep the most recently read part of a 108 MB buffer across ~45-100 MB of other traffic
// (LRU-like), so that a second pass in REVERSE time order hits it?  Pass A: read W front to back
// (all CTAs sweep the same band at the same time).  Pass T: read+write `trash` MB of another
// buffer.  Pass B: read+write W front-to-back or back-to-front.  Prints pass B's time for both.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_order l2_order.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) sweep(float4* buf, size_t n4, int reverse, int write, float* sink) {
  float acc = 0.f;
  const size_t per_iter = (size_t)gridDim.x * blockDim.x * 4;
  const size_t iters = (n4 + per_iter - 1) / per_iter;
  for (size_t it = 0; it < iters; ++it) {
    const size_t band = reverse ? iters - 1 - it : it;
    const size_t i = band * per_iter + (size_t)blockIdx.x * blockDim.x * 4 + threadIdx.x;
    float4 v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const size_t j = i + (size_t)q * blockDim.x;
      v[q] = j < n4 ? buf[j] : make_float4(0, 0, 0, 0);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const size_t j = i + (size_t)q * blockDim.x;
      acc += v[q].x + v[q].w;
      if (write && j < n4) { v[q].x += 1.f; buf[j] = v[q]; }
    }
  }
  if (acc == 123.456f) *sink = acc;
}

int main() {
  const size_t MB = 1 << 20, wsz = 108 * MB;
  float4 *W, *M;
  float* sink;
  CK(cudaMalloc(&W, wsz)); CK(cudaMalloc(&M, 256 * MB)); CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(W, 0, wsz)); CK(cudaMemset(M, 0, 256 * MB));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = 148 * 2;
  for (size_t trash : {0, 20, 45, 80, 216})
    for (int bwrite = 0; bwrite < 2; ++bwrite)
      for (int rev = 0; rev < 2; ++rev) {
        float tb = 0;
        const int reps = 6;
        for (int r = 0; r < reps + 2; ++r) {
          sweep<<<grid, 256>>>(W, wsz / 16, 0, 0, sink);
          if (trash) sweep<<<grid, 256>>>(M, trash * MB / 16, 0, 1, sink);
          cudaEventRecord(e0);
          sweep<<<grid, 256>>>(W, wsz / 16, rev, bwrite, sink);
          cudaEventRecord(e1);
          CK(cudaEventSynchronize(e1));
          float a;
          cudaEventElapsedTime(&a, e0, e1);
          if (r >= 2) tb += a;
        }
        printf("trash %3zu MB, pass B %s %s: %.1f us\n", trash, bwrite ? "r+w" : "r  ", rev ? "reverse" : "forward", tb / reps * 1e3);
      }
  return 0;
}
