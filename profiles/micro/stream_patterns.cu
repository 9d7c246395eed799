// Micro-benchmark: achieved HBM read bandwidth of cp.async streaming with the access patterns the
// split-NN kernels use (rows x bytes-per-row per warp step, row stride 21,632 B) vs a linear read.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_patterns stream_patterns.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// W[N][K] floats. A warp step covers R rows x (32*4*C) contiguous floats per row (C float4 per lane per row).
// steps are ordered (row group, k position) and split evenly over all warps. ring depth D steps.
template <int R, int C, int D>
__global__ void __launch_bounds__(256, 1) pattern_kernel(const float* W, int N, int K, float* sink) {
  extern __shared__ float4 ring[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float4* my = ring + warp * (D * R * C * 32);
  const int kw = 128 * C;                        // floats per row per step
  const int ksteps = (K + kw - 1) / kw, groups = N / R;
  const long long T = (long long)groups * ksteps;
  const int gw = blockIdx.x * 8 + warp, nw = gridDim.x * 8;
  long long t0 = T * gw / nw, t1 = T * (gw + 1) / nw;
  long long ti = t0;
  auto issue = [&](int slot) {
    if (ti < t1) {
      int g = (int)(ti / ksteps), ks = (int)(ti - (long long)g * ksteps);
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < C; ++c) {
          int k = ks * kw + c * 128 + lane * 4;
          if (k < K) cp_async16(my + ((slot * R + r) * C + c) * 32 + lane, W + (size_t)(g * R + r) * K + k);
        }
      ++ti;
    }
    commit();
  };
  for (int d = 0; d < D - 1; ++d) issue(d);
  float acc = 0.f;
  int slot = 0;
  for (long long t = t0; t < t1; ++t) {
    issue((slot + D - 1) % D);
    wait<D - 1>();
#pragma unroll
    for (int q = 0; q < R * C; ++q) { float4 v = my[(slot * R * C + q) * 32 + lane]; acc += v.x + v.y + v.z + v.w; }
    slot = (slot + 1) % D;
  }
  wait<0>();
  if (acc == 123.456f) sink[0] = acc;
}

template <int R, int C, int D>
void run(const char* name, const float* W, int N, int K, float* sink) {
  size_t smem = (size_t)8 * D * R * C * 32 * 16;
  cudaFuncSetAttribute(pattern_kernel<R, C, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) pattern_kernel<R, C, D><<<148, 256, smem>>>(W, N, K, sink);
  cudaEventRecord(a);
  const int reps = 20;
  for (int i = 0; i < reps; ++i) pattern_kernel<R, C, D><<<148, 256, smem>>>(W + (size_t)(i % 4) * N * K, N, K, sink);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double us = 1e3 * ms / reps;
  printf("%-44s smem/CTA %6zu KB  %8.2f us  %8.1f GB/s  (%s)\n", name, smem / 1024, us, (double)N * K * 4 / us / 1e3, cudaGetErrorString(cudaGetLastError()));
}

// read-modify-write of two arrays (weights + momentum), per-thread cp.async ring of D rows:
// the memory pattern of bwd_update_kernel without its math.  grid (kgroups, nsplit), 8 warps:
// kw warps along k (128 floats each), 8/kw along rows.
template <int D>
__global__ void __launch_bounds__(256, 1) rmw_kernel(float* W, float* S, int N, int K, int kw, int rows_per_cta) {
  extern __shared__ float4 ring[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wk = warp % kw, wr = warp / kw, rw = 8 / kw;
  const int nb0 = blockIdx.y * rows_per_cta, nb1 = min(N, nb0 + rows_per_cta);
  const int k0 = (blockIdx.x * kw + wk) * 128 + lane * 4;
  const bool kok = k0 < K;
  auto issue = [&](int slot, int row) {
    if (kok && row < nb1) {
      size_t o = (size_t)row * K + k0;
      cp_async16(ring + (0 * D + slot) * 256 + tid, W + o);
      cp_async16(ring + (1 * D + slot) * 256 + tid, S + o);
    }
    commit();
  };
  int n = nb0 + wr;
  for (int d = 0; d < D - 1; ++d) issue(d, n + d * rw);
  int slot = 0;
  for (; n < nb1; n += rw) {
    issue((slot + D - 1) % D, n + (D - 1) * rw);
    wait<D - 1>();
    if (kok) {
      float4 w = ring[(0 * D + slot) * 256 + tid], m = ring[(1 * D + slot) * 256 + tid];
      m.x = 0.9f * m.x + w.x; m.y = 0.9f * m.y + w.y; m.z = 0.9f * m.z + w.z; m.w = 0.9f * m.w + w.w;
      w.x -= 1e-9f * m.x; w.y -= 1e-9f * m.y; w.z -= 1e-9f * m.z; w.w -= 1e-9f * m.w;
      size_t o = (size_t)n * K + k0;
      *reinterpret_cast<float4*>(W + o) = w;
      __stcs(reinterpret_cast<float4*>(S + o), m);
    }
    slot = (slot + 1) % D;
  }
  wait<0>();
}

template <int D>
void run_rmw(const char* name, float* W, float* S, int N, int K, int kw, int nsplit) {
  int kgroups = (K / 128 + kw) / kw;  // ceil(43/kw)
  kgroups = ((K + 127) / 128 + kw - 1) / kw;
  int rows = (N + nsplit - 1) / nsplit;
  size_t smem = (size_t)2 * D * 256 * 16;
  cudaFuncSetAttribute(rmw_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) rmw_kernel<D><<<dim3(kgroups, nsplit), 256, smem>>>(W, S, N, K, kw, rows);
  cudaEventRecord(a);
  const int reps = 20;
  for (int i = 0; i < reps; ++i) rmw_kernel<D><<<dim3(kgroups, nsplit), 256, smem>>>(W + (size_t)(i % 2) * N * K, S + (size_t)(i % 2) * N * K, N, K, kw, rows);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double us = 1e3 * ms / reps;
  printf("%-52s grid %3dx%-3d  %8.2f us  %8.1f GB/s (r+w)  (%s)\n", name, kgroups, nsplit, us, (double)N * K * 16 / us / 1e3, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int N = 5000, K = 5408;
  float *W, *sink;
  cudaMalloc(&W, (size_t)4 * N * K * 4);       // 4 copies (432 MB) so consecutive reps miss L2
  cudaMalloc(&sink, 16);
  cudaMemset(W, 0, (size_t)4 * N * K * 4);
  run<8, 1, 4>("8 rows x 512 B, ring 4 (fwd kernel today)", W, N, K, sink);
  run<8, 1, 6>("8 rows x 512 B, ring 6", W, N, K, sink);
  run<4, 2, 4>("4 rows x 1 KB,  ring 4", W, N, K, sink);
  run<2, 4, 4>("2 rows x 2 KB,  ring 4", W, N, K, sink);
  run<1, 8, 4>("1 row  x 4 KB,  ring 4", W, N, K, sink);
  run<1, 8, 6>("1 row  x 4 KB,  ring 6", W, N, K, sink);
  run<8, 2, 3>("8 rows x 1 KB,  ring 3", W, N, K, sink);
  run<1, 1, 8>("1 row  x 512 B, ring 8 (bwd kernel today)", W, N, K, sink);
  run<1, 1, 16>("1 row  x 512 B, ring 16", W, N, K, sink);
  float* S; cudaMalloc(&S, (size_t)2 * N * K * 4); cudaMemset(S, 0, (size_t)2 * N * K * 4);
  run_rmw<8>("RMW W+momentum, kw=4 nsplit=13, ring 8 (bwd today)", W, S, N, K, 4, 13);
  run_rmw<16>("RMW W+momentum, kw=4 nsplit=13, ring 16", W, S, N, K, 4, 13);
  run_rmw<8>("RMW W+momentum, kw=8 nsplit=24, ring 8", W, S, N, K, 8, 24);
  run_rmw<8>("RMW W+momentum, kw=2 nsplit=6,  ring 8", W, S, N, K, 2, 6);
  run_rmw<8>("RMW W+momentum, kw=1 nsplit=3,  ring 8", W, S, N, K, 1, 3);
  run_rmw<12>("RMW W+momentum, kw=8 nsplit=24, ring 12", W, S, N, K, 8, 24);
  return 0;
}
