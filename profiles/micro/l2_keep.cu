// How much of a re-read buffer survives in B200's L2 across a large streaming pass, as a function of
// the eviction-priority hint on its lines and of the persisting-L2 set-aside?
//   pass K (keep):   read W (size S) with policy P        - timed on the 2nd, 3rd ... iteration
//   pass T (trash):  read+write M (216 MB) with evict_first / normal
// Reported: effective GB/s of pass K (>> HBM rate means L2 hits) and of pass T.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_keep l2_keep.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t policy(int kind) {
  uint64_t p;
  if (kind == 1) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  else if (kind == 2) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ float4 ld_hint(const float4* p, uint64_t pol) {
  float4 v;
  asm volatile("ld.global.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ void st_hint(float4* p, float4 v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol) : "memory");
}

// read n4 float4s; rows below keep4 use policy kind_keep, the rest kind_rest; optional write-back
__global__ void __launch_bounds__(256) stream_kernel(float4* buf, size_t n4, size_t keep4, int kind_keep, int kind_rest,
                                                     int write, float* sink) {
  const uint64_t pk = policy(kind_keep), pr = policy(kind_rest);
  float acc = 0.f;
  const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x * 4 + threadIdx.x; i < n4; i += stride) {
    float4 v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const size_t j = i + (size_t)q * blockDim.x;
      v[q] = j < n4 ? ld_hint(buf + j, j < keep4 ? pk : pr) : make_float4(0, 0, 0, 0);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const size_t j = i + (size_t)q * blockDim.x;
      acc += v[q].x + v[q].w;
      if (write && j < n4) {
        v[q].x += 1.f;
        st_hint(buf + j, v[q], j < keep4 ? pk : pr);
      }
    }
  }
  if (acc == 123.456f) *sink = acc;
}

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("L2 %d MB, persistingL2CacheMaxSize %d MB, accessPolicyMaxWindowSize %d MB\n", prop.l2CacheSize >> 20,
         prop.persistingL2CacheMaxSize >> 20, prop.accessPolicyMaxWindowSize >> 20);
  size_t lim = 0;
  CK(cudaDeviceGetLimit(&lim, cudaLimitPersistingL2CacheSize));
  printf("default cudaLimitPersistingL2CacheSize = %zu MB\n", lim >> 20);
  const size_t MB = 1 << 20;
  const size_t wmax = 128 * MB, msz = 216 * MB;
  float4 *W, *M;
  float* sink;
  CK(cudaMalloc(&W, wmax));
  CK(cudaMalloc(&M, msz));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(W, 0, wmax));
  CK(cudaMemset(M, 0, msz));
  cudaEvent_t e0, e1, e2;
  cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
  const int grid = 148 * 4;
  for (size_t sa_mb : {0, 16, 23, 32, 40, 48, 56, 64, 79}) {
    CK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, sa_mb * MB));
    CK(cudaDeviceGetLimit(&lim, cudaLimitPersistingL2CacheSize));
    printf("--- set-aside %zu MB (asked %zu)\n", lim >> 20, sa_mb);
    const int wwrite = 1, kind = 1;
    for (size_t smb : {16, 24, 32, 48, 64}) {
      const size_t n4 = smb * MB / 16;
      float tk = 0, tt = 0;
      const int reps = 6;
      for (int r = 0; r < reps + 2; ++r) {
        cudaEventRecord(e0);
        stream_kernel<<<grid, 256>>>(W, n4, n4, kind, kind, wwrite, sink);
        cudaEventRecord(e1);
        stream_kernel<<<grid, 256>>>(M, msz / 16, 0, 2, 2, 1, sink);
        cudaEventRecord(e2);
        CK(cudaEventSynchronize(e2));
        float a, b;
        cudaEventElapsedTime(&a, e0, e1);
        cudaEventElapsedTime(&b, e1, e2);
        if (r >= 2) { tk += a; tt += b; }
      }
      const double gk = 2.0 * smb * MB / (tk / reps * 1e-3) / 1e9, gt = 2.0 * msz / (tt / reps * 1e-3) / 1e9;
      printf("  W r+w evict_last S=%3zu MB: keep pass %7.0f GB/s (%.1f us)   trash pass %6.0f GB/s (%.1f us)\n", smb, gk,
             tk / reps * 1e3, gt, tt / reps * 1e3);
    }
  }
  CK(cudaCtxResetPersistingL2Cache());
  return 0;
}
