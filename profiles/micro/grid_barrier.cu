// What does a phase boundary cost inside ONE persistent kernel (grid barrier) against a kernel
// boundary under programmatic dependent launch?  The question behind north_star's "persistent
// whole-MLP kernel" (DESIGN.md 3): at batch 16 the activations of the server MLP (400 KB) do not fit
// one SM, so the layers of a persistent kernel must meet at grid-wide barriers.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/micro/grid_barrier profiles/micro/grid_barrier.cu
//   ./profiles/micro/grid_barrier
//
// Measures, per boundary, with one 256-thread CTA per SM (148):
//   (a) grid barrier, counter form: __syncthreads; thread 0: __threadfence, atomicAdd on one global
//       counter, spin on it with ld.acquire until all CTAs of this generation have arrived; __syncthreads
//   (b) the same with a little work between barriers: every CTA writes a 4 KB slice that ANOTHER CTA
//       reads after the barrier (what a layer boundary does: data really crosses SMs)
//   (c) a chain of empty kernels launched with programmatic stream serialization, each doing
//       griddepcontrol.wait / launch_dependents like the library's kernels
//   (d) the same chain where each kernel also does (b)'s work
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    unsigned v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    } while (v < target);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256, 1) barrier_only_kernel(unsigned* counter, int phases) {
  for (int p = 1; p <= phases; ++p) grid_barrier(counter, (unsigned)p * gridDim.x);
}

__global__ void __launch_bounds__(256, 1) barrier_work_kernel(unsigned* counter, int phases, float* buf, float* sink) {
  float acc = 0.f;
  const int n = gridDim.x;
  for (int p = 1; p <= phases; ++p) {
    // write my slice (1024 floats), barrier, read my neighbour's slice
    float4* mine = reinterpret_cast<float4*>(buf + (size_t)blockIdx.x * 1024);
    mine[threadIdx.x] = make_float4(p + acc, 1.f, 2.f, 3.f);
    grid_barrier(counter, (unsigned)p * n);
    const float4* other = reinterpret_cast<const float4*>(buf + (size_t)((blockIdx.x + 37) % n) * 1024);
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(other + threadIdx.x));
    acc += v.x * 1e-9f;
    // (a second barrier would be needed before the slices are overwritten; slices alternate instead)
    buf += (size_t)n * 1024 * ((p & 1) ? 1 : -1);
  }
  if (acc == 12345.f) sink[0] = acc;
}

__global__ void __launch_bounds__(256, 1) pdl_empty_kernel(float* sink) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (sink == nullptr && threadIdx.x == 1024) printf("never\n");
}

__global__ void __launch_bounds__(256, 1) pdl_work_kernel(float* buf_w, const float* buf_r, float* sink, int p) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int n = gridDim.x;
  const float4* other = reinterpret_cast<const float4*>(buf_r + (size_t)((blockIdx.x + 37) % n) * 1024);
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(other + threadIdx.x));
  float4* mine = reinterpret_cast<float4*>(buf_w + (size_t)blockIdx.x * 1024);
  mine[threadIdx.x] = make_float4(p + v.x * 1e-9f, 1.f, 2.f, 3.f);
  if (v.x == 12345.f) sink[0] = v.x;
}

template <typename... P, typename... A>
static void launch_pdl(void (*k)(P...), int grid, cudaStream_t st, A... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(256);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, k, static_cast<P>(args)...));
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  unsigned* counter;
  float *buf, *sink;
  CK(cudaMalloc(&counter, 4));
  CK(cudaMalloc(&buf, (size_t)2 * sms * 1024 * 4));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(buf, 0, (size_t)2 * sms * 1024 * 4));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const int phases = 2000, reps = 5;
  float best[4] = {1e30f, 1e30f, 1e30f, 1e30f};
  for (int r = 0; r < reps; ++r) {
    float ms;
    CK(cudaMemsetAsync(counter, 0, 4, st));
    CK(cudaEventRecord(e0, st));
    barrier_only_kernel<<<sms, 256, 0, st>>>(counter, phases);
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    best[0] = fminf(best[0], ms);
    CK(cudaMemsetAsync(counter, 0, 4, st));
    CK(cudaEventRecord(e0, st));
    barrier_work_kernel<<<sms, 256, 0, st>>>(counter, phases, buf, sink);
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    best[1] = fminf(best[1], ms);
    CK(cudaEventRecord(e0, st));
    for (int p = 0; p < phases; ++p) launch_pdl(pdl_empty_kernel, sms, st, sink);
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    best[2] = fminf(best[2], ms);
    CK(cudaEventRecord(e0, st));
    for (int p = 0; p < phases; ++p)
      launch_pdl(pdl_work_kernel, sms, st, buf + (size_t)(p & 1) * sms * 1024, (const float*)(buf + (size_t)((p + 1) & 1) * sms * 1024), sink, p);
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    best[3] = fminf(best[3], ms);
  }
  CK(cudaGetLastError());
  printf("SMs %d, %d boundaries, best of %d\n", sms, phases, reps);
  printf("grid barrier, nothing else          : %.3f us per boundary\n", 1e3f * best[0] / phases);
  printf("grid barrier + 4 KB/CTA cross-SM    : %.3f us per boundary\n", 1e3f * best[1] / phases);
  printf("PDL kernel boundary, empty kernels  : %.3f us per boundary\n", 1e3f * best[2] / phases);
  printf("PDL kernel boundary + the same work : %.3f us per boundary\n", 1e3f * best[3] / phases);
  return 0;
}
