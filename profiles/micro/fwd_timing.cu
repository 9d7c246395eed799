// Phase timing of fwd_skinny_kernel (clock64 per warp): prologue / X staging / barrier wait /
// main loop (of which row-group reductions).  Built with -DSN_TIMING against the real kernel source.
#define SN_TIMING 1
#include "../../splitlearning_b200/csrc/mlp_kernels.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
namespace sn { thread_local std::string g_last_error; std::atomic<uint64_t> g_launches{0}; }
using namespace sn;
int main(int argc, char** argv) {
  int N = 5000, K = 5408, B = 16, kc = 1408, ksplit = 4, nx = 37;
  if (argc > 5) { N = atoi(argv[1]); K = atoi(argv[2]); kc = atoi(argv[3]); ksplit = atoi(argv[4]); nx = atoi(argv[5]); }
  float *W, *X, *P;
  cudaMalloc(&W, (size_t)N * K * 4); cudaMalloc(&X, (size_t)B * K * 4); cudaMalloc(&P, (size_t)2 * ksplit * 16 * N * 4);
  cudaMemset(W, 0, (size_t)N * K * 4); cudaMemset(X, 0, (size_t)B * K * 4);
  float* flush; cudaMalloc(&flush, 256 << 20);
  size_t smem = (size_t)16 * kc * 4 + (size_t)8 * kFwdRing * 8 * 32 * 16;
  cudaFuncSetAttribute(fwd_skinny_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9;
  for (int r = 0; r < 5; ++r) {
    cudaMemset(flush, r, 256 << 20);
    cudaEventRecord(a);
    fwd_skinny_kernel<8><<<dim3(nx, ksplit, 1), 256, smem>>>(X, K, B, W, K, N, K, kc, P, 16, ksplit, 0);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  printf("kernel %.2f us (%s)\n", best * 1e3, cudaGetErrorString(cudaGetLastError()));
  std::vector<long long> h(148 * 8 * 8);
  cudaMemcpyFromSymbol(h.data(), g_fwd_timing, h.size() * 8);
  double s[6] = {0}; long long tmin = 1LL << 62, tmax = 0, emax = 0, emin = 1LL << 62;
  int n = nx * ksplit * 8;
  for (int i = 0; i < n; ++i) { for (int j = 0; j < 6; ++j) s[j] += h[i * 8 + j]; tmin = std::min(tmin, h[i*8+6]); tmax = std::max(tmax, h[i*8+6]); emax = std::max(emax, h[i*8+7]); emin = std::min(emin, h[i*8+7]); }
  printf("avg cycles per warp: prologue %.0f | staging %.0f | barrier wait %.0f | main loop %.0f (reductions %.0f) | tiles %.1f\n",
         s[0]/n, s[1]/n, s[2]/n, s[3]/n, s[4]/n, s[5]/n);
  printf("(clock64 is per-SM; start spread and end spread are not comparable across SMs)\n");
  return 0;
}
