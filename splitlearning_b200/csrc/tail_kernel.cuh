// The small tail of a batch-16 server step as ONE persistent kernel (north_star: "a persistent
// whole-MLP kernel").  For the 5408-5000-1000-100 server the step is two big weight streams (fc1
// forward, fc1 backward+update) and, between them, seven small launches that move < 1 MB each or
// stream the 20 MB of fc2: epilogue(fc1) -> fc2 forward -> epilogue(fc2) -> fc3 forward + CE ->
// fc3 backward -> fc2 backward+update -> dA reduction.  As separate launches they cost 51 us of the
// 165 us vanilla step (62 of SISA's 185), about three times their traffic floor: each boundary is a
// kernel drain + launch + prologue (~3 us under programmatic dependent launch; 2.0-2.4 us for EMPTY
// kernels, profiles/r2_grid_barrier.txt).  Here the same phases run inside one launch of one CTA per
// SM and meet at grid-wide barriers (1.2-1.6 us measured with data crossing SMs).  The activations
// (H1 320 KB, H2 64 KB) do not fit one SM, so they live in L2 between phases, not in shared memory.
//
// The phases are the bodies of the stand-alone kernels (mlp_kernels.cuh), compiled with GridSync in
// place of PdlSync: identical arithmetic in identical order, so a step through the tail kernel is
// bit-identical to the launch chain (tests/test_gpu_kernels.py runs both).  Anything a phase reads
// that an earlier phase wrote is loaded with ld.global.cg (ld_dep), never through L1 or the
// read-only path.  Weights of the NEXT phase's stream are prefetched before the barrier in front of
// it, as the stand-alone kernels do before their dependency wait.
//
// Co-residency: the grid is one CTA per SM and every CTA of the grid must be running for a barrier to
// complete.  The launch is an ordinary (programmatically serialized) one: the CTAs wait for the
// previous kernel with griddepcontrol.wait before touching the first barrier, by which time that
// kernel has left the SMs; the dependents are released only after the last barrier.
#pragma once
#include "mlp_kernels.cuh"

namespace sn {

__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    // poll with back-off: in phases that occupy a third of the grid the other hundred CTAs wait here
    // for 10-20 us, and a hundred pollers hammering one L2 line slowed every phase down (measured:
    // fc3 backward 5.9 -> 19 us inside the tail kernel with a tight polling loop)
    unsigned v, ns = 32;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
      if ((int)(v - target) >= 0) break;
      __nanosleep(ns);
      if (ns < 512) ns *= 2;
    }
    __threadfence();
#ifdef SN_TRACE
    if (blockIdx.x == 0) {                       // trace build: one stamp per phase boundary (see common.cuh)
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      g_trace[atomicAdd(&g_trace_n, 1u) & 8191u] = t;
    }
#endif
  }
  __syncthreads();
}

// Phase boundary inside the tail kernel.  The arrival count only ever grows (wrap-safe signed
// compare): launch i of a handle uses the targets base + G, base + 2G, ...
struct GridSync {
  static constexpr bool kCoherent = true;
  unsigned* counter;
  unsigned* target;      // running target of this CTA (thread-uniform, lives in registers of the caller)
  unsigned G;
  __device__ __forceinline__ void operator()() const {
    *target += G;
    grid_barrier(counter, *target);
  }
};
// The first phase only waits for the previous KERNEL.
struct FirstSync {
  static constexpr bool kCoherent = true;
  __device__ __forceinline__ void operator()() const { pdl_wait(); }
};

// The last phase: after its barrier the next kernel may start its prologue.
struct LastSync {
  static constexpr bool kCoherent = true;
  GridSync g;
  __device__ __forceinline__ void operator()() const {
    g();
    pdl_launch();
  }
};

constexpr int kTailBarriers = 6;

struct TailParams {
  int B;
  float keep, scale;               // dropout keep probability (1 = none / mask bytes given) and 1/keep
  uint64_t seed, offset;
  OptParams op;
  unsigned* bar;                   // arrival counter (device), never reset
  unsigned bar_base;               // arrivals before this launch
  // phase 0: H1 = dropout(relu(sum of the first layer's partials + b0))
  const float* P0; int parts0, bpad0; const float* b0; const uint8_t* mask0; float* H1; int n1;
  // phase 1: second layer's split-K partials
  const float* W1; int n2, kc1, ksplit1, nx1, l2_keep1; float* P1;
  // phase 2: H2
  const float* b1; const uint8_t* mask1; float* H2;
  // phase 3: logits = H2 . W2^T + b2
  const float* W2; const float* b2; int n3; float* H3; float* logits_out;
  // phase 4: cross-entropy + last layer's backward/update (dZ2 gated in place)
  CeTail ce; float* W2m; float* b2m; float *s1_2, *s2_2, *bs1_2, *bs2_2; float gscale; float* dZ2;
  // phase 5: second layer's backward/update, input-gradient partials
  float* W1m; float* b1m; float *s1_1, *s2_1, *bs1_1, *bs2_1; int kw1, rows1, kgroups1, nsplit1; float* PA;
  // phase 6: dZ1 = (sum of partials) gated by H1
  float* dZ1;
};

template <int OPT>
__global__ void __launch_bounds__(256, 1) tail_kernel(const __grid_constant__ TailParams p) {
  extern __shared__ float4 smem4[];
  const unsigned cta = blockIdx.x, G = gridDim.x;
  unsigned target = p.bar_base;
  const GridSync gs{p.bar, &target, G};
  const float kp = p.keep;

  // ---- phase 0: first layer's epilogue (waits for the kernel that produced the partials)
  fwd_epilogue_body(cta, G, FirstSync{}, p.P0, p.parts0, p.bpad0, p.B, p.n1, p.b0, 1, p.mask0, p.mask0 ? 1.f : kp, p.scale, p.seed,
                    p.offset, 0, p.H1, p.n1, (float*)nullptr, 0, 0);

  // ---- phase 1: second layer forward (its weight ring starts before the barrier)
  __syncthreads();
  if ((int)cta < p.nx1 * p.ksplit1)
    fwd_skinny_body<8>(smem4, dim3(cta % p.nx1, cta / p.nx1, 0), p.nx1, gs, p.H1, p.n1, p.B, p.W1, p.n1, p.n2, p.n1, p.kc1, p.P1, kBT,
                       p.ksplit1, 1, p.l2_keep1);
  else
    gs();

  // ---- phase 2: second layer's epilogue
  fwd_epilogue_body(cta, G, gs, p.P1, kFwdSlots * p.ksplit1, kBT, p.B, p.n2, p.b1, 1, p.mask1, p.mask1 ? 1.f : kp, p.scale, p.seed,
                    p.offset, 1, p.H2, p.n2, (float*)nullptr, 0, 0);

  // ---- phase 3: last layer forward (two output rows per CTA)
  if ((int)cta < (p.n3 + kHeadRows - 1) / kHeadRows)
    fwd_small_body(dim3(cta, 0, 0), 0u, gs, p.H2, p.n2, p.B, p.W2, p.n2, p.n3, p.n2, p.b2, 0, (const uint8_t*)nullptr, 1.f, 1.f,
                   p.seed, p.offset, 2, p.H3, p.n3, p.logits_out, 1, CeTail{});
  else
    gs();

  // ---- phase 4: loss + last layer backward (every participating CTA forms the loss gradient itself)
  if ((int)cta < (p.n2 + 31) / 32)
    bwd_small_body<OPT, true>(cta, gs, p.W2m, p.n2, p.n3, p.n2, p.b2m, p.s1_2, p.s2_2, p.bs1_2, p.bs2_2, (const float*)nullptr, p.n3,
                              p.H2, p.n2, p.B, p.op, 1, p.gscale, p.dZ2, p.n2, p.H3, p.ce);
  else
    gs();

  // ---- phase 5: second layer backward + update (weight / state ring starts before the barrier)
  __syncthreads();
  if ((int)cta < p.kgroups1 * p.nsplit1)
    bwd_update_body<OPT, true, false, false>(smem4, dim3(cta % p.kgroups1, cta / p.kgroups1, 0), gs, p.W1m, p.n1, p.n2, p.n1, p.b1m,
                                             p.s1_1, p.s2_1, p.bs1_1, p.bs2_1, p.dZ2, p.n2, p.H1, p.n1, p.B, p.PA, p.kw1, p.rows1,
                                             p.op, (const float*)nullptr, 0, (float*)nullptr, p.l2_keep1, (const float*)nullptr,
                                             1.f);
  else
    gs();

  // ---- phase 6: dZ1; the next kernel may start its prologue from here on
  bwd_epilogue_body(cta, G, LastSync{gs}, p.PA, p.nsplit1, p.B, p.n1, p.H1, p.n1, p.gscale, p.dZ1, p.n1, (float* const*)nullptr, 1);
}

}  // namespace sn
