// Shared helpers for libsplitnn (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <string>

namespace sn {

constexpr int kBT = 16;          // batch tile of the skinny (HBM-bound) kernels
constexpr int kD0 = 5408;        // 32*13*13 cut-layer width
constexpr int kConvParams = 320; // 32*9 weights + 32 biases

extern thread_local std::string g_last_error;
extern std::atomic<uint64_t> g_launches;
extern thread_local cudaError_t g_launch_status;

inline int fail(const char* what, cudaError_t e, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof buf, "%s failed: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  g_last_error = buf;
  return 1;
}
inline int fail_msg(const std::string& msg) {
  g_last_error = msg;
  return 2;
}

#define SN_CUDA(expr)                                                   \
  do {                                                                  \
    cudaError_t e_ = (expr);                                            \
    if (e_ != cudaSuccess) return sn::fail(#expr, e_, __FILE__, __LINE__); \
  } while (0)

// after a launch_k(): the launch call's own status first, then any sticky error; only a launch
// that went through is counted
#define SN_LAUNCH_CHECK()                                               \
  do {                                                                  \
    cudaError_t e_ = sn::g_launch_status;                               \
    sn::g_launch_status = cudaSuccess;                                  \
    if (e_ == cudaSuccess) e_ = cudaGetLastError(); else cudaGetLastError(); \
    if (e_ != cudaSuccess) return sn::fail("kernel launch", e_, __FILE__, __LINE__); \
    sn::g_launches.fetch_add(1, std::memory_order_relaxed);             \
  } while (0)

#define SN_REQUIRE(cond, msg)                                           \
  do {                                                                  \
    if (!(cond)) return sn::fail_msg(std::string("libsplitnn: ") + (msg)); \
  } while (0)

// Optimizer hyper-parameters with the per-step scalars torch computes in host double
// (torch optim/adam.py _single_tensor_adam; optim/sgd.py _single_tensor_sgd).
struct OptParams {
  int type;          // SN_OPT_*
  float lr;
  float momentum;
  float one_minus_b1;
  float beta2;
  float one_minus_b2;
  float eps;
  float wd;
  float step_size;     // lr / (1 - beta1^t)
  float bc2_sqrt;      // sqrt(1 - beta2^t)
  float inv_bc2_sqrt;  // 1 / sqrt(1 - beta2^t), formed in double
  int acc_first;       // gradient-accumulation launches (OPT == 3 in bwd_update_kernel): first batch tile
};

template <int OPT>
__device__ __forceinline__ void opt_update(const OptParams& op, float& p, float g, float& s1, float& s2) {
  if (OPT == 1) {  // SGD momentum: buf = mom*buf + g ; p -= lr*buf   (buf starts at 0 == "buf <- g")
    s1 = __fadd_rn(__fmul_rn(op.momentum, s1), g);
    p = __fmaf_rn(-op.lr, s1, p);
  } else if (OPT == 2) {  // Adam, coupled L2
    if (op.wd != 0.f) g = __fmaf_rn(op.wd, p, g);
    s1 = __fmaf_rn(op.one_minus_b1, g - s1, s1);                      // lerp_
    s2 = __fmaf_rn(op.one_minus_b2 * g, g, __fmul_rn(s2, op.beta2));   // mul_ then addcmul_
    // denom = sqrt(v)/sqrt(bc2) + eps ; p -= step_size * m/denom.  The two divisions and the square
    // root are the MUFU forms (sqrt.approx, multiply by the host-double reciprocal, __fdividef):
    // <= 2 ulp each, ~1e-7 relative on the update, far inside the 1e-4 parity bound, and they
    // keep the Adam traversal HBM-bound instead of issue-bound (IEEE div/sqrt are ~10x the
    // instructions).
    float sq;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(sq) : "f"(s2));
    const float denom = __fmaf_rn(sq, op.inv_bc2_sqrt, op.eps);
    p = __fmaf_rn(-op.step_size, __fdividef(s1, denom), p);           // addcdiv_
  }
}

// Programmatic dependent launch (PDL).  Every kernel of a step is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization: it may start while its predecessor in the
// stream is still running, must call pdl_wait() before touching anything the predecessor writes,
// and calls pdl_launch() right AFTER its own wait (so a dependent can never overtake the
// predecessor's predecessor: whatever a kernel does before its wait may only touch data produced
// two or more kernels back).  Without the launch attribute both are no-ops.
#ifdef SN_TRACE
// Debug build only (profiles/micro/step_trace.py): CTA 0 of every kernel stamps %globaltimer right
// after its dependency wait, i.e. at the moment its predecessor has fully completed; consecutive
// differences are each kernel's cost inside the PDL chain, which a profiler cannot see.
__device__ unsigned long long g_trace[8192];
__device__ unsigned g_trace_n;
__device__ long long g_trace_phase[16];        // clock64 phase stamps of one CTA of the kernel under study
#define SN_TP(i) do { if (threadIdx.x == 0) tp_[i] = clock64(); } while (0)
#define SN_TP_DECL long long tp_[16] = {0}
#ifdef SN_TRACE_TC   // the tcgen05 skinny kernel owns g_trace_phase in this build
#define SN_TP_FLUSH() do {} while (0)
#else
#define SN_TP_FLUSH() do { if (threadIdx.x == 0) for (int q_ = 0; q_ < 16; ++q_) g_trace_phase[q_] = tp_[q_]; } while (0)
#endif
#else
#define SN_TP(i) do {} while (0)
#define SN_TP_DECL do {} while (0)
#define SN_TP_FLUSH() do {} while (0)
#endif
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
#ifdef SN_TRACE
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_trace[atomicAdd(&g_trace_n, 1u) & 8191u] = t;
  }
#endif
}
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Cross-GPU flags (peer-memory mailboxes, include/splitnn.h): gate kernels of one CTA.  The waiter
// polls a flag in ITS OWN GPU's memory that a kernel on ANOTHER GPU sets over NVLink, so the two
// never compete for the same SMs; being PDL launches, the kernel queued behind a waiter runs its
// prologue (weight prefetch) while the waiter polls.  A stream-ordered cuStreamWaitValue32 pair in
// the same place costs ~20 us per hop (front-end semaphore latency), these ~3.
// thread i < count polls flag[i * stride_words]: one launch can wait for `count` producers
// The wait is bounded: a producer that died would otherwise wedge this GPU for good.  After
// `timeout_ns` of %globaltimer the waiter gives up, writes {1, value, last value seen, index} to
// `err` (pinned host memory mapped into every device, read by the next library call on the host:
// sn_flag_status) and lets the stream run on - whatever follows computes on stale data, and the
// host is told.
__global__ void flag_wait_kernel(const unsigned* flag, unsigned value, int count, int stride_words,
                                 unsigned long long timeout_ns, unsigned* err) {
  pdl_wait();
  pdl_launch();
  if ((int)threadIdx.x < count) {
    const unsigned* f = flag + (size_t)threadIdx.x * stride_words;
    unsigned cur;
    unsigned long long t0 = 0, now = 0;
    unsigned spins = 0;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(cur) : "l"(f) : "memory");
      if ((int)(cur - value) >= 0) break;
      if ((++spins & 1023u) == 0 && timeout_ns) {       // the timer is read once per ~1k polls
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > timeout_ns) {
          if (err) {
            err[1] = value; err[2] = cur; err[3] = threadIdx.x;
            __threadfence_system();
            *reinterpret_cast<volatile unsigned*>(err) = 1u;
            __threadfence_system();
          }
          break;
        }
      }
    }
  }
}
// optional small copy (labels) into the peer's mailbox, then the flag: everything the stream did
// before this launch, and the copy, are visible to the peer before it sees the new value
__global__ void flag_signal_kernel(unsigned* flag, unsigned value, unsigned long long* dst,
                                   const unsigned long long* src, int words, unsigned* const* flag_table, int nflags) {
  pdl_wait();
  pdl_launch();
  for (int i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0 && flag) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
  // several consumers (one flag in each one's own memory): thread i releases flag_table[i]
  if ((int)threadIdx.x < nflags)
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag_table[threadIdx.x]), "r"(value) : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Philox4x32-10 (Salmon et al. 2011): counter-based RNG for perf-mode dropout.
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
// the four 32-bit words that decide elements 4*idx4 .. 4*idx4+3 of dropout layer `layer` at step `offset`
__device__ __forceinline__ uint4 philox_words(uint64_t seed, uint64_t offset, int layer, uint64_t idx4) {
  uint4 c = make_uint4((uint32_t)idx4, (uint32_t)(idx4 >> 32), (uint32_t)offset, ((uint32_t)(offset >> 32) << 4) | (uint32_t)layer);
  return philox4x32(c, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
}
__device__ __forceinline__ bool philox_word_keeps(uint32_t v, float keep_prob) { return (v >> 8) * (1.0f / 16777216.0f) < keep_prob; }
// keep-decision for element `idx` of dropout layer `layer` at step `offset`
__device__ __forceinline__ bool philox_keep(uint64_t seed, uint64_t offset, int layer, uint64_t idx, float keep_prob) {
  uint4 r = philox_words(seed, offset, layer, idx >> 2);
  uint32_t v = (idx & 3) == 0 ? r.x : (idx & 3) == 1 ? r.y : (idx & 3) == 2 ? r.z : r.w;
  return philox_word_keeps(v, keep_prob);
}

}  // namespace sn
