// libsplitnn: C ABI (include/splitnn.h) over the sm_100a kernels.  Host logic only: handle
// bookkeeping, launch plans, per-step optimizer scalars.  No CPU compute path exists here.
#include "../../include/splitnn.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <functional>
#include <mutex>
#include <vector>

#include "alice_kernels.cuh"
#include "common.cuh"
#include "gemm_tc.cuh"
#include "mlp_kernels.cuh"
#include "tail_kernel.cuh"

namespace sn {
thread_local std::string g_last_error;
std::atomic<uint64_t> g_launches{0};

// large-batch matrix products: 0 = 3xTF32 on tcgen05 (default, fp32-accurate, parity mode);
// 1 = TF32 on tcgen05 (perf mode); 2 = fp32 SIMT fallback everywhere
std::atomic<int> g_matmul_precision{0};

static int sm_count(int device) {
  static int cache[64] = {0};
  if (device >= 0 && device < 64 && cache[device]) return cache[device];
  int n = 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) { n = 148; cudaGetLastError(); }
  if (device >= 0 && device < 64) cache[device] = n;
  return n;
}
static inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int ew_grid(size_t n, int sms) {
  size_t g = (n + 255) / 256;
  return (int)std::max<size_t>(1, std::min<size_t>(g, (size_t)sms * 8));
}

// ---------------------------------------------------------------- launch plans
constexpr int kMaxKc = 1536;        // X chunk [16][kc] fp32 = 96 KB of shared memory (+128 KB of W rings)
constexpr int kSkinnyFwdMaxB = 64;  // above this the forward uses the general-batch path
constexpr size_t kSmallLayerElems = 256 * 1024;   // <= 1 MB of weights: one-warp-per-row kernel
constexpr int kSmallLayerMaxK = 2048;             // its X tile [16][K] fp32 must fit in shared memory

struct FwdPlan { int kc, ksplit, nx; };
struct BwdPlan { int kw, kgroups, nsplit, rows_per_cta; };

// rows per tile of the FMA weight-stream forward: 8 (8 warps per CTA) or 4 (16 warps per CTA)
static int fwd_rows() {
  static const int r = (getenv("SN_FWD_ROWS") && atoi(getenv("SN_FWD_ROWS")) == 4) ? 4 : 8;
  return r;
}

// largest batch whose backward goes through the fp32 traversal in tiles of 16 rows with a gradient
// buffer; above it: tcgen05 GEMMs (3xTF32, each exact to ~1e-6).  Default 16, i.e. the tiled form is
// off: since the dW GEMM's optimizer epilogue streams through shared memory the GEMMs win from 17 rows
// up (H 8192: 0.46 vs 0.67 ms at 32 rows, 0.60 vs 1.12 ms at 64).  sn_set_bwd_tiled_max /
// SN_BWD_TILED_MAX (<= 64) bring it back for A/B runs and its parity test.
std::atomic<int> g_bwd_tiled_max{-1};
static int bwd_tiled_max() {
  int v = g_bwd_tiled_max.load();
  if (v < 0) {
    v = getenv("SN_BWD_TILED_MAX") ? atoi(getenv("SN_BWD_TILED_MAX")) : kBT;
    v = std::max(kBT, std::min(v, kSkinnyFwdMaxB));
    g_bwd_tiled_max.store(v);
  }
  return v;
}

static FwdPlan plan_fwd(int K, int N, int sms) {
  // k-chunks of whole 128-float slices, as equal as possible; CTAs per chunk capped so that a
  // warp's tile range is at least one row group long (=> at most two partials per group)
  const int R = fwd_rows(), WPC = 64 / R;
  const int slices = ceil_div(K, 128), groups = ceil_div(N, R);
  FwdPlan best{128 * std::min(slices, kMaxKc / 128), 0, 1};
  best.ksplit = ceil_div(K, best.kc);
  double best_score = -1;
  for (int s = 1; s <= slices; ++s) {
    int kc = 128 * ceil_div(slices, s);
    if (kc > kMaxKc) continue;
    int actual = ceil_div(K, kc);
    int nx = std::max(1, std::min(sms / actual, groups / WPC));
    double iters = kc / 128.0;
    double tp = groups * iters / (nx * (double)WPC);     // tiles per warp
    double util = std::min(1.0, (double)nx * actual / sms);
    double balance = (double)K / ((double)actual * kc);
    double score = util * balance * (tp / (tp + 2.0)) * (iters / (iters + 0.25));
    if (score > best_score) { best_score = score; best = FwdPlan{kc, actual, nx}; }
  }
  return best;
}

static BwdPlan plan_bwd(int N, int K, int sms, bool need_da, bool adam) {
  BwdPlan best{1, ceil_div(K, 128), 1, N};
  double best_score = -1;
  const int kslices = ceil_div(K, 128);
  const double main_bytes = (double)N * K * 4.0 * (adam ? 6 : 4);
  for (int kw = 1; kw <= 8; kw *= 2) {
    int rw = 8 / kw;
    int kgroups = ceil_div(kslices, kw);
    int nsplit = std::max(1, sms / kgroups);
    int rows = std::max(rw, ceil_div(N, nsplit));
    rows = std::min(rows, 1024);
    nsplit = ceil_div(N, rows);
    double ctas = (double)kgroups * nsplit;
    double waves = ceil(ctas / sms);
    double util = ctas / (waves * sms);
    double kwaste = (double)K / ((double)kgroups * kw * 128);
    double part = need_da ? (double)nsplit * 16.0 * K * 4.0 * 2.0 : 0.0;
    double score = util * kwaste * main_bytes / (main_bytes + part);
    // near-ties go to the wider CTA: kw*512 contiguous bytes per row instead of 512 (DRAM page
    // locality; the 37856-wide concat layer ran at 4.35 TB/s with kw=1)
    score *= 1.0 + 0.01 * (kw == 2 ? 1 : kw == 4 ? 2 : kw == 8 ? 3 : 0);
    if (score > best_score) { best_score = score; best = BwdPlan{kw, kgroups, nsplit, rows}; }
  }
  return best;
}

constexpr size_t kMaxSmemPerCta = 232448;    // 227 KB
// shared-memory collection of the fused next-forward partials (bwd_update_kernel, pf_smem): fits?
static size_t bwd_pf_smem_bytes(const BwdPlan& p) { return (size_t)p.kw * p.rows_per_cta * (kBT + 1) * 4; }

static size_t bwd_smem_bytes(const BwdPlan& p, bool need_da, bool adam) {
  // dZ tile + max(per-thread cp.async ring, dA cross-warp reduction buffer) which alias
  const size_t ring = (size_t)(adam ? 3 : 2) * kRing * 256 * 16;
  const size_t red = need_da ? (size_t)8 * kBT * 32 * 16 : 0;
  return (size_t)((p.rows_per_cta * kBT + 3) / 4) * 16 + std::max(ring, red);
}

// ---------------------------------------------------------------- optimizer scalars
struct Hyper {
  int type = SN_OPT_NONE;
  double lr = 0, momentum = 0, beta1 = 0.9, beta2 = 0.999, eps = 1e-8, wd = 0;
};
static OptParams make_opt(const Hyper& h, int64_t step) {
  // every scalar is formed in double and rounded to fp32 where torch rounds it
  OptParams o{};
  o.type = h.type;
  o.lr = (float)h.lr;
  o.momentum = (float)h.momentum;
  o.one_minus_b1 = (float)(1.0 - h.beta1);
  o.beta2 = (float)h.beta2;
  o.one_minus_b2 = (float)(1.0 - h.beta2);
  o.eps = (float)h.eps;
  o.wd = (float)h.wd;
  if (h.type == SN_OPT_ADAM) {
    const double bc1 = 1.0 - pow(h.beta1, (double)step), bc2 = 1.0 - pow(h.beta2, (double)step);
    o.step_size = (float)(h.lr / bc1);
    o.bc2_sqrt = (float)sqrt(bc2);
    o.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
  }
  return o;
}

template <typename F>
static int set_smem(F func, size_t bytes) {
  if (bytes > 48 * 1024) SN_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

// ---------------------------------------------------------------- per-kernel timing (bench only)
// When enabled on a handle, the two streaming kernels are bracketed by CUDA events on the
// launching stream; bench.py reads the per-launch averages for the roofline report.
struct Profiler {
  bool on = false;
  struct Rec { int kind, layer; cudaEvent_t a, b; };
  std::vector<Rec> recs;
  cudaEvent_t begin(cudaStream_t st, int kind, int layer) {
    if (!on) return nullptr;
    Rec r{kind, layer, nullptr, nullptr};
    cudaEventCreate(&r.a);
    cudaEventCreate(&r.b);
    cudaEventRecord(r.a, st);
    recs.push_back(r);
    return r.b;
  }
  static void end(cudaEvent_t e, cudaStream_t st) { if (e) cudaEventRecord(e, st); }
  void clear() {
    for (auto& r : recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    recs.clear();
  }
};

// ---------------------------------------------------------------- launches (PDL)
// All kernels go through launch_k: cudaLaunchKernelEx with programmatic stream serialization, so a
// kernel's prologue overlaps its predecessor's tail (see pdl_wait/pdl_launch in common.cuh).
// SN_NO_PDL=1 in the environment turns the attribute off for A/B timing.
// Status of the most recent cudaLaunchKernelEx on this thread; SN_LAUNCH_CHECK reports it.
thread_local cudaError_t g_launch_status = cudaSuccess;

// Which weights did the most recent library launch on each device write?  A streaming kernel may
// start reading its weights BEFORE its dependency wait only when the launch right before it on the
// device cannot be writing them.  The record is per device, not per handle: several handles may be
// bound to the same tensors (a training engine and an inference engine over one module).  Every
// launch clears it; a launch that updates weights sets it again afterwards (mark_written).  One host
// thread drives a device and the library's launches on it go to one stream at a time; callers that
// interleave streams on one device must order them themselves (include/splitnn.h).
struct WrittenRange { const char* lo; const char* hi; };
static WrittenRange g_written[64] = {};
static inline int cur_device() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess) { cudaGetLastError(); d = 0; }
  return d;
}
static inline void mark_written(const void* p, size_t bytes) {
  const int d = cur_device();
  if (d >= 0 && d < 64) g_written[d] = WrittenRange{static_cast<const char*>(p), static_cast<const char*>(p) + bytes};
}
static inline bool just_written(const void* p, size_t bytes) {
  const int d = cur_device();
  if (d < 0 || d >= 64) return true;
  const char* lo = static_cast<const char*>(p);
  return g_written[d].lo && lo < g_written[d].hi && g_written[d].lo < lo + bytes;
}

// cluster_x > 1: thread-block clusters of (cluster_x, 1, 1) (the tcgen05 CTA pairs)
template <typename... P, typename... A>
static inline void launch_kc(void (*kernel)(P...), int cluster_x, dim3 grid, dim3 block, size_t smem, cudaStream_t st, A... args) {
  static const bool pdl = getenv("SN_NO_PDL") == nullptr;
  {
    const int d = cur_device();
    if (d >= 0 && d < 64) g_written[d] = WrittenRange{nullptr, nullptr};
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (pdl) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  g_launch_status = cudaLaunchKernelEx(&cfg, kernel, static_cast<P>(args)...);
}
template <typename... P, typename... A>
static inline void launch_k(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A... args) {
  launch_kc(kernel, 1, grid, block, smem, st, args...);
}

// ---------------------------------------------------------------- shared launch helpers
// C[M,N] (+epilogue) = sum_r A(m,r) B(n,r).  Operand memory: K-major = [extent, R] row-major,
// MN-major = [R, extent] row-major; lda/ldb are the row strides of those matrices.
template <bool AMN, bool BMN, int EPI, int OPT, bool X3, int BN, int STG = tc::stages_for(BN, X3)>
static int launch_tc_gemm_p(const float* A, int lda, int M, const float* Bm, int ldb, int N, int R, float* C, int ldc,
                            const tc::Epilogue& ep, cudaStream_t st, int splits = 1, size_t split_stride = 0) {
  CUtensorMap ma, mb;
  if (AMN ? tc::make_map(&ma, A, R, M, lda, 32, true) : tc::make_map(&ma, A, M, R, lda, tc::BM)) return 1;
  if (BMN ? tc::make_map(&mb, Bm, R, N, ldb, 32, true) : tc::make_map(&mb, Bm, N, R, ldb, BN)) return 1;
  auto kern = tc::gemm_tf32_kernel<AMN, BMN, EPI, OPT, X3, BN, STG>;
  constexpr size_t smem = tc::smem_bytes<X3, BN, STG>();
  if (set_smem(kern, smem)) return 1;
  const int kblocks = ceil_div(R, tc::BK);
  const int per = ceil_div(kblocks, std::max(1, splits));
  dim3 grid(ceil_div(N, BN), ceil_div(M, tc::BM), ceil_div(kblocks, per));
  launch_k(kern, grid, dim3(tc::kThreads), smem, st, ma, mb, M, N, R, C, ldc, ep, per, split_stride);
  SN_LAUNCH_CHECK();
  return 0;
}
// CTA pairs (cta_group::2): 256 x 256 output tile per pair, each CTA staging half of B.  3xTF32: A through
// tensor memory, 4-deep ring; one-pass TF32: both operands from shared memory, 6-deep ring.
template <bool AMN, bool BMN, int EPI, int OPT, bool X3>
static int launch_tc_gemm_pair(const float* A, int lda, int M, const float* Bm, int ldb, int N, int R, float* C, int ldc,
                               const tc::Epilogue& ep, cudaStream_t st) {
  CUtensorMap ma, mb;
  if (AMN ? tc::make_map(&ma, A, R, M, lda, 32, true) : tc::make_map(&ma, A, M, R, lda, tc::BM)) return 1;
  if (BMN ? tc::make_map(&mb, Bm, R, N, ldb, 32, true) : tc::make_map(&mb, Bm, N, R, ldb, 128)) return 1;
  constexpr int STG = X3 ? 4 : 6;
  auto kern = tc::gemm_tf32_kernel<AMN, BMN, EPI, OPT, X3, 256, STG, 2>;
  constexpr size_t smem = tc::smem_bytes<X3, 128, STG>();
  if (set_smem(kern, smem)) return 1;
  const int kblocks = ceil_div(R, tc::BK);
  dim3 grid(2 * ceil_div(M, 256), ceil_div(N, 256), 1);
  launch_kc(kern, 2, grid, dim3(tc::kThreads), smem, st, ma, mb, M, N, R, C, ldc, ep, kblocks, (size_t)0);
  SN_LAUNCH_CHECK();
  return 0;
}
// precision 1 = TF32 (one pass), anything else that reaches here = 3xTF32 (fp32-accurate)
template <bool AMN, bool BMN, int EPI, int OPT>
static int launch_tc_gemm(const float* A, int lda, int M, const float* Bm, int ldb, int N, int R, float* C, int ldc,
                          const tc::Epilogue& ep, cudaStream_t st, int splits = 1, size_t split_stride = 0) {
  const long long sms = sm_count(cur_device());
  if (g_matmul_precision.load() == 1) {
    // CTA pairs once a wave of them exists (SN_TF32_PAIR=0 disables, 1 forces)
    static const int tpair = getenv("SN_TF32_PAIR") ? atoi(getenv("SN_TF32_PAIR")) : -1;
    if (splits <= 1 && (tpair == 1 || (tpair < 0 && 2LL * ceil_div(M, 256) * ceil_div(N, 256) >= sms && R >= 256)))
      return launch_tc_gemm_pair<AMN, BMN, EPI, OPT, false>(A, lda, M, Bm, ldb, N, R, C, ldc, ep, st);
    // wide tiles once there is enough work to fill the SMs with them (SN_TC_BN=128 forces narrow)
    static const bool wide_ok = !(getenv("SN_TC_BN") && atoi(getenv("SN_TC_BN")) == 128);
    if (wide_ok && (long long)ceil_div(N, 256) * ceil_div(M, tc::BM) >= sms && R >= 256)
      return launch_tc_gemm_p<AMN, BMN, EPI, OPT, false, 256>(A, lda, M, Bm, ldb, N, R, C, ldc, ep, st, splits, split_stride);
    return launch_tc_gemm_p<AMN, BMN, EPI, OPT, false, 128>(A, lda, M, Bm, ldb, N, R, C, ldc, ep, st, splits, split_stride);
  }
  if constexpr (EPI == tc::EPI_UPDATE) {
    // two CTAs per SM with a 2-deep ring each (SN_DW_STAGES=4: one CTA with a 4-deep ring)
    static const bool deep = getenv("SN_DW_STAGES") && atoi(getenv("SN_DW_STAGES")) == 4;
    if (!deep) return launch_tc_gemm_p<AMN, BMN, EPI, OPT, true, 128, 2>(A, lda, M, Bm, ldb, N, R, C, ldc, ep, st, splits, split_stride);
  } else {
    // CTA pairs (256 x 256 per pair) when whole waves of them cost less than whole waves of single
    // 128 x 128 tiles: measured per k block of a wave 1.32 us against 0.76 us for a quarter of the
    // work (H 8192, batch 1024 and 4096; profiles/r2y_gemm_breakdown_pairs.txt).  SN_X3_PAIR=1 forces
    // pairs, 0 disables them.
    static const int pair = getenv("SN_X3_PAIR") ? atoi(getenv("SN_X3_PAIR")) : -1;
    const long long pair_ctas = 2LL * ceil_div(M, 256) * ceil_div(N, 256);
    const long long single_ctas = (long long)ceil_div(M, tc::BM) * ceil_div(N, 128);
    const double cost_pair = 1.32 * (double)((pair_ctas + sms - 1) / sms), cost_single = 0.76 * (double)((single_ctas + sms - 1) / sms);
    if (splits <= 1 && (pair == 1 || (pair < 0 && cost_pair < cost_single)))
      return launch_tc_gemm_pair<AMN, BMN, EPI, OPT, true>(A, lda, M, Bm, ldb, N, R, C, ldc, ep, st);
    // the same pairing pays for the other forms once the grid is several waves deep (batch 4096, H 8192:
    // 186 -> 220 TFLOP/s); on a grid of one or two waves it only halves the SMs in use
    static const int force = getenv("SN_X3_STAGES") ? atoi(getenv("SN_X3_STAGES")) : 0;
    const long long ctas = (long long)ceil_div(N, 128) * ceil_div(M, tc::BM) * std::max(1, splits);
    if (force == 2 || (force == 0 && ctas >= 4 * sms))
      return launch_tc_gemm_p<AMN, BMN, EPI, OPT, true, 128, 2>(A, lda, M, Bm, ldb, N, R, C, ldc, ep, st, splits, split_stride);
  }
  return launch_tc_gemm_p<AMN, BMN, EPI, OPT, true, 128>(A, lda, M, Bm, ldb, N, R, C, ldc, ep, st, splits, split_stride);
}

// Split-K for GEMMs whose output has too few 128x128 tiles to occupy the SMs (a narrow layer at a
// moderate batch: 26 tiles for [256 x 1636]): number of reduction slices, each at least 8 k blocks deep.
// The slices write raw partial tiles; an element-wise launch adds them and applies the epilogue.
static int plan_splitk(int M, int N, int R, int sms, size_t scratch_floats) {
  static const bool off = getenv("SN_NO_SPLITK") && atoi(getenv("SN_NO_SPLITK")) == 1;
  const int tiles = ceil_div(M, tc::BM) * ceil_div(N, 128), kblocks = ceil_div(R, tc::BK);
  if (off || tiles * 2 > sms) return 1;
  int s = std::min(sms / tiles, kblocks / 8);
  while (s > 1 && (size_t)s * M * N > scratch_floats) --s;
  return std::max(1, s);
}

// tcgen05 skinny forward (gemm_tc.cuh): grid and partial-slot count for a [N x K] layer
struct SkPlan { int grid, qmax; };
constexpr size_t kSkinnyTcMinElems = 16u << 20;     // below this the launch is latency, not bandwidth
std::atomic<int> g_skinny_tensor{-1};               // -1: not set yet (SN_TC_SKINNY=1 in the environment turns it on)
static bool sk_enabled() {
  int v = g_skinny_tensor.load();
  if (v < 0) {
    v = (getenv("SN_TC_SKINNY") && atoi(getenv("SN_TC_SKINNY")) != 0) ? 1 : 0;
    g_skinny_tensor.store(v);
  }
  return v == 1 && g_matmul_precision.load() != 2;
}
std::atomic<int> g_persistent_tail{-1};             // -1: not set yet (SN_TAIL=1 in the environment turns it on)
static bool tail_enabled() {
  int v = g_persistent_tail.load();
  if (v < 0) {
    v = (getenv("SN_TAIL") && atoi(getenv("SN_TAIL")) != 0) ? 1 : 0;
    g_persistent_tail.store(v);
  }
  return v == 1;
}
static SkPlan plan_sk(int K, int N, int sms) {
  const long long KB = ceil_div(K, tc::BK), RT = ceil_div(N, tc::BM), U = KB * RT;
  long long G = std::min<long long>(sms, U);
  G = std::max(G, RT);                              // a range may touch two row tiles at most
  int qmax = 1;
  for (long long rt = 0; rt < RT; ++rt)
    qmax = std::max(qmax, tc::sk_cta_of_unit((rt + 1) * KB - 1, U, G) - tc::sk_cta_of_unit(rt * KB, U, G) + 1);
  return SkPlan{(int)G, qmax};
}
static bool sk_applies(int B, int K, int N, const float* X, int ldx, const float* W) {
  return sk_enabled() && B <= tc::kSkNB && (size_t)N * K >= kSkinnyTcMinElems && tc::usable(X, ldx, W, K, K);
}

static int launch_linear_fwd(const float* X, int ldx, int B, const float* W, const float* bias, int K, int N,
                             int relu, const uint8_t* mask, float keep_prob, float scale, uint64_t seed,
                             uint64_t offset, int layer, float* H, int ldh, float* out2, float* P, int sms,
                             cudaStream_t st, Profiler* prof = nullptr, const CeTail* ce = nullptr,
                             bool* ce_fused = nullptr, int l2_keep_w = 0, bool prefetch_w = false,
                             int skinny_max = kSkinnyFwdMaxB, float* sk = nullptr, size_t sk_floats = 0) {
  // skinny_max: batches up to this many rows go through the fp32 weight-stream kernels in row tiles
  // of 16 (every row's arithmetic is then independent of the batch it arrives in)
  if (ce_fused) *ce_fused = false;
  if (B <= skinny_max && (size_t)N * K <= kSmallLayerElems && K <= kSmallLayerMaxK) {
    dim3 grid(ceil_div(N, kHeadRows), ceil_div(B, kBT));
    const size_t smem_small = 0;
    cudaEvent_t pe = prof ? prof->begin(st, 0, layer) : nullptr;
    // the loss can ride along when the logits are dense rows (the tail reads H as [rows][C])
    const bool tail = ce && ce->labels && ldh == N;
    launch_k(fwd_small_kernel, dim3(grid), dim3(256), smem_small, st, X, ldx, B, W, K, N, K, bias, relu, mask, keep_prob, scale, seed, offset,
                                          layer, H, ldh, out2, layer > 0 ? 1 : 0, tail ? *ce : CeTail{});
    if (ce_fused) *ce_fused = tail;
    Profiler::end(pe, st);
    SN_LAUNCH_CHECK();
  } else if (sk_applies(B, K, N, X, ldx, W)) {
    // big layer, B <= 16: the weight stream goes through the tensor cores (3xTF32, fp32-accurate)
    const SkPlan p = plan_sk(K, N, sms);
    CUtensorMap mw, mx;
    if (tc::make_map(&mw, W, N, K, K, tc::BM) || tc::make_map(&mx, X, B, K, ldx, tc::kSkNB)) return 1;
    if (set_smem(tc::skinny_fwd_x3_kernel, tc::sk_smem_bytes())) return 1;
    cudaEvent_t pe = prof ? prof->begin(st, 0, layer) : nullptr;
    launch_k(tc::skinny_fwd_x3_kernel, dim3(p.grid), dim3(tc::kThreads), tc::sk_smem_bytes(), st, mw, mx, N, K, B, P, kBT, p.qmax,
             prefetch_w ? 1 : 0, l2_keep_w);
    Profiler::end(pe, st);
    SN_LAUNCH_CHECK();
    launch_k(fwd_epilogue_kernel, dim3(ew_grid((size_t)B * N, sms)), dim3(256), 0, st, P, p.qmax, kBT, B, N, bias, relu, mask, keep_prob,
             scale, seed, offset, layer, H, ldh, out2, 0, 0);
    SN_LAUNCH_CHECK();
  } else if (B <= skinny_max && (K % 4) == 0) {
    FwdPlan p = plan_fwd(K, N, sms);
    int tiles = ceil_div(B, kBT), Bpad = tiles * kBT;
    size_t smem = (size_t)kBT * p.kc * 4 + (size_t)8 * kFwdRing * 8 * 32 * 16;   // X chunk + per-warp W rings
    dim3 grid(p.nx, p.ksplit, tiles);
    cudaEvent_t pe = prof ? prof->begin(st, 0, layer) : nullptr;
    // the weight ring starts before the dependency wait unless the previous launch may have written W
    if (fwd_rows() == 4) {
      if (set_smem(fwd_skinny_kernel<4>, smem)) return 1;
      launch_k(fwd_skinny_kernel<4>, dim3(grid), dim3(512), smem, st, X, ldx, B, W, K, N, K, p.kc, P, Bpad, p.ksplit,
               (layer > 0 || prefetch_w) ? 1 : 0, l2_keep_w);
    } else {
      if (set_smem(fwd_skinny_kernel<8>, smem)) return 1;
      launch_k(fwd_skinny_kernel<8>, dim3(grid), dim3(256), smem, st, X, ldx, B, W, K, N, K, p.kc, P, Bpad, p.ksplit,
               (layer > 0 || prefetch_w) ? 1 : 0, l2_keep_w);
    }
    Profiler::end(pe, st);
    SN_LAUNCH_CHECK();
    launch_k(fwd_epilogue_kernel, dim3(ew_grid((size_t)B * N, sms)), dim3(256), 0, st, P, kFwdSlots * p.ksplit, Bpad, B, N, bias, relu, mask,
                                                                    keep_prob, scale, seed, offset, layer, H, ldh, out2, 0, 0);
    SN_LAUNCH_CHECK();
  } else if (g_matmul_precision.load() != 2 && tc::usable(X, ldx, W, K, K)) {
    // tensor-bound regime: tcgen05 TF32 GEMM with the bias / ReLU / dropout epilogue fused
    tc::Epilogue ep{};
    ep.bias = bias; ep.relu = relu; ep.mask = mask; ep.keep_prob = keep_prob; ep.scale = scale;
    ep.seed = seed; ep.offset = offset; ep.layer = layer; ep.out2 = out2;
    cudaEvent_t pe = prof ? prof->begin(st, 0, layer) : nullptr;
    const int splits = sk ? plan_splitk(B, N, K, sms, sk_floats) : 1;
    int rc;
    if (splits > 1) {
      tc::Epilogue raw{};
      rc = launch_tc_gemm<false, false, tc::EPI_GATE, 0>(X, ldx, B, W, K, N, K, sk, N, raw, st, splits, (size_t)B * N);
      if (!rc) {
        const int nsl = ceil_div(ceil_div(K, tc::BK), ceil_div(ceil_div(K, tc::BK), splits));
        launch_k(fwd_epilogue_kernel, dim3(ew_grid((size_t)B * N, sms)), dim3(256), 0, st, sk, nsl, B, B, N, bias, relu, mask, keep_prob,
                 scale, seed, offset, layer, H, ldh, out2, 0, 0);
        SN_LAUNCH_CHECK();
      }
    } else {
      rc = launch_tc_gemm<false, false, tc::EPI_FWD, 0>(X, ldx, B, W, K, N, K, H, ldh, ep, st);
    }
    Profiler::end(pe, st);
    if (rc) return rc;
  } else {
    dim3 grid(ceil_div(N, 64), ceil_div(B, 64));
    launch_k(sgemm_strided_kernel, dim3(grid), dim3(256), 0, st, X, ldx, 1, W, 1, K, H, ldh, B, N, K);
    SN_LAUNCH_CHECK();
    SN_REQUIRE(ldh == N, "general-batch forward needs a dense output");
    launch_k(fwd_epilogue_kernel, dim3(ew_grid((size_t)B * N, sms)), dim3(256), 0, st, H, 1, 0, B, N, bias, relu, mask, keep_prob,
                                                                    scale, seed, offset, layer, H, ldh, out2, 0, 0);
    SN_LAUNCH_CHECK();
  }
  return 0;
}

static size_t fwd_partial_floats(int B, int K, int N, int sms, int skinny_max = kSkinnyFwdMaxB) {
  if (B > skinny_max || ((size_t)N * K <= kSmallLayerElems && K <= kSmallLayerMaxK)) return 0;
  if ((K % 4) != 0) return 0;
  FwdPlan p = plan_fwd(K, N, sms);
  size_t need = (size_t)kFwdSlots * p.ksplit * round_up(B, kBT) * N;
  if ((size_t)N * K >= kSkinnyTcMinElems) need = std::max(need, (size_t)plan_sk(K, N, sms).qmax * kBT * N);
  return need;
}

}  // namespace sn

using namespace sn;

// A handle's calls run on the handle's device and leave the caller's current device as it was (the
// host program - torch - keeps its own idea of the current device).
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess) { cudaGetLastError(); cur = -1; }
    if (cur != dev) {
      cudaSetDevice(dev);
      prev = cur;
    }
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
  DeviceGuard(const DeviceGuard&) = delete;
  DeviceGuard& operator=(const DeviceGuard&) = delete;
};


// =============================================================================== handles
struct sn_bob {
  int device = 0, L = 0, relu_last = 0, n_slots = 1, max_batch = 16, Bpad = 16, sms = 148;
  float dropout_p = 0.f;
  Hyper hyper;
  std::vector<int> dims;
  std::vector<float*> W, b;
  struct Slot {
    std::vector<float*> s1, s2;
    int64_t step = 0;
  };
  std::vector<Slot> slots;
  std::vector<float*> H, dZ;   // index l = 1..L ([Bpad][dims[l]])
  float* P = nullptr;          // forward split-K partials
  int stream_rows = kSkinnyFwdMaxB;   // forward batches up to this many rows use the weight-stream kernels
  float* PA = nullptr;         // backward dA partials
  float* SK = nullptr;         // split-K partial tiles of the tcgen05 GEMMs (max_batch > 16)
  size_t SK_floats = 0;
  float* BG = nullptr;         // bias-gradient chunk sums [<= 128 chunks][widest layer] (max_batch > 16)
  size_t PA_floats = 0;
  float* dWbuf = nullptr;      // general-batch path only (lazy)
  float* row_loss = nullptr;
  unsigned* ticket = nullptr;
  float* const* cut_table = nullptr;   // tensor-parallel head: per-consumer destinations of the cut gradient's columns
  int cut_cols = 1;
  int cut_first = -1;                  // >= 0: consumer 0 owns this many leading columns, the others cut_cols each
  // tensor-parallel head: produce and hand out the cut gradient BEFORE the first layer's update
  // traversal (a read-only pass over its weights), then call this (the flag release), then update
  std::function<int(cudaStream_t)> after_cut;
  bool stashed = false;
  int stashed_B = 0;
  int heads = 1;
  int row_loss_rows = 0;
  Profiler prof;
  float* PF = nullptr;         // fused next-forward partials of layer 0 [k slices][16][dims[1]] (lazy)
  bool pf_valid = false;       // PF holds layer 0's pre-activation partials for the coming step
  int pf_parts = 0;            // how many partials per output PF holds (k groups, or 128-wide k slices)
  float* Ptail = nullptr;      // tail kernel: the second layer's split-K partials (lazy)
  unsigned* tail_bar = nullptr;   // tail kernel: grid-barrier arrival counter (never reset)
  unsigned tail_arrivals = 0;     // arrivals all earlier launches have added to it
  std::vector<int> l2_keep;    // per layer: tag the weight lines evict_last in L2 (see l2_policy)
};

struct sn_alice {
  int device = 0, max_batch = 16, sms = 148;
  Hyper hyper;
  float *w = nullptr, *b = nullptr;
  float *s1 = nullptr, *s2 = nullptr;   // [320] each: 288 weights then 32 biases
  int64_t step = 0;
  float* partial = nullptr;
  unsigned* ticket = nullptr;           // [0] conv ticket, [1] CE ticket
  float* act = nullptr;                 // local step: [max_batch][5408]
  float* dact = nullptr;
  float* row_loss = nullptr;
  int gx = 1;
};

// One row slice of a Linear layer with its optimizer state: the unit of the tensor-parallel first
// layer (include/splitnn.h, sn_shard_*).
struct sn_shard {
  int device = 0, K = 0, n_total = 0, n_off = 0, n_rows = 0, relu = 1, max_batch = 16, sms = 148;
  float dropout_p = 0.f;
  Hyper hyper;
  int64_t step = 0;
  float *W = nullptr, *b = nullptr;
  float *s1 = nullptr, *s2 = nullptr, *bs1 = nullptr, *bs2 = nullptr;
  float* P = nullptr;        // forward split-K partials
  float* PF = nullptr;       // fused next-forward partials [k slices][16][n_rows]
  float* Hloc = nullptr;     // dense copy of the slice's activation [16][n_rows] (the backward's gate)
  bool pf_valid = false, h_valid = false;
  int pf_parts = 0;
  int l2_keep = 0;
};

static int conv_bwd_gx(int B, int sms) { return std::max(1, std::min(B, sms * 2)); }

// Alice's forward: the 16-sample latency kernel, or - from kConvStreamMinB samples up (a whole
// shard: SISA's cache fill, evaluation) - the persistent streaming kernel.  Same numbers either way.
static int conv_stream_min() {
  static const int v = getenv("SN_CONV_STREAM_MIN") ? atoi(getenv("SN_CONV_STREAM_MIN")) : 512;
  return v;
}
static int launch_conv_fwd(const float* x, const float* w, const float* b, int B, float* act, size_t ld_act, int sms,
                           cudaStream_t st, float* x_mirror = nullptr, const long long* y_src = nullptr,
                           long long* y_dst = nullptr) {
  if (B >= conv_stream_min() && !x_mirror && !y_dst) {
    // SN_CONV_STREAM=cta selects the CTA-per-image form (A/B timing); default: warp-autonomous images
    static const bool warp_form = !(getenv("SN_CONV_STREAM") && !strcmp(getenv("SN_CONV_STREAM"), "cta"));
    // resident CTAs per SM of the CTA form: 3 (54 registers) by default; SN_CONV_CTAS=2|4 for A/B timing
    static const int ctas = getenv("SN_CONV_CTAS") ? atoi(getenv("SN_CONV_CTAS")) : 3;
    if (warp_form) {
      if (set_smem(conv_pool_stream_warp_kernel, kStreamWarpSmem)) return 1;
      static const int wctas = getenv("SN_CONV_WCTAS") ? std::max(1, atoi(getenv("SN_CONV_WCTAS"))) : 2;   // 2: 0.162 ms, 3: 0.172, 4: 0.168
      launch_k(conv_pool_stream_warp_kernel, dim3(std::min(ceil_div(B, kStreamWarps), sms * wctas)), dim3(32 * kStreamWarps),
               kStreamWarpSmem, st, x, w, b, B, act, ld_act);
    } else if (ctas == 2)
      launch_k(conv_pool_stream_kernel<2>, dim3(std::min(B, sms * 2)), dim3(kStreamThreads), 0, st, x, w, b, B, act, ld_act);
    else if (ctas == 4)
      launch_k(conv_pool_stream_kernel<4>, dim3(std::min(B, sms * 4)), dim3(kStreamThreads), 0, st, x, w, b, B, act, ld_act);
    else
      launch_k(conv_pool_stream_kernel<3>, dim3(std::min(B, sms * 3)), dim3(kStreamThreads), 0, st, x, w, b, B, act, ld_act);
  } else {
    launch_k(conv_pool_fwd_kernel, dim3(std::min(B, sms * 4), 4), dim3(256), 0, st, x, w, b, B, act, ld_act, x_mirror, y_src, y_dst);
  }
  SN_LAUNCH_CHECK();
  return 0;
}

extern "C" {

int sn_version(void) { return 1; }
int sn_set_matmul_precision(int mode) {
  SN_REQUIRE(mode >= 0 && mode <= 2, "sn_set_matmul_precision: mode must be 0 (3xTF32), 1 (TF32) or 2 (fp32 SIMT)");
  g_matmul_precision.store(mode);
  return 0;
}
int sn_get_matmul_precision(void) { return g_matmul_precision.load(); }
int sn_set_skinny_tensor_forward(int on) {
  g_skinny_tensor.store(on ? 1 : 0);
  return 0;
}
int sn_get_skinny_tensor_forward(void) { return sk_enabled() ? 1 : 0; }
int sn_set_persistent_tail(int on) {
  g_persistent_tail.store(on ? 1 : 0);
  return 0;
}
int sn_get_persistent_tail(void) { return tail_enabled() ? 1 : 0; }
int sn_set_bwd_tiled_max(int rows) {
  SN_REQUIRE(rows >= kBT && rows <= kSkinnyFwdMaxB, "sn_set_bwd_tiled_max: rows must be in [16, 64]");
  g_bwd_tiled_max.store(rows);
  return 0;
}
int sn_get_bwd_tiled_max(void) { return bwd_tiled_max(); }
const char* sn_last_error(void) { return g_last_error.c_str(); }
uint64_t sn_launch_count(void) { return g_launches.load(); }

// ------------------------------------------------------------------ stateless operators
int sn_conv_pool_forward(const float* x, const float* w, const float* b, int B, float* act, void* stream) {
  SN_REQUIRE(x && w && b && act && B > 0, "sn_conv_pool_forward: bad arguments");
  int dev = 0;
  SN_CUDA(cudaGetDevice(&dev));
  return launch_conv_fwd(x, w, b, B, act, kD0, sm_count(dev), (cudaStream_t)stream);
}

int sn_conv_weights_copy(float* dst_w, float* dst_b, const float* src_w, const float* src_b, void* stream) {
  SN_REQUIRE(dst_w && dst_b && src_w && src_b, "sn_conv_weights_copy: bad arguments");
  launch_k(conv_weights_copy_kernel, dim3(1), dim3(128), 0, (cudaStream_t)stream, dst_w, dst_b, src_w, src_b);
  SN_LAUNCH_CHECK();
  mark_written(dst_w, 288 * 4);
  return 0;
}

size_t sn_conv_pool_backward_scratch_bytes(int B) {
  int gx = conv_bwd_gx(std::max(B, 1), sm_count(cur_device()));
  return (size_t)gx * kConvParams * 4 + 256;
}

int sn_conv_pool_backward(const float* x, const float* w, const float* b, const float* dact, int B, float* dw,
                          float* db, void* scratch, void* stream) {
  SN_REQUIRE(x && w && b && dact && dw && db && scratch && B > 0, "sn_conv_pool_backward: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  int gx = conv_bwd_gx(B, sm_count(cur_device()));
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  SN_CUDA(cudaMemsetAsync(ticket, 0, 256, st));
  OptParams op{};
  launch_k(conv_pool_bwd_kernel<0>, dim3(gx, 4), dim3(256), 0, st, x, const_cast<float*>(w), const_cast<float*>(b), dact, B,
                                                      partial, ticket, dw, db, nullptr, nullptr, op, 0);
  SN_LAUNCH_CHECK();
  return 0;
}

size_t sn_linear_forward_scratch_bytes(int B, int K, int N) {
  if (B <= 0 || K <= 0 || N <= 0) return 16;
  return std::max<size_t>(16, fwd_partial_floats(B, K, N, sm_count(cur_device())) * 4);
}

int sn_linear_forward(const float* x, const float* W, const float* bias, int B, int K, int N, int relu,
                      const uint8_t* keep_mask, float mask_scale, float* y, void* scratch, size_t scratch_bytes,
                      void* stream) {
  SN_REQUIRE(x && W && bias && y && B > 0 && K > 0 && N > 0, "sn_linear_forward: bad arguments");
  int dev = 0;
  SN_CUDA(cudaGetDevice(&dev));
  int sms = sm_count(dev);
  size_t need = fwd_partial_floats(B, K, N, sms) * 4;
  SN_REQUIRE(need == 0 || (scratch && scratch_bytes >= need), "sn_linear_forward: scratch too small");
  return launch_linear_fwd(x, K, B, W, bias, K, N, relu, keep_mask, 1.f, mask_scale, 0, 0, 0, y, N, nullptr,
                           reinterpret_cast<float*>(scratch), sms, (cudaStream_t)stream);
}

// ---------------------------------------------------------------- allocation-free linear forward
// Scratch for sn_linear_forward_noalloc: a fixed array inside the loaded module, so the entry point
// never calls an allocator.  4 MiB; every GPU that loads the library carries it until the process
// exits.  On the host `g_noalloc_scratch` is only a handle for the symbol, not a device address:
// the address is looked up once per device with cudaGetSymbolAddress (below) and that pointer is
// what the kernels receive.
constexpr size_t kNoAllocScratchBytes = 4u << 20;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchBytes / sizeof(float)];

static std::mutex g_noalloc_mu;
static float* g_noalloc_ptr[64] = {nullptr};

// Device address of the array on the CURRENT device (looked up on first use per device).
static int noalloc_scratch(float** out, int* device) {
  int dev = 0;
  SN_CUDA(cudaGetDevice(&dev));
  SN_REQUIRE(dev >= 0 && dev < 64, "sn_linear_forward_noalloc: device index out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    float* p = nullptr;
    SN_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    SN_REQUIRE(p != nullptr, "sn_linear_forward_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  *device = dev;
  return 0;
}

size_t sn_linear_forward_noalloc_max_scratch_bytes(void) { return kNoAllocScratchBytes; }

int sn_linear_forward_noalloc(const float* x, const float* W, const float* bias, int B, int K, int N, int relu,
                              const uint8_t* keep_mask, float mask_scale, float* y, void* stream) {
  SN_REQUIRE(x && W && bias && y && B > 0 && K > 0 && N > 0, "sn_linear_forward_noalloc: bad arguments");
  float* scratch = nullptr;
  int dev = 0;
  if (noalloc_scratch(&scratch, &dev)) return 1;
  const int sms = sm_count(dev);
  const size_t need = fwd_partial_floats(B, K, N, sms) * sizeof(float);
  if (need > kNoAllocScratchBytes) {
    return fail_msg("libsplitnn: sn_linear_forward_noalloc: this shape needs " + std::to_string(need) +
                    " bytes of split-K scratch but the built-in scratch holds " +
                    std::to_string(kNoAllocScratchBytes) + " bytes; use sn_linear_forward with caller scratch");
  }
  // Nothing in the scratch has to start at zero: every partial sum the epilogue reads is written
  // by the forward kernel of the same call first.
  return launch_linear_fwd(x, K, B, W, bias, K, N, relu, keep_mask, 1.f, mask_scale, 0, 0, 0, y, N, nullptr, scratch,
                           sms, (cudaStream_t)stream);
}

static int launch_ce(const float* z, int ldz, const int64_t* y, int B, int C, float* dz, int lddz, float* row_loss,
                     unsigned* ticket, float* loss, float* loss2, cudaStream_t st) {
  constexpr int kWideNV = 24;                       // ce_wide_kernel holds rows of up to 6144 classes in registers
  if (C > 128 && C <= 256 * kWideNV && !loss2) {
    launch_k(ce_wide_kernel<kWideNV>, dim3(B), dim3(256), 0, st, z, ldz, y, B, C, dz, lddz, row_loss, ticket, loss);
  } else {
    launch_k(ce_kernel, dim3(ceil_div(B, 8)), dim3(256), 0, st, z, ldz, y, B, C, dz, lddz, row_loss, ticket, loss, loss2);
  }
  SN_LAUNCH_CHECK();
  return 0;
}

int sn_cross_entropy(const float* logits, const int64_t* labels, int B, int C, float* loss, float* dlogits,
                     void* stream) {
  SN_REQUIRE(logits && labels && loss && B > 0 && C > 0, "sn_cross_entropy: bad arguments");
  // stateless form: row losses and the ticket live in a small per-call allocation
  cudaStream_t st = (cudaStream_t)stream;
  float* tmp = nullptr;
  SN_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&tmp), (size_t)B * 4 + 256, st));
  SN_CUDA(cudaMemsetAsync(tmp, 0, 256, st));
  int rc = launch_ce(logits, C, labels, B, C, dlogits, C, tmp + 64, reinterpret_cast<unsigned*>(tmp), loss, nullptr, st);
  SN_CUDA(cudaFreeAsync(tmp, st));
  return rc;
}

int sn_eval_counts(const float* logits, const int64_t* labels, int B, int C, int64_t omit_label, int64_t* counters6,
                   void* stream) {
  SN_REQUIRE(logits && labels && counters6 && B > 0 && C > 0, "sn_eval_counts: bad arguments");
  launch_k(eval_counts_kernel, dim3(ceil_div(B, 8)), dim3(256), 0, (cudaStream_t)stream, 
      logits, labels, B, C, omit_label, reinterpret_cast<unsigned long long*>(counters6));
  SN_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------ Bob
int sn_bob_create(sn_bob** out, int device, const int* dims, int n_layers, int relu_last, double dropout_p,
                  int optimizer, double lr, double momentum, double beta1, double beta2, double eps,
                  double weight_decay, int n_opt_slots, int max_batch, float* const* W, float* const* b) {
  SN_REQUIRE(out && dims && W && b && n_layers >= 1 && n_layers <= 16, "sn_bob_create: bad arguments");
  SN_REQUIRE(optimizer >= SN_OPT_NONE && optimizer <= SN_OPT_ADAM, "sn_bob_create: unknown optimizer");
  SN_REQUIRE(n_opt_slots >= 1 && max_batch >= 1, "sn_bob_create: need n_opt_slots >= 1 and max_batch >= 1");
  SN_REQUIRE(dropout_p >= 0.0 && dropout_p < 1.0, "sn_bob_create: dropout_p must be in [0,1)");
  int ndev = 0;
  SN_CUDA(cudaGetDeviceCount(&ndev));
  SN_REQUIRE(device >= 0 && device < ndev, "sn_bob_create: no such CUDA device (a B200 is required; there is no CPU path)");
  DeviceGuard device_guard_(device);
  sn_bob* h = new sn_bob();
  h->device = device;
  h->L = n_layers;
  h->relu_last = relu_last;
  h->dropout_p = (float)dropout_p;
  h->hyper.type = optimizer;
  h->hyper.lr = lr;
  h->hyper.momentum = momentum;
  h->hyper.beta1 = beta1;
  h->hyper.beta2 = beta2;
  h->hyper.eps = eps;
  h->hyper.wd = weight_decay;
  h->n_slots = n_opt_slots;
  h->max_batch = max_batch;
  h->Bpad = round_up(max_batch, kBT);
  h->sms = sm_count(device);
  h->dims.assign(dims, dims + n_layers + 1);
  for (int l = 0; l <= n_layers; ++l) {
    if (dims[l] <= 0) { delete h; return fail_msg("libsplitnn: sn_bob_create: non-positive layer width"); }
  }
  h->W.assign(W, W + n_layers);
  h->b.assign(b, b + n_layers);
  auto fail_free = [&](cudaError_t e) { sn_bob_destroy(h); return fail("cudaMalloc", e, __FILE__, __LINE__); };
  cudaError_t e;
  h->H.assign(n_layers + 1, nullptr);
  h->dZ.assign(n_layers + 1, nullptr);
  size_t p_floats = 16, pa_floats = 16;
  for (int l = 0; l < n_layers; ++l) {
    const int K = dims[l], N = dims[l + 1];
    if ((e = cudaMalloc(&h->H[l + 1], (size_t)h->Bpad * N * 4)) != cudaSuccess) return fail_free(e);
    if ((e = cudaMalloc(&h->dZ[l + 1], (size_t)h->Bpad * N * 4)) != cudaSuccess) return fail_free(e);
    p_floats = std::max(p_floats, fwd_partial_floats(std::min(max_batch, kSkinnyFwdMaxB), K, N, h->sms));
    BwdPlan bp = plan_bwd(N, K, h->sms, true, optimizer == SN_OPT_ADAM);
    pa_floats = std::max(pa_floats, (size_t)bp.nsplit * kBT * K);
    if (max_batch > kBT) pa_floats = std::max(pa_floats, (size_t)h->Bpad * K);
  }
  h->PA_floats = pa_floats;
  {
    // The leading rows of the largest weight matrix are tagged evict_last in L2 (l2_policy): they
    // are read by the forward and read+written by the backward of every step, and a tagged set
    // that fits the L2's persisting carve-out survives the streaming traffic in between.  The
    // carve-out is left at the driver's default (23 MB on B200: raising it with cudaDeviceSetLimit
    // slows all other streaming traffic by 20-55%, profiles/micro/l2_keep.cu); the tagged size
    // follows it.  Measured on the vanilla step: 78.8 -> 70.5 us for the fc1 traversal at 20-24 MB,
    // less on either side.  SN_L2_KEEP_MB overrides (0 = off).
    h->l2_keep.assign(n_layers, 0);
    const char* env = getenv("SN_L2_KEEP_MB");
    size_t carve = 0;
    if (cudaDeviceGetLimit(&carve, cudaLimitPersistingL2CacheSize) != cudaSuccess) { carve = 0; cudaGetLastError(); }
    const double keep_mb = env ? atof(env) : 0.95 * (double)carve / (1 << 20);
    int big = 0;
    for (int l = 1; l < n_layers; ++l)
      if ((size_t)dims[l] * dims[l + 1] > (size_t)dims[big] * dims[big + 1]) big = l;
    const double row_mb = dims[big] * 4.0 / (1 << 20);
    if (keep_mb > 0 && (size_t)dims[big] * dims[big + 1] > kSmallLayerElems)
      h->l2_keep[big] = (int)std::min<double>(dims[big + 1], keep_mb / row_mb);
  }
  if ((e = cudaMalloc(&h->P, p_floats * 4)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMalloc(&h->PA, pa_floats * 4)) != cudaSuccess) return fail_free(e);
  if (max_batch > kBT) {
    // a split GEMM has at most sms / slices tiles, so sms full tiles hold any set of partials
    h->SK_floats = (size_t)h->sms * tc::BM * 128;
    if ((e = cudaMalloc(&h->SK, h->SK_floats * 4)) != cudaSuccess) return fail_free(e);
    int widest = 0;
    for (int l = 1; l <= n_layers; ++l) widest = std::max(widest, dims[l]);
    if ((e = cudaMalloc(&h->BG, (size_t)129 * widest * 4)) != cudaSuccess) return fail_free(e);
  }
  h->row_loss_rows = h->Bpad;
  if ((e = cudaMalloc(&h->row_loss, (size_t)h->Bpad * 4)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMalloc(&h->ticket, 256)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMemset(h->ticket, 0, 256)) != cudaSuccess) return fail_free(e);
  h->slots.resize(n_opt_slots);
  if (optimizer != SN_OPT_NONE) {
    for (auto& s : h->slots) {
      s.s1.assign(2 * n_layers, nullptr);
      s.s2.assign(2 * n_layers, nullptr);
      for (int l = 0; l < n_layers; ++l) {
        size_t nw = (size_t)dims[l] * dims[l + 1], nb = dims[l + 1];
        if ((e = cudaMalloc(&s.s1[2 * l], nw * 4)) != cudaSuccess) return fail_free(e);
        if ((e = cudaMalloc(&s.s1[2 * l + 1], nb * 4)) != cudaSuccess) return fail_free(e);
        if ((e = cudaMemset(s.s1[2 * l], 0, nw * 4)) != cudaSuccess) return fail_free(e);
        if ((e = cudaMemset(s.s1[2 * l + 1], 0, nb * 4)) != cudaSuccess) return fail_free(e);
        if (optimizer == SN_OPT_ADAM) {
          if ((e = cudaMalloc(&s.s2[2 * l], nw * 4)) != cudaSuccess) return fail_free(e);
          if ((e = cudaMalloc(&s.s2[2 * l + 1], nb * 4)) != cudaSuccess) return fail_free(e);
          if ((e = cudaMemset(s.s2[2 * l], 0, nw * 4)) != cudaSuccess) return fail_free(e);
          if ((e = cudaMemset(s.s2[2 * l + 1], 0, nb * 4)) != cudaSuccess) return fail_free(e);
        }
      }
    }
  }
  *out = h;
  return 0;
}

int sn_bob_destroy(sn_bob* h) {
  if (!h) return 0;
  DeviceGuard device_guard_(h->device);
  for (auto p : h->H) cudaFree(p);
  for (auto p : h->dZ) cudaFree(p);
  cudaFree(h->P);
  cudaFree(h->PA);
  cudaFree(h->SK);
  cudaFree(h->BG);
  cudaFree(h->dWbuf);
  cudaFree(h->PF);
  cudaFree(h->Ptail);
  cudaFree(h->tail_bar);
  cudaFree(h->row_loss);
  cudaFree(h->ticket);
  h->prof.clear();
  for (auto& s : h->slots) {
    for (auto p : s.s1) cudaFree(p);
    for (auto p : s.s2) cudaFree(p);
  }
  delete h;
  return 0;
}

#define SN_BOB_ENTER(h)                                     \
  SN_REQUIRE(h != nullptr, "null sn_bob handle");           \
  DeviceGuard device_guard_(h->device);

int sn_bob_reset_opt_slot(sn_bob* h, int slot, void* stream) {
  SN_BOB_ENTER(h);
  SN_REQUIRE(slot >= 0 && slot < h->n_slots, "sn_bob_reset_opt_slot: slot out of range");
  auto& s = h->slots[slot];
  s.step = 0;
  for (int l = 0; l < h->L && h->hyper.type != SN_OPT_NONE; ++l) {
    size_t nw = (size_t)h->dims[l] * h->dims[l + 1], nb = h->dims[l + 1];
    SN_CUDA(cudaMemsetAsync(s.s1[2 * l], 0, nw * 4, (cudaStream_t)stream));
    SN_CUDA(cudaMemsetAsync(s.s1[2 * l + 1], 0, nb * 4, (cudaStream_t)stream));
    if (h->hyper.type == SN_OPT_ADAM) {
      SN_CUDA(cudaMemsetAsync(s.s2[2 * l], 0, nw * 4, (cudaStream_t)stream));
      SN_CUDA(cudaMemsetAsync(s.s2[2 * l + 1], 0, nb * 4, (cudaStream_t)stream));
    }
  }
  return 0;
}

int sn_bob_opt_state(sn_bob* h, int slot, int which, int param, float** dptr, size_t* count) {
  SN_BOB_ENTER(h);
  SN_REQUIRE(slot >= 0 && slot < h->n_slots && param >= 0 && param < 2 * h->L && dptr && count,
             "sn_bob_opt_state: bad arguments");
  SN_REQUIRE(h->hyper.type != SN_OPT_NONE && (which == 0 || (which == 1 && h->hyper.type == SN_OPT_ADAM)),
             "sn_bob_opt_state: no such state for this optimizer");
  int l = param / 2;
  *count = (param & 1) ? (size_t)h->dims[l + 1] : (size_t)h->dims[l] * h->dims[l + 1];
  *dptr = which == 0 ? h->slots[slot].s1[param] : h->slots[slot].s2[param];
  return 0;
}
int sn_bob_get_opt_step(sn_bob* h, int slot, int64_t* step) {
  SN_REQUIRE(h && step && slot >= 0 && slot < h->n_slots, "sn_bob_get_opt_step: bad arguments");
  *step = h->slots[slot].step;
  return 0;
}
int sn_bob_set_opt_step(sn_bob* h, int slot, int64_t step) {
  SN_REQUIRE(h && slot >= 0 && slot < h->n_slots && step >= 0, "sn_bob_set_opt_step: bad arguments");
  h->slots[slot].step = step;
  return 0;
}
int sn_bob_set_lr(sn_bob* h, double lr) {
  SN_REQUIRE(h, "null sn_bob handle");
  h->hyper.lr = lr;
  return 0;
}
int sn_bob_profile(sn_bob* h, int enable) {
  SN_BOB_ENTER(h);
  h->prof.clear();
  h->prof.on = enable != 0;
  return 0;
}
int sn_bob_profile_read(sn_bob* h, int kind, int layer, double* total_ms, int* launches) {
  SN_BOB_ENTER(h);
  SN_REQUIRE(total_ms && launches, "sn_bob_profile_read: bad arguments");
  double t = 0;
  int n = 0;
  for (auto& r : h->prof.recs) {
    if (r.kind != kind || r.layer != layer) continue;
    SN_CUDA(cudaEventSynchronize(r.b));
    float ms = 0;
    SN_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
    t += ms;
    ++n;
  }
  *total_ms = t;
  *launches = n;
  return 0;
}
int sn_bob_set_stream_rows(sn_bob* h, int rows, size_t max_scratch_bytes, int* rows_granted) {
  SN_BOB_ENTER(h);
  SN_REQUIRE(rows >= 1, "sn_bob_set_stream_rows: rows must be positive");
  rows = std::min(rows, h->max_batch);
  auto need = [&](int r) {
    size_t f = 16;
    for (int l = 0; l < h->L; ++l) f = std::max(f, fwd_partial_floats(r, h->dims[l], h->dims[l + 1], h->sms, r));
    return f * 4;
  };
  int r = std::max(rows, kSkinnyFwdMaxB);
  r = std::min(round_up(r, kBT), std::max(kSkinnyFwdMaxB, h->max_batch));
  if (max_scratch_bytes)
    while (r > kSkinnyFwdMaxB && need(r) > max_scratch_bytes) r = std::max(kSkinnyFwdMaxB, (r / 2 + kBT - 1) / kBT * kBT);
  const size_t bytes = std::max(need(r), need(std::min(h->max_batch, kSkinnyFwdMaxB)));
  float* p = nullptr;
  SN_CUDA(cudaMalloc(&p, bytes));
  SN_CUDA(cudaDeviceSynchronize());           // the old buffer may still be in use by queued kernels
  cudaFree(h->P);
  h->P = p;
  h->stream_rows = r;
  if (rows_granted) *rows_granted = r;
  return 0;
}

int sn_bob_set_ce_heads(sn_bob* h, int heads) {
  SN_BOB_ENTER(h);
  SN_REQUIRE(heads >= 1 && h->dims[h->L] % heads == 0, "sn_bob_set_ce_heads: heads must divide the output width");
  if (h->Bpad * heads > h->row_loss_rows) {
    SN_CUDA(cudaFree(h->row_loss));
    h->row_loss = nullptr;
    SN_CUDA(cudaMalloc(&h->row_loss, (size_t)h->Bpad * heads * 4));
    h->row_loss_rows = h->Bpad * heads;
  }
  h->heads = heads;
  return 0;
}

static int bob_forward_impl(sn_bob* h, const float* act, int B, int train, const uint8_t* const* masks, uint64_t seed,
                            uint64_t offset, float* out, cudaStream_t st, const CeTail* ce = nullptr,
                            bool* ce_fused = nullptr, int skinny_max = kSkinnyFwdMaxB) {
  if (ce_fused) *ce_fused = false;
  const float keep = 1.f - h->dropout_p, scale = 1.f / keep;
  const float* X = act;
  int ldx = h->dims[0];
  for (int l = 0; l < h->L; ++l) {
    const int K = h->dims[l], N = h->dims[l + 1];
    const bool last = l == h->L - 1;
    const int relu = (!last || h->relu_last) ? 1 : 0;
    const bool drop = train && !last && h->dropout_p > 0.f;
    const uint8_t* mask = (drop && masks) ? masks[l] : nullptr;
    const float kp = (drop && !masks) ? keep : 1.f;
    if (l == 0 && h->pf_valid) {
      // layer 0 of this step was already multiplied out by the previous step's fused traversal
      h->pf_valid = false;
      launch_k(fwd_epilogue_kernel, dim3(ew_grid((size_t)B * N, h->sms)), dim3(256), 0, st, h->PF, h->pf_parts, kBT, B, N,
               h->b[l], relu, mask, kp, scale, seed, offset, l, h->H[l + 1], N, last ? out : nullptr, 0, 0);
      SN_LAUNCH_CHECK();
    } else if (launch_linear_fwd(X, ldx, B, h->W[l], h->b[l], K, N, relu, mask, kp, scale, seed, offset, l, h->H[l + 1], N,
                                 last ? out : nullptr, h->P, h->sms, st, &h->prof, last ? ce : nullptr,
                                 last ? ce_fused : nullptr, h->l2_keep[l], !just_written(h->W[l], (size_t)N * K * 4),
                                 skinny_max, h->SK, h->SK_floats))
      return 1;
    X = h->H[l + 1];
    ldx = N;
  }
  return 0;
}

}  // extern "C" (templates need C++ linkage)
// l_top: the highest layer to process (the tail kernel has already done the layers above 0);
// op_in: the step's optimizer scalars when the caller has already advanced the step counter.
template <int OPT>
static int bob_backward_impl(sn_bob* h, int slot, const float* act, int B, float* cut_grad, cudaStream_t st,
                             const float* next_act = nullptr, int next_B = 0, int l_top = -1,
                             const OptParams* op_in = nullptr) {
  auto& s = h->slots[slot];
  if (!op_in) s.step += 1;
  const OptParams op = op_in ? *op_in : make_opt(h->hyper, s.step);
  const float scale = 1.f / (1.f - h->dropout_p);
  for (int l = (l_top >= 0 ? l_top : h->L - 1); l >= 0; --l) {
    const int K = h->dims[l], N = h->dims[l + 1];
    const float* A = l == 0 ? act : h->H[l];
    const float* dZ = h->dZ[l + 1];
    const bool need_da = l > 0 || cut_grad != nullptr;
    float* dst = l > 0 ? h->dZ[l] : cut_grad;
    const float* gateH = l > 0 ? h->H[l] : nullptr;
    const float gscale = (l > 0 && h->dropout_p > 0.f) ? scale : 1.f;
    // every branch ends with mark_written(W[l]) unless a non-updating launch (the dA reduction) comes last
    const size_t wbytes = (size_t)N * K * 4;
    if (B <= kBT && (size_t)N * K <= kSmallLayerElems && N <= 8 * kSmallBwdRows) {
      // tiny layer: one launch, no n split, gate applied in place (bwd_small_kernel)
      cudaEvent_t pe = h->prof.begin(st, 1, l);
      if (need_da)
        launch_k(bwd_small_kernel<OPT, true>, dim3(ceil_div(K, 32)), dim3(256), 0, st, h->W[l], K, N, K, h->b[l], s.s1[2 * l],
                 s.s2[2 * l], s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B, op, gateH ? 1 : 0, gscale, dst, K);
      else
        launch_k(bwd_small_kernel<OPT, false>, dim3(ceil_div(K, 32)), dim3(256), 0, st, h->W[l], K, N, K, h->b[l], s.s1[2 * l],
                 s.s2[2 * l], s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B, op, 0, 1.f, (float*)nullptr, K);
      Profiler::end(pe, st);
      SN_LAUNCH_CHECK();
      mark_written(h->W[l], wbytes);
    } else if (B <= kBT && (K % 4) == 0) {
      BwdPlan p = plan_bwd(N, K, h->sms, need_da, OPT == 2);
      size_t smem = bwd_smem_bytes(p, need_da, OPT == 2);
      dim3 grid(p.kgroups, p.nsplit);
      if (need_da && l == 0 && h->after_cut) {
        // Tensor-parallel head: the shards wait for this layer's input gradient and for nothing else,
        // and it depends only on dZ and the OLD weights.  So: a read-only pass forms it (4*N*K bytes),
        // its reduction pushes each consumer's columns into that consumer's memory, the flags are
        // released - and only then does the update traversal (16-24 bytes per weight) run, while the
        // shards are already streaming their slices.
        if (set_smem(bwd_update_kernel<4, true>, smem)) return 1;
        launch_k(bwd_update_kernel<4, true>, dim3(grid), dim3(256), smem, st, h->W[l], K, N, K, h->b[l], s.s1[2 * l], s.s2[2 * l],
                 s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B, h->PA, p.kw, p.rows_per_cta, op, (const float*)nullptr, 0,
                 (float*)nullptr, 0, (const float*)nullptr, 1.f);
        SN_LAUNCH_CHECK();
        launch_k(bwd_epilogue_kernel, dim3(ew_grid((size_t)B * K, h->sms)), dim3(256), 0, st, h->PA, p.nsplit, B, K, gateH, K,
                 gscale, dst, K, h->cut_table, h->cut_cols, h->cut_first);
        SN_LAUNCH_CHECK();
        if (int rc = h->after_cut(st)) return rc;
        const BwdPlan pu = plan_bwd(N, K, h->sms, false, OPT == 2);
        const size_t smem_u = bwd_smem_bytes(pu, false, OPT == 2);
        if (set_smem(bwd_update_kernel<OPT, false>, smem_u)) return 1;
        cudaEvent_t pe = h->prof.begin(st, 1, l);
        launch_k(bwd_update_kernel<OPT, false>, dim3(pu.kgroups, pu.nsplit), dim3(256), smem_u, st, h->W[l], K, N, K, h->b[l],
                 s.s1[2 * l], s.s2[2 * l], s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B, h->PA, pu.kw, pu.rows_per_cta, op,
                 (const float*)nullptr, 0, (float*)nullptr, h->l2_keep[l], (const float*)nullptr, 1.f);
        Profiler::end(pe, st);
        SN_LAUNCH_CHECK();
        mark_written(h->W[l], wbytes);
      } else if (need_da) {
        if (set_smem(bwd_update_kernel<OPT, true>, smem)) return 1;
        cudaEvent_t pe = h->prof.begin(st, 1, l);
        launch_k(bwd_update_kernel<OPT, true>, dim3(grid), dim3(256), smem, st, h->W[l], K, N, K, h->b[l], s.s1[2 * l], s.s2[2 * l],
                                                              s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B, h->PA,
                                                              p.kw, p.rows_per_cta, op, (const float*)nullptr, 0, (float*)nullptr,
                                                              h->l2_keep[l], (const float*)nullptr, 1.f);
        Profiler::end(pe, st);
        SN_LAUNCH_CHECK();
        launch_k(bwd_epilogue_kernel, dim3(ew_grid((size_t)B * K, h->sms)), dim3(256), 0, st, h->PA, p.nsplit, B, K, gateH, K,
                 gscale, dst, K, l == 0 ? h->cut_table : (float* const*)nullptr, h->cut_cols, h->cut_first);
        SN_LAUNCH_CHECK();
      } else if (l == 0 && next_act && next_B > 0 && next_B <= kBT && (size_t)N * K > kSmallLayerElems) {
        // SISA: fuse the next step's layer-0 forward into this traversal
        if (!h->PF) SN_CUDA(cudaMalloc(&h->PF, (size_t)ceil_div(K, 128) * kBT * N * 4));
        // the partials are collected in shared memory when they fit next to the ring (one partial per
        // k GROUP then, not per 128-wide slice)
        const int pf_smem = smem + bwd_pf_smem_bytes(p) <= kMaxSmemPerCta && !getenv("SN_NO_PF_SMEM") ? 1 : 0;
        const size_t smem_f = smem + (pf_smem ? bwd_pf_smem_bytes(p) : 0);
        h->pf_parts = pf_smem ? p.kgroups : ceil_div(K, 128);
        cudaEvent_t pe = h->prof.begin(st, 1, l);
        if (pf_smem) {
          if (set_smem(bwd_update_kernel<OPT, false, true, true>, smem_f)) return 1;
          launch_k(bwd_update_kernel<OPT, false, true, true>, dim3(grid), dim3(256), smem_f, st, h->W[l], K, N, K, h->b[l],
                   s.s1[2 * l], s.s2[2 * l], s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B, h->PA, p.kw, p.rows_per_cta, op,
                   next_act, next_B, h->PF, h->l2_keep[l], (const float*)nullptr, 1.f);
        } else {
          if (set_smem(bwd_update_kernel<OPT, false, true, false>, smem_f)) return 1;
          launch_k(bwd_update_kernel<OPT, false, true, false>, dim3(grid), dim3(256), smem_f, st, h->W[l], K, N, K, h->b[l],
                   s.s1[2 * l], s.s2[2 * l], s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B, h->PA, p.kw, p.rows_per_cta, op,
                   next_act, next_B, h->PF, h->l2_keep[l], (const float*)nullptr, 1.f);
        }
        Profiler::end(pe, st);
        SN_LAUNCH_CHECK();
        mark_written(h->W[l], wbytes);
        h->pf_valid = true;
      } else {
        if (set_smem(bwd_update_kernel<OPT, false>, smem)) return 1;
        cudaEvent_t pe = h->prof.begin(st, 1, l);
        launch_k(bwd_update_kernel<OPT, false>, dim3(grid), dim3(256), smem, st, h->W[l], K, N, K, h->b[l], s.s1[2 * l], s.s2[2 * l],
                                                               s.s1[2 * l + 1], s.s2[2 * l + 1], dZ, N, A, K, B,
                                                               h->PA, p.kw, p.rows_per_cta, op, (const float*)nullptr, 0, (float*)nullptr,
                                                               h->l2_keep[l], (const float*)nullptr, 1.f);
        Profiler::end(pe, st);
        SN_LAUNCH_CHECK();
        mark_written(h->W[l], wbytes);
      }
    } else if (B <= bwd_tiled_max() && (K % 4) == 0 && g_matmul_precision.load() != 2) {
      // 17..64 rows (--batch_size 32/64): still a weight stream, far below the tensor ridge.  The
      // batch goes through the same fp32 traversal in tiles of 16 rows: each tile's launch forms its
      // dA rows and ADDS its dW into a gradient buffer (OPT=3); one elementwise pass then applies
      // the optimizer.  ~3x the bytes of the B<=16 form, but 2-3x faster than 3xTF32 GEMMs whose
      // reduction dimension is only B.
      if (!h->dWbuf) {
        size_t mx = 0;
        for (int q = 0; q < h->L; ++q) mx = std::max(mx, (size_t)h->dims[q] * h->dims[q + 1]);
        SN_CUDA(cudaMalloc(&h->dWbuf, mx * 4));
      }
      BwdPlan p = plan_bwd(N, K, h->sms, need_da, false);
      size_t smem = bwd_smem_bytes(p, need_da, false);
      dim3 grid(p.kgroups, p.nsplit);
      OptParams acc = op;
      cudaEvent_t pe = h->prof.begin(st, 1, l);
      for (int r0 = 0; r0 < B; r0 += kBT) {
        const int nt = std::min(kBT, B - r0);
        acc.acc_first = r0 == 0 ? 1 : 0;
        if (need_da) {
          if (set_smem(bwd_update_kernel<3, true>, smem)) return 1;
          launch_k(bwd_update_kernel<3, true>, dim3(grid), dim3(256), smem, st, h->W[l], K, N, K, (float*)nullptr, h->dWbuf,
                   (float*)nullptr, (float*)nullptr, (float*)nullptr, dZ + (size_t)r0 * N, N, A + (size_t)r0 * K, K, nt, h->PA, p.kw,
                   p.rows_per_cta, acc, (const float*)nullptr, 0, (float*)nullptr, 0, (const float*)nullptr, 1.f);
          SN_LAUNCH_CHECK();
          launch_k(bwd_epilogue_kernel, dim3(ew_grid((size_t)nt * K, h->sms)), dim3(256), 0, st, h->PA, p.nsplit, nt, K,
                   gateH ? gateH + (size_t)r0 * K : (const float*)nullptr, K, gscale, dst + (size_t)r0 * K, K,
                   (float* const*)nullptr, 1, -1);
          SN_LAUNCH_CHECK();
        } else {
          if (set_smem(bwd_update_kernel<3, false>, smem)) return 1;
          launch_k(bwd_update_kernel<3, false>, dim3(grid), dim3(256), smem, st, h->W[l], K, N, K, (float*)nullptr, h->dWbuf,
                   (float*)nullptr, (float*)nullptr, (float*)nullptr, dZ + (size_t)r0 * N, N, A + (size_t)r0 * K, K, nt, h->PA, p.kw,
                   p.rows_per_cta, acc, (const float*)nullptr, 0, (float*)nullptr, 0, (const float*)nullptr, 1.f);
          SN_LAUNCH_CHECK();
        }
      }
      launch_k(opt_elementwise_kernel<OPT>, dim3(ew_grid((size_t)N * K, h->sms)), dim3(256), 0, st, h->W[l], h->dWbuf, s.s1[2 * l],
               s.s2[2 * l], (size_t)N * K, op);
      SN_LAUNCH_CHECK();
      launch_k(bias_grad_update_kernel<OPT>, dim3(ceil_div(N, 256)), dim3(256), 0, st, dZ, N, B, N, h->b[l], s.s1[2 * l + 1],
               s.s2[2 * l + 1], op);
      SN_LAUNCH_CHECK();
      Profiler::end(pe, st);
    } else if (g_matmul_precision.load() != 2 && (K % 4) == 0 && (N % 4) == 0) {
      // tensor-bound regime (perf mode): dA, then dW fused with the optimizer, both on tcgen05
      if (need_da) {
        tc::Epilogue ep{};
        ep.gate = gateH; ep.ldgate = K; ep.scale = gscale;
        cudaEvent_t pe = h->prof.begin(st, 2, l);
        const int splits = h->SK ? plan_splitk(B, K, N, h->sms, h->SK_floats) : 1;
        int rc;
        if (splits > 1) {
          tc::Epilogue raw{};
          rc = launch_tc_gemm<false, true, tc::EPI_GATE, 0>(dZ, N, B, h->W[l], K, K, N, h->SK, K, raw, st, splits, (size_t)B * K);
          if (!rc) {
            const int nsl = ceil_div(ceil_div(N, tc::BK), ceil_div(ceil_div(N, tc::BK), splits));
            launch_k(splitk_gate_kernel, dim3(ew_grid((size_t)B * K / 4, h->sms)), dim3(256), 0, st, h->SK, nsl, (size_t)B * K, B, K, gateH,
                     K, gscale, dst, K);
            SN_LAUNCH_CHECK();
          }
        } else {
          rc = launch_tc_gemm<false, true, tc::EPI_GATE, 0>(dZ, N, B, h->W[l], K, K, N, dst, K, ep, st);
        }
        Profiler::end(pe, st);
        if (rc) return rc;
      }
      if (h->BG) {
        const int rpc = std::max(64, round_up(ceil_div(B, 128), 8)), chunks = ceil_div(B, rpc);
        launch_k(tc::bias_grad_partial_kernel, dim3(ceil_div(N, 32), chunks), dim3(256), 0, st, dZ, N, B, N, rpc, h->BG);
        SN_LAUNCH_CHECK();
        launch_k(tc::bias_opt_from_partials_kernel<OPT>, dim3(ceil_div(N, 256)), dim3(256), 0, st, (const float*)h->BG, chunks, N, h->b[l],
                 s.s1[2 * l + 1], s.s2[2 * l + 1], op);
      } else {
        launch_k(bias_grad_update_kernel<OPT>, dim3(ceil_div(N, 256)), dim3(256), 0, st, dZ, N, B, N, h->b[l],
                 s.s1[2 * l + 1], s.s2[2 * l + 1], op);
      }
      SN_LAUNCH_CHECK();
      tc::Epilogue eu{};
      eu.s1 = s.s1[2 * l]; eu.s2 = s.s2[2 * l]; eu.op = op;
      cudaEvent_t pe = h->prof.begin(st, 1, l);
      // a narrow layer at a huge batch (e.g. [48 x 256] at 65536 rows: 2 tiles, 2048 k blocks): split the
      // reduction, then reduce the slices and apply the optimizer in one element-wise pass
      const int splits_w = h->SK ? plan_splitk(N, K, B, h->sms, h->SK_floats) : 1;
      int rc;
      if (splits_w > 1) {
        tc::Epilogue raw{};
        rc = launch_tc_gemm<true, true, tc::EPI_GATE, 0>(dZ, N, N, A, K, K, B, h->SK, K, raw, st, splits_w, (size_t)N * K);
        if (!rc) {
          const int nsl = ceil_div(ceil_div(B, tc::BK), ceil_div(ceil_div(B, tc::BK), splits_w));
          launch_k(tc::splitk_opt_kernel<OPT>, dim3(ew_grid((size_t)N * K / 4, h->sms)), dim3(256), 0, st, h->W[l], (const float*)h->SK, nsl,
                   (size_t)N * K, s.s1[2 * l], s.s2[2 * l], (size_t)N * K / 4, op);
          SN_LAUNCH_CHECK();
        }
      } else {
        rc = launch_tc_gemm<true, true, tc::EPI_UPDATE, OPT>(dZ, N, N, A, K, K, B, h->W[l], K, eu, st);
      }
      Profiler::end(pe, st);
      if (rc) return rc;
      mark_written(h->W[l], wbytes);
    } else {
      // general-batch path: dA from W_old first, then dW into a buffer, then the updates
      if (!h->dWbuf) {
        size_t mx = 0;
        for (int q = 0; q < h->L; ++q) mx = std::max(mx, (size_t)h->dims[q] * h->dims[q + 1]);
        SN_CUDA(cudaMalloc(&h->dWbuf, mx * 4));
      }
      if (need_da) {
        SN_REQUIRE((size_t)B * K <= h->PA_floats, "batch larger than max_batch given to sn_bob_create");
        launch_k(sgemm_strided_kernel, dim3(ceil_div(K, 64), ceil_div(B, 64)), dim3(256), 0, st, dZ, N, 1, h->W[l], K, 1, h->PA, K,
                                                                                    B, K, N);
        SN_LAUNCH_CHECK();
      }
      launch_k(sgemm_strided_kernel, dim3(ceil_div(K, 64), ceil_div(N, 64)), dim3(256), 0, st, dZ, 1, N, A, K, 1, h->dWbuf, K, N,
                                                                                  K, B);
      SN_LAUNCH_CHECK();
      launch_k(opt_elementwise_kernel<OPT>, dim3(ew_grid((size_t)N * K, h->sms)), dim3(256), 0, st, h->W[l], h->dWbuf, s.s1[2 * l],
                                                                                 s.s2[2 * l], (size_t)N * K, op);
      SN_LAUNCH_CHECK();
      launch_k(bias_grad_update_kernel<OPT>, dim3(ceil_div(N, 256)), dim3(256), 0, st, dZ, N, B, N, h->b[l], s.s1[2 * l + 1],
                                                                    s.s2[2 * l + 1], op);
      SN_LAUNCH_CHECK();
      if (need_da) {
        // PA holds one dense [B][K] "split"; bwd_epilogue indexes splits with a 16-row stride
        launch_k(bwd_epilogue_kernel, dim3(ew_grid((size_t)B * K, h->sms)), dim3(256), 0, st, h->PA, 1, B, K, gateH, K, gscale, dst, K,
                 l == 0 ? h->cut_table : (float* const*)nullptr, h->cut_cols, h->cut_first);
        SN_LAUNCH_CHECK();
      }
    }
  }
  return 0;
}

// ---------------------------------------------------------------- persistent tail (tail_kernel.cuh)
// Applies to the reference's server shape class: three layers, a big first layer, a second layer on
// the weight-stream path and a small last layer with at most 128 classes, batch <= 16.
static bool tail_applies(const sn_bob* h, int B, const float* act) {
  if (!tail_enabled() || h->prof.on || h->L != 3 || B > kBT || h->relu_last || h->heads != 1 || fwd_rows() != 8) return false;
  if (h->hyper.type != SN_OPT_SGD_MOMENTUM && h->hyper.type != SN_OPT_ADAM) return false;
  const int d0 = h->dims[0], n1 = h->dims[1], n2 = h->dims[2], n3 = h->dims[3];
  auto small = [](int K, int N) { return (size_t)N * K <= kSmallLayerElems && K <= kSmallLayerMaxK; };
  if (small(d0, n1) || (d0 % 4) != 0) return false;                            // layer 0 must leave split-K partials
  if (small(n1, n2) || (n1 % 4) != 0 || sk_applies(B, n1, n2, act, n1, h->W[1])) return false;
  if (!small(n2, n3) || n3 > 8 * kSmallBwdRows || n3 > 128) return false;
  const FwdPlan f1 = plan_fwd(n1, n2, h->sms);
  const BwdPlan b1 = plan_bwd(n2, n1, h->sms, true, h->hyper.type == SN_OPT_ADAM);
  return f1.nx * f1.ksplit <= h->sms && b1.kgroups * b1.nsplit <= h->sms && ceil_div(n3, kHeadRows) <= h->sms &&
         ceil_div(n2, 32) <= h->sms;
}

template <int OPT>
static int bob_train_step_tail(sn_bob* h, int slot, const float* act, const int64_t* labels, int B,
                               const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset, float* cut_grad,
                               float* logits_out, float* loss_out, cudaStream_t st, const float* next_act, int next_B) {
  const int d0 = h->dims[0], n1 = h->dims[1], n2 = h->dims[2], n3 = h->dims[3];
  const bool drop = h->dropout_p > 0.f;
  TailParams tp{};
  tp.B = B;
  tp.keep = (drop && !drop_masks) ? 1.f - h->dropout_p : 1.f;
  tp.scale = 1.f / (1.f - h->dropout_p);
  tp.seed = seed;
  tp.offset = offset;
  // ---- first layer: leave its split-K partials for the tail's first phase
  if (h->pf_valid) {                     // multiplied out by the previous step's fused traversal
    h->pf_valid = false;
    tp.P0 = h->PF; tp.parts0 = h->pf_parts; tp.bpad0 = kBT;
  } else if (sk_applies(B, d0, n1, act, d0, h->W[0])) {
    const SkPlan p = plan_sk(d0, n1, h->sms);
    CUtensorMap mw, mx;
    if (tc::make_map(&mw, h->W[0], n1, d0, d0, tc::BM) || tc::make_map(&mx, act, B, d0, d0, tc::kSkNB)) return 1;
    if (set_smem(tc::skinny_fwd_x3_kernel, tc::sk_smem_bytes())) return 1;
    launch_k(tc::skinny_fwd_x3_kernel, dim3(p.grid), dim3(tc::kThreads), tc::sk_smem_bytes(), st, mw, mx, n1, d0, B, h->P, kBT, p.qmax,
             just_written(h->W[0], (size_t)n1 * d0 * 4) ? 0 : 1, h->l2_keep[0]);
    SN_LAUNCH_CHECK();
    tp.P0 = h->P; tp.parts0 = p.qmax; tp.bpad0 = kBT;
  } else {
    const FwdPlan p = plan_fwd(d0, n1, h->sms);
    const size_t smem = (size_t)kBT * p.kc * 4 + (size_t)8 * kFwdRing * 8 * 32 * 16;
    if (set_smem(fwd_skinny_kernel<8>, smem)) return 1;
    const int early = just_written(h->W[0], (size_t)n1 * d0 * 4) ? 0 : 1;
    launch_k(fwd_skinny_kernel<8>, dim3(p.nx, p.ksplit, 1), dim3(256), smem, st, act, d0, B, h->W[0], d0, n1, d0, p.kc, h->P, kBT,
             p.ksplit, early, h->l2_keep[0]);
    SN_LAUNCH_CHECK();
    tp.P0 = h->P; tp.parts0 = kFwdSlots * p.ksplit; tp.bpad0 = kBT;
  }
  auto& s = h->slots[slot];
  s.step += 1;
  const OptParams op = make_opt(h->hyper, s.step);
  tp.op = op;
  if (!h->tail_bar) {
    SN_CUDA(cudaMalloc(&h->tail_bar, 256));
    SN_CUDA(cudaMemsetAsync(h->tail_bar, 0, 256, st));
    h->tail_arrivals = 0;
  }
  const FwdPlan f1 = plan_fwd(n1, n2, h->sms);
  const BwdPlan b1 = plan_bwd(n2, n1, h->sms, true, OPT == 2);
  if (!h->Ptail) SN_CUDA(cudaMalloc(&h->Ptail, (size_t)kFwdSlots * f1.ksplit * kBT * n2 * 4));
  SN_REQUIRE((size_t)b1.nsplit * kBT * n1 <= h->PA_floats, "internal: tail kernel input-gradient partials exceed the scratch");
  const int grid = h->sms;
  tp.bar = h->tail_bar;
  tp.bar_base = h->tail_arrivals;
  h->tail_arrivals += (unsigned)kTailBarriers * (unsigned)grid;
  tp.b0 = h->b[0]; tp.mask0 = drop_masks ? drop_masks[0] : nullptr; tp.H1 = h->H[1]; tp.n1 = n1;
  tp.W1 = h->W[1]; tp.n2 = n2; tp.kc1 = f1.kc; tp.ksplit1 = f1.ksplit; tp.nx1 = f1.nx; tp.l2_keep1 = h->l2_keep[1]; tp.P1 = h->Ptail;
  tp.b1 = h->b[1]; tp.mask1 = drop_masks ? drop_masks[1] : nullptr; tp.H2 = h->H[2];
  tp.W2 = h->W[2]; tp.b2 = h->b[2]; tp.n3 = n3; tp.H3 = h->H[3]; tp.logits_out = logits_out;
  tp.ce = CeTail{labels, h->dZ[3], h->row_loss, h->ticket, loss_out, B, n3};
  tp.W2m = h->W[2]; tp.b2m = h->b[2]; tp.s1_2 = s.s1[4]; tp.s2_2 = s.s2[4]; tp.bs1_2 = s.s1[5]; tp.bs2_2 = s.s2[5];
  tp.gscale = drop ? tp.scale : 1.f;
  tp.dZ2 = h->dZ[2];
  tp.W1m = h->W[1]; tp.b1m = h->b[1]; tp.s1_1 = s.s1[2]; tp.s2_1 = s.s2[2]; tp.bs1_1 = s.s1[3]; tp.bs2_1 = s.s2[3];
  tp.kw1 = b1.kw; tp.rows1 = b1.rows_per_cta; tp.kgroups1 = b1.kgroups; tp.nsplit1 = b1.nsplit; tp.PA = h->PA;
  tp.dZ1 = h->dZ[1];
  const size_t smem = std::max((size_t)kBT * f1.kc * 4 + (size_t)8 * kFwdRing * 8 * 32 * 16, bwd_smem_bytes(b1, true, OPT == 2));
  if (set_smem(tail_kernel<OPT>, smem)) return 1;
  launch_k(tail_kernel<OPT>, dim3(grid), dim3(256), smem, st, tp);
  SN_LAUNCH_CHECK();
  // ---- first layer's backward + update (and the cut-layer gradient): the ordinary launches
  return bob_backward_impl<OPT>(h, slot, act, B, cut_grad, st, next_act, next_B, 0, &op);
}

extern "C" {
static int bob_backward_dispatch(sn_bob* h, int slot, const float* act, int B, float* cut_grad, cudaStream_t st,
                                 const float* next_act = nullptr, int next_B = 0) {
  switch (h->hyper.type) {
    case SN_OPT_SGD_MOMENTUM: return bob_backward_impl<1>(h, slot, act, B, cut_grad, st, next_act, next_B);
    case SN_OPT_ADAM: return bob_backward_impl<2>(h, slot, act, B, cut_grad, st, next_act, next_B);
    default: return fail_msg("libsplitnn: this sn_bob was created without an optimizer");
  }
}

static int bob_train_step_impl(sn_bob* h, int slot, const float* act, const int64_t* labels, int B,
                               const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset, float* cut_grad,
                               float* logits_out, float* loss_out, cudaStream_t st, const float* next_act, int next_B) {
  SN_REQUIRE(act && labels && B > 0, "sn_bob_train_step: bad arguments");
  SN_REQUIRE(B <= h->max_batch, "sn_bob_train_step: batch larger than max_batch given to sn_bob_create");
  SN_REQUIRE(slot >= 0 && slot < h->n_slots, "sn_bob_train_step: optimizer slot out of range");
  h->stashed = true;
  h->stashed_B = B;
  if (h->relu_last) return fail_msg("libsplitnn: sn_bob_train_step needs a linear last layer (relu_last=0)");
  if (tail_applies(h, B, act)) {
    const int rc = h->hyper.type == SN_OPT_ADAM
                       ? bob_train_step_tail<2>(h, slot, act, labels, B, drop_masks, seed, offset, cut_grad, logits_out, loss_out, st,
                                                next_act, next_B)
                       : bob_train_step_tail<1>(h, slot, act, labels, B, drop_masks, seed, offset, cut_grad, logits_out, loss_out, st,
                                                next_act, next_B);
    h->stashed = false;
    return rc;
  }
  const int C = h->dims[h->L];
  float* z = h->H[h->L];
  // [B][heads*Ch] row-major is [B*heads][Ch]: multi-head CE is the same kernel over more rows
  const int Ch = C / h->heads;
  // a small last layer computes the loss in its own launch (fwd_small_kernel's tail)
  const CeTail ce{labels, h->dZ[h->L], h->row_loss, h->ticket, loss_out, B * h->heads, Ch};
  bool ce_fused = false;
  if (bob_forward_impl(h, act, B, 1, drop_masks, seed, offset, logits_out, st, &ce, &ce_fused)) return 1;
  if (!ce_fused &&
      launch_ce(z, Ch, labels, B * h->heads, Ch, h->dZ[h->L], Ch, h->row_loss, h->ticket, loss_out, nullptr, st))
    return 1;
  const int rc = bob_backward_dispatch(h, slot, act, B, cut_grad, st, next_act, next_B);
  h->stashed = false;            // the stashed activations have been consumed
  return rc;
}

int sn_bob_train_step(sn_bob* h, int slot, const float* act, const int64_t* labels, int B,
                      const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset, float* cut_grad,
                      float* logits_out, float* loss_out, void* stream) {
  SN_BOB_ENTER(h);
  h->pf_valid = false;
  return bob_train_step_impl(h, slot, act, labels, B, drop_masks, seed, offset, cut_grad, logits_out, loss_out,
                             (cudaStream_t)stream, nullptr, 0);
}

int sn_bob_forward(sn_bob* h, const float* act, int B, int train_mode, const uint8_t* const* drop_masks, uint64_t seed,
                   uint64_t offset, float* out, void* stream) {
  SN_BOB_ENTER(h);
  SN_REQUIRE(act && out && B > 0, "sn_bob_forward: bad arguments");
  SN_REQUIRE(B <= h->max_batch, "sn_bob_forward: batch larger than max_batch given to sn_bob_create");
  h->pf_valid = false;
  h->stashed = train_mode != 0;
  h->stashed_B = B;
  return bob_forward_impl(h, act, B, train_mode, drop_masks, seed, offset, out, (cudaStream_t)stream, nullptr, nullptr,
                          h->stream_rows);
}

int sn_bob_backward_update(sn_bob* h, int slot, const float* act, const float* dout, int B, float* cut_grad,
                           void* stream) {
  SN_BOB_ENTER(h);
  SN_REQUIRE(act && dout && B > 0, "sn_bob_backward_update: bad arguments");
  SN_REQUIRE(h->stashed && h->stashed_B == B, "sn_bob_backward_update: no matching train-mode sn_bob_forward");
  SN_REQUIRE(slot >= 0 && slot < h->n_slots, "sn_bob_backward_update: optimizer slot out of range");
  cudaStream_t st = (cudaStream_t)stream;
  const int C = h->dims[h->L];
  const size_t n = (size_t)B * C;
  if (h->relu_last) {
    launch_k(relu_gate_kernel, dim3(ew_grid(n, h->sms)), dim3(256), 0, st, dout, h->H[h->L], n, h->dZ[h->L]);
    SN_LAUNCH_CHECK();
  } else {
    SN_CUDA(cudaMemcpyAsync(h->dZ[h->L], dout, n * 4, cudaMemcpyDeviceToDevice, st));
  }
  // one backward per stashed forward: a second one would, for a single-layer net, start reading the
  // weights before its predecessor - the first backward - has finished writing them (PDL prologue)
  h->stashed = false;
  return bob_backward_dispatch(h, slot, act, B, cut_grad, st);
}

int sn_bob_server_epoch(sn_bob* h, int slot, const float* cache, const int64_t* labels, const int64_t* batch_offsets,
                        int n_batches, const uint8_t* const* drop_masks, uint64_t seed, uint64_t offset,
                        float* loss_per_step, void* stream) {
  SN_BOB_ENTER(h);
  if (n_batches == 0) return 0;       // an empty shard (an Alice whose data was all unlearned) is a no-op
  SN_REQUIRE(cache && labels && batch_offsets && n_batches > 0, "sn_bob_server_epoch: bad arguments");
  const int d0 = h->dims[0], nm = h->L - 1;
  std::vector<const uint8_t*> m(std::max(nm, 1), nullptr);
  h->pf_valid = false;
  // SN_NO_FUSE_FWD=1 keeps the two-traversal form (A/B timing and parity tests of the fusion)
  static const bool fuse = getenv("SN_NO_FUSE_FWD") == nullptr;
  for (int i = 0; i < n_batches; ++i) {
    const int64_t r0 = batch_offsets[i], r1 = batch_offsets[i + 1];
    SN_REQUIRE(r1 > r0 && r1 - r0 <= h->max_batch, "sn_bob_server_epoch: bad batch_offsets");
    if (drop_masks)
      for (int l = 0; l < nm; ++l) m[l] = drop_masks[l] + (size_t)r0 * h->dims[l + 1];
    // the next batch is already resident in the cache: let this step's layer-0 traversal also
    // produce the next step's layer-0 forward (one read of W per step instead of two)
    const float* next_act = nullptr;
    int next_B = 0;
    if (fuse && i + 1 < n_batches) {
      next_act = cache + (size_t)batch_offsets[i + 1] * d0;
      next_B = (int)(batch_offsets[i + 2] - batch_offsets[i + 1]);
      if (next_B <= 0 || next_B > h->max_batch) { next_act = nullptr; next_B = 0; }
    }
    int rc = bob_train_step_impl(h, slot, cache + (size_t)r0 * d0, labels + r0 * h->heads, (int)(r1 - r0),
                                 drop_masks ? m.data() : nullptr, seed, offset + (uint64_t)i, nullptr, nullptr,
                                 loss_per_step ? loss_per_step + i : nullptr, (cudaStream_t)stream, next_act, next_B);
    if (rc) { h->pf_valid = false; return rc; }
  }
  h->pf_valid = false;
  return 0;
}

// ------------------------------------------------------------------ Alice
int sn_alice_create(sn_alice** out, int device, int optimizer, double lr, double momentum, double beta1,
                    double beta2, double eps, double weight_decay, int max_batch, float* w, float* b) {
  SN_REQUIRE(out && w && b && max_batch >= 1, "sn_alice_create: bad arguments");
  SN_REQUIRE(optimizer >= SN_OPT_NONE && optimizer <= SN_OPT_ADAM, "sn_alice_create: unknown optimizer");
  int ndev = 0;
  SN_CUDA(cudaGetDeviceCount(&ndev));
  SN_REQUIRE(device >= 0 && device < ndev, "sn_alice_create: no such CUDA device (a B200 is required; there is no CPU path)");
  DeviceGuard device_guard_(device);
  sn_alice* h = new sn_alice();
  h->device = device;
  h->max_batch = max_batch;
  h->sms = sm_count(device);
  h->hyper.type = optimizer;
  h->hyper.lr = lr;
  h->hyper.momentum = momentum;
  h->hyper.beta1 = beta1;
  h->hyper.beta2 = beta2;
  h->hyper.eps = eps;
  h->hyper.wd = weight_decay;
  h->w = w;
  h->b = b;
  h->gx = conv_bwd_gx(max_batch, h->sms);
  cudaError_t e;
  auto fail_free = [&](cudaError_t err) { sn_alice_destroy(h); return fail("cudaMalloc", err, __FILE__, __LINE__); };
  if ((e = cudaMalloc(&h->s1, kConvParams * 4)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMalloc(&h->s2, kConvParams * 4)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMemset(h->s1, 0, kConvParams * 4)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMemset(h->s2, 0, kConvParams * 4)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMalloc(&h->partial, (size_t)h->gx * kConvParams * 4)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMalloc(&h->ticket, 256)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMemset(h->ticket, 0, 256)) != cudaSuccess) return fail_free(e);
  if ((e = cudaMalloc(&h->row_loss, (size_t)max_batch * 4)) != cudaSuccess) return fail_free(e);
  *out = h;
  return 0;
}

int sn_alice_destroy(sn_alice* h) {
  if (!h) return 0;
  DeviceGuard device_guard_(h->device);
  cudaFree(h->s1);
  cudaFree(h->s2);
  cudaFree(h->partial);
  cudaFree(h->ticket);
  cudaFree(h->act);
  cudaFree(h->dact);
  cudaFree(h->row_loss);
  delete h;
  return 0;
}

#define SN_ALICE_ENTER(h)                                   \
  SN_REQUIRE(h != nullptr, "null sn_alice handle");         \
  DeviceGuard device_guard_(h->device);

int sn_alice_reset_opt(sn_alice* h, void* stream) {
  SN_ALICE_ENTER(h);
  h->step = 0;
  SN_CUDA(cudaMemsetAsync(h->s1, 0, kConvParams * 4, (cudaStream_t)stream));
  SN_CUDA(cudaMemsetAsync(h->s2, 0, kConvParams * 4, (cudaStream_t)stream));
  return 0;
}
int sn_alice_opt_state(sn_alice* h, int which, int param, float** dptr, size_t* count) {
  SN_REQUIRE(h && dptr && count && (which == 0 || which == 1) && (param == 0 || param == 1),
             "sn_alice_opt_state: bad arguments");
  float* base = which == 0 ? h->s1 : h->s2;
  *dptr = param == 0 ? base : base + 288;
  *count = param == 0 ? 288 : 32;
  return 0;
}
int sn_alice_get_opt_step(sn_alice* h, int64_t* step) {
  SN_REQUIRE(h && step, "sn_alice_get_opt_step: bad arguments");
  *step = h->step;
  return 0;
}
int sn_alice_set_opt_step(sn_alice* h, int64_t step) {
  SN_REQUIRE(h && step >= 0, "sn_alice_set_opt_step: bad arguments");
  h->step = step;
  return 0;
}

int sn_alice_forward(sn_alice* h, const float* x, int B, float* act, void* stream) {
  SN_ALICE_ENTER(h);
  SN_REQUIRE(x && act && B > 0, "sn_alice_forward: bad arguments");
  return launch_conv_fwd(x, h->w, h->b, B, act, kD0, h->sms, (cudaStream_t)stream);
}

int sn_alice_forward_host(sn_alice* h, const float* host_x, const int64_t* host_y, int B, float* act, float* x_dev,
                          int64_t* y_dev, void* stream) {
  SN_ALICE_ENTER(h);
  SN_REQUIRE(host_x && act && x_dev && B > 0 && (!host_y == !y_dev), "sn_alice_forward_host: bad arguments");
  static_assert(sizeof(long long) == sizeof(int64_t), "int64 labels");
  return launch_conv_fwd(host_x, h->w, h->b, B, act, kD0, h->sms, (cudaStream_t)stream, x_dev,
                         reinterpret_cast<const long long*>(host_y), reinterpret_cast<long long*>(y_dev));
}

int sn_alice_forward_ld(sn_alice* h, const float* x, int B, float* act, size_t ld_act, void* stream) {
  SN_ALICE_ENTER(h);
  SN_REQUIRE(x && act && B > 0 && ld_act >= (size_t)kD0, "sn_alice_forward_ld: bad arguments (ld_act >= 5408)");
  return launch_conv_fwd(x, h->w, h->b, B, act, ld_act, h->sms, (cudaStream_t)stream);
}

static int alice_bwd(sn_alice* h, const float* x, const float* dact, int B, cudaStream_t st) {
  SN_REQUIRE(B <= h->max_batch, "batch larger than max_batch given to sn_alice_create");
  h->step += 1;
  const OptParams op = make_opt(h->hyper, h->step);
  dim3 grid(conv_bwd_gx(B, h->sms), 4);
  SN_REQUIRE((int)grid.x <= h->gx, "internal: conv backward grid exceeds scratch");
  // the kernel may read the weights before its dependency wait unless the launch right before it
  // could be this Alice's previous weight update (two backward calls in a row)
  const int early = just_written(h->w, 288 * 4) ? 0 : 1;
  switch (h->hyper.type) {
    case SN_OPT_SGD_MOMENTUM:
      launch_k(conv_pool_bwd_kernel<1>, dim3(grid), dim3(256), 0, st, x, h->w, h->b, dact, B, h->partial, h->ticket, nullptr, nullptr,
                                                   h->s1, h->s2, op, early);
      break;
    case SN_OPT_ADAM:
      launch_k(conv_pool_bwd_kernel<2>, dim3(grid), dim3(256), 0, st, x, h->w, h->b, dact, B, h->partial, h->ticket, nullptr, nullptr,
                                                   h->s1, h->s2, op, early);
      break;
    default:
      return fail_msg("libsplitnn: this sn_alice was created without an optimizer");
  }
  SN_LAUNCH_CHECK();
  mark_written(h->w, 288 * 4);
  return 0;
}

int sn_alice_backward_update(sn_alice* h, const float* x, const float* cut_grad, int B, void* stream) {
  SN_ALICE_ENTER(h);
  SN_REQUIRE(x && cut_grad && B > 0, "sn_alice_backward_update: bad arguments");
  return alice_bwd(h, x, cut_grad, B, (cudaStream_t)stream);
}

int sn_alice_local_step(sn_alice* h, const float* x, const int64_t* labels, int B, float* loss_out, void* stream) {
  SN_ALICE_ENTER(h);
  SN_REQUIRE(x && labels && B > 0 && B <= h->max_batch, "sn_alice_local_step: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (!h->act) {
    SN_CUDA(cudaMalloc(&h->act, (size_t)h->max_batch * kD0 * 4));
    SN_CUDA(cudaMalloc(&h->dact, (size_t)h->max_batch * kD0 * 4));
  }
  if (launch_conv_fwd(x, h->w, h->b, B, h->act, kD0, h->sms, st)) return 1;
  if (launch_ce(h->act, kD0, labels, B, kD0, h->dact, kD0, h->row_loss, h->ticket + 1, loss_out, nullptr, st)) return 1;
  return alice_bwd(h, x, h->dact, B, st);
}

int sn_alice_local_epoch(sn_alice* h, const float* x, const int64_t* labels, int64_t n, int batch, float* loss_per_step,
                         void* stream) {
  SN_ALICE_ENTER(h);
  SN_REQUIRE(x && labels && n >= 0 && batch > 0 && batch <= h->max_batch, "sn_alice_local_epoch: bad arguments");
  int64_t step = 0;
  for (int64_t s = 0; s < n; s += batch, ++step) {
    const int b = (int)std::min<int64_t>(batch, n - s);
    int rc = sn_alice_local_step(h, x + (size_t)s * 784, labels + s, b, loss_per_step ? loss_per_step + step : nullptr,
                                 stream);
    if (rc) return rc;
  }
  return 0;
}

int sn_vanilla_step(sn_alice* alice, sn_bob* bob, int slot, const float* x, const int64_t* labels, int B, uint64_t seed,
                    uint64_t offset, float* act, float* cut_grad, float* loss_out, void* stream) {
  SN_REQUIRE(alice && bob && x && labels && act && cut_grad && B > 0, "sn_vanilla_step: bad arguments");
  SN_REQUIRE(alice->device == bob->device, "sn_vanilla_step: Alice and Bob must share a device (use the separate calls across GPUs)");
  int rc = sn_alice_forward(alice, x, B, act, stream);
  if (rc) return rc;
  rc = sn_bob_train_step(bob, slot, act, labels, B, nullptr, seed, offset, cut_grad, nullptr, loss_out, stream);
  if (rc) return rc;
  return sn_alice_backward_update(alice, x, cut_grad, B, stream);
}

int sn_vanilla_step_host(sn_alice* alice, sn_bob* bob, int slot, const float* host_x, const int64_t* host_y, int B,
                         uint64_t seed, uint64_t offset, float* x_dev, int64_t* y_dev, float* act, float* cut_grad,
                         float* loss_out, void* stream) {
  SN_REQUIRE(alice && bob && host_x && host_y && x_dev && y_dev && act && cut_grad && B > 0, "sn_vanilla_step_host: bad arguments");
  SN_REQUIRE(alice->device == bob->device, "sn_vanilla_step_host: Alice and Bob must share a device");
  int rc = sn_alice_forward_host(alice, host_x, host_y, B, act, x_dev, y_dev, stream);
  if (rc) return rc;
  rc = sn_bob_train_step(bob, slot, act, y_dev, B, nullptr, seed, offset, cut_grad, nullptr, loss_out, stream);
  if (rc) return rc;
  return sn_alice_backward_update(alice, x_dev, cut_grad, B, stream);
}

#ifdef SN_TRACE
// debug build only: copy out and reset the kernel-start trace (see common.cuh)
int sn_trace_read(unsigned long long* out, unsigned* n) {
  SN_CUDA(cudaDeviceSynchronize());
  SN_CUDA(cudaMemcpyFromSymbol(out, sn::g_trace, sizeof(unsigned long long) * 8192));
  SN_CUDA(cudaMemcpyFromSymbol(n, sn::g_trace_n, sizeof(unsigned)));
  const unsigned zero = 0;
  SN_CUDA(cudaMemcpyToSymbol(sn::g_trace_n, &zero, sizeof zero));
  return 0;
}
int sn_trace_phase_read(long long* out16) {
  SN_CUDA(cudaDeviceSynchronize());
  SN_CUDA(cudaMemcpyFromSymbol(out16, sn::g_trace_phase, sizeof(long long) * 16));
  return 0;
}
#endif

// ------------------------------------------------------------------ NVLink plumbing
int sn_p2p_enable_all(int n_devices) {
  int ndev = 0;
  SN_CUDA(cudaGetDeviceCount(&ndev));
  n_devices = std::min(n_devices, ndev);
  int keep = 0;
  SN_CUDA(cudaGetDevice(&keep));
  for (int i = 0; i < n_devices; ++i) {
    SN_CUDA(cudaSetDevice(i));
    for (int j = 0; j < n_devices; ++j) {
      if (i == j) continue;
      int can = 0;
      SN_CUDA(cudaDeviceCanAccessPeer(&can, i, j));
      if (!can) continue;
      cudaError_t e = cudaDeviceEnablePeerAccess(j, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); continue; }
      if (e != cudaSuccess) return fail("cudaDeviceEnablePeerAccess", e, __FILE__, __LINE__);
    }
  }
  SN_CUDA(cudaSetDevice(keep));
  return 0;
}

int sn_p2p_copy(void* dst, int dst_device, const void* src, int src_device, size_t bytes, void* stream) {
  SN_REQUIRE(dst && src, "sn_p2p_copy: bad arguments");
  SN_CUDA(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, bytes, (cudaStream_t)stream));
  return 0;
}

// ------------------------------------------------------------------ tensor-parallel layer slices
#define SN_SHARD_ENTER(h)                                              \
  SN_REQUIRE((h) != nullptr, "null sn_shard handle");                  \
  DeviceGuard device_guard_((h)->device)

int sn_shard_destroy(sn_shard* h) {
  if (!h) return 0;
  DeviceGuard device_guard_(h->device);
  cudaFree(h->s1); cudaFree(h->s2); cudaFree(h->bs1); cudaFree(h->bs2);
  cudaFree(h->P); cudaFree(h->PF); cudaFree(h->Hloc);
  delete h;
  return 0;
}

int sn_shard_create(sn_shard** out, int device, int K, int n_total, int n_off, int n_rows, int relu, double dropout_p,
                    int optimizer, double lr, double momentum, double beta1, double beta2, double eps,
                    double weight_decay, float* W_slice, float* b_slice) {
  SN_REQUIRE(out && W_slice && b_slice, "sn_shard_create: bad arguments");
  SN_REQUIRE(K > 0 && (K % 4) == 0 && n_rows > 0 && n_off >= 0 && n_off + n_rows <= n_total,
             "sn_shard_create: need K % 4 == 0 and 0 <= n_off, n_off + n_rows <= n_total");
  SN_REQUIRE((size_t)n_rows * K > kSmallLayerElems, "sn_shard_create: the slice is too small to be worth a GPU (<= 1 MB of weights)");
  SN_REQUIRE(optimizer == SN_OPT_SGD_MOMENTUM || optimizer == SN_OPT_ADAM, "sn_shard_create: optimizer must be SGD-momentum or Adam");
  SN_REQUIRE(dropout_p >= 0.0 && dropout_p < 1.0, "sn_shard_create: dropout_p must be in [0,1)");
  int ndev = 0;
  SN_CUDA(cudaGetDeviceCount(&ndev));
  SN_REQUIRE(device >= 0 && device < ndev, "sn_shard_create: no such CUDA device");
  DeviceGuard device_guard_(device);
  sn_shard* h = new sn_shard();
  h->device = device; h->K = K; h->n_total = n_total; h->n_off = n_off; h->n_rows = n_rows; h->relu = relu;
  h->dropout_p = (float)dropout_p;
  h->hyper.type = optimizer; h->hyper.lr = lr; h->hyper.momentum = momentum; h->hyper.beta1 = beta1;
  h->hyper.beta2 = beta2; h->hyper.eps = eps; h->hyper.wd = weight_decay;
  h->sms = sm_count(device);
  h->W = W_slice; h->b = b_slice;
  const size_t nw = (size_t)n_rows * K;
  cudaError_t e = cudaSuccess;
  auto alloc0 = [&](float** p, size_t n) {
    if (e == cudaSuccess) e = cudaMalloc(p, n * 4);
    if (e == cudaSuccess) e = cudaMemset(*p, 0, n * 4);
  };
  alloc0(&h->s1, nw); alloc0(&h->bs1, n_rows);
  if (optimizer == SN_OPT_ADAM) { alloc0(&h->s2, nw); alloc0(&h->bs2, n_rows); }
  alloc0(&h->P, std::max<size_t>(16, fwd_partial_floats(kBT, K, n_rows, h->sms)));
  alloc0(&h->PF, (size_t)ceil_div(K, 128) * kBT * n_rows);
  alloc0(&h->Hloc, (size_t)kBT * n_rows);
  if (e != cudaSuccess) { sn_shard_destroy(h); return fail("sn_shard_create", e, __FILE__, __LINE__); }
  size_t carve = 0;
  if (cudaDeviceGetLimit(&carve, cudaLimitPersistingL2CacheSize) != cudaSuccess) { carve = 0; cudaGetLastError(); }
  const char* env = getenv("SN_L2_KEEP_MB");
  const double keep_mb = env ? atof(env) : 0.95 * (double)carve / (1 << 20);
  h->l2_keep = (int)std::min<double>(n_rows, keep_mb / (K * 4.0 / (1 << 20)));
  *out = h;
  return 0;
}

int sn_shard_reset_opt(sn_shard* h, void* stream) {
  SN_SHARD_ENTER(h);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t nw = (size_t)h->n_rows * h->K;
  SN_CUDA(cudaMemsetAsync(h->s1, 0, nw * 4, st));
  SN_CUDA(cudaMemsetAsync(h->bs1, 0, (size_t)h->n_rows * 4, st));
  if (h->s2) {
    SN_CUDA(cudaMemsetAsync(h->s2, 0, nw * 4, st));
    SN_CUDA(cudaMemsetAsync(h->bs2, 0, (size_t)h->n_rows * 4, st));
  }
  h->step = 0;
  h->pf_valid = false;
  return 0;
}

int sn_shard_opt_state(sn_shard* h, int which, int param, float** dptr, size_t* count) {
  SN_REQUIRE(h && dptr && count && (which == 0 || which == 1) && (param == 0 || param == 1), "sn_shard_opt_state: bad arguments");
  SN_REQUIRE(which == 0 || h->s2, "sn_shard_opt_state: no second moment for this optimizer");
  *dptr = which == 0 ? (param == 0 ? h->s1 : h->bs1) : (param == 0 ? h->s2 : h->bs2);
  *count = param == 0 ? (size_t)h->n_rows * h->K : (size_t)h->n_rows;
  return 0;
}
int sn_shard_get_opt_step(sn_shard* h, int64_t* step) {
  SN_REQUIRE(h && step, "sn_shard_get_opt_step: bad arguments");
  *step = h->step;
  return 0;
}
int sn_shard_set_opt_step(sn_shard* h, int64_t step) {
  SN_REQUIRE(h && step >= 0, "sn_shard_set_opt_step: bad arguments");
  h->step = step;
  h->pf_valid = false;       // whoever sets the step has rewritten the state: fused partials are stale
  return 0;
}

int sn_shard_forward(sn_shard* h, const float* X, int B, int train_mode, const uint8_t* keep_mask, uint64_t seed,
                     uint64_t offset, float* H_out, int ld_out, void* stream) {
  SN_SHARD_ENTER(h);
  SN_REQUIRE(X && H_out && B > 0 && B <= kBT && ld_out >= h->n_off + h->n_rows, "sn_shard_forward: bad arguments (B <= 16)");
  cudaStream_t st = (cudaStream_t)stream;
  const int K = h->K, N = h->n_rows;
  const bool drop = train_mode && h->dropout_p > 0.f;
  const float keep = 1.f - h->dropout_p, scale = 1.f / keep;
  const uint8_t* mask = drop ? keep_mask : nullptr;
  const float kp = (drop && !keep_mask) ? keep : 1.f;
  const float* P = h->P;
  int nparts = 0;
  if (h->pf_valid) {                 // the previous backward traversal already multiplied this batch out
    P = h->PF;
    nparts = h->pf_parts;
    h->pf_valid = false;
  } else {
    FwdPlan p = plan_fwd(K, N, h->sms);
    size_t smem = (size_t)kBT * p.kc * 4 + (size_t)8 * kFwdRing * 8 * 32 * 16;
    if (fwd_rows() == 4) {
      if (set_smem(fwd_skinny_kernel<4>, smem)) return 1;
      launch_k(fwd_skinny_kernel<4>, dim3(p.nx, p.ksplit, 1), dim3(512), smem, st, X, K, B, h->W, K, N, K, p.kc, h->P, kBT, p.ksplit, 0,
               h->l2_keep);
    } else {
      if (set_smem(fwd_skinny_kernel<8>, smem)) return 1;
      launch_k(fwd_skinny_kernel<8>, dim3(p.nx, p.ksplit, 1), dim3(256), smem, st, X, K, B, h->W, K, N, K, p.kc, h->P, kBT, p.ksplit, 0,
               h->l2_keep);
    }
    SN_LAUNCH_CHECK();
    nparts = kFwdSlots * p.ksplit;
  }
  launch_k(fwd_epilogue_kernel, dim3(ew_grid((size_t)B * N, h->sms)), dim3(256), 0, st, P, nparts, kBT, B, N, h->b, h->relu, mask, kp,
           scale, seed, offset, 0, H_out + h->n_off, ld_out, h->Hloc, h->n_total, h->n_off);
  SN_LAUNCH_CHECK();
  h->h_valid = train_mode != 0;
  return 0;
}

}  // extern "C" (templates need C++ linkage)
template <int OPT>
static int shard_backward_impl(sn_shard* h, const float* X, const float* dH, int ld_dh, int B, const float* X_next,
                               int B_next, cudaStream_t st) {
  h->step += 1;
  const OptParams op = make_opt(h->hyper, h->step);
  const int K = h->K, N = h->n_rows;
  const float gscale = h->dropout_p > 0.f ? 1.f / (1.f - h->dropout_p) : 1.f;
  BwdPlan p = plan_bwd(N, K, h->sms, false, OPT == 2);
  size_t smem = bwd_smem_bytes(p, false, OPT == 2);
  dim3 grid(p.kgroups, p.nsplit);
  const float* dZ = dH + h->n_off;
  if (X_next && B_next > 0) {
    const int pf_smem = smem + bwd_pf_smem_bytes(p) <= kMaxSmemPerCta && !getenv("SN_NO_PF_SMEM") ? 1 : 0;
    const size_t smem_f = smem + (pf_smem ? bwd_pf_smem_bytes(p) : 0);
    h->pf_parts = pf_smem ? p.kgroups : ceil_div(K, 128);
    if (pf_smem) {
      if (set_smem(bwd_update_kernel<OPT, false, true, true>, smem_f)) return 1;
      launch_k(bwd_update_kernel<OPT, false, true, true>, dim3(grid), dim3(256), smem_f, st, h->W, K, N, K, h->b, h->s1, h->s2, h->bs1,
               h->bs2, dZ, ld_dh, X, K, B, (float*)nullptr, p.kw, p.rows_per_cta, op, X_next, B_next, h->PF, h->l2_keep,
               (const float*)h->Hloc, gscale);
    } else {
      if (set_smem(bwd_update_kernel<OPT, false, true, false>, smem_f)) return 1;
      launch_k(bwd_update_kernel<OPT, false, true, false>, dim3(grid), dim3(256), smem_f, st, h->W, K, N, K, h->b, h->s1, h->s2, h->bs1,
               h->bs2, dZ, ld_dh, X, K, B, (float*)nullptr, p.kw, p.rows_per_cta, op, X_next, B_next, h->PF, h->l2_keep,
               (const float*)h->Hloc, gscale);
    }
    h->pf_valid = true;
  } else {
    if (set_smem(bwd_update_kernel<OPT, false, false>, smem)) return 1;
    launch_k(bwd_update_kernel<OPT, false, false>, dim3(grid), dim3(256), smem, st, h->W, K, N, K, h->b, h->s1, h->s2, h->bs1, h->bs2,
             dZ, ld_dh, X, K, B, (float*)nullptr, p.kw, p.rows_per_cta, op, (const float*)nullptr, 0, (float*)nullptr,
             h->l2_keep, (const float*)h->Hloc, gscale);
    h->pf_valid = false;
  }
  SN_LAUNCH_CHECK();
  mark_written(h->W, (size_t)N * K * 4);
  h->h_valid = false;
  return 0;
}
extern "C" {

int sn_shard_backward_update(sn_shard* h, const float* X, const float* dH, int ld_dh, int B, const float* X_next,
                             int B_next, void* stream) {
  SN_SHARD_ENTER(h);
  SN_REQUIRE(X && dH && B > 0 && B <= kBT && ld_dh >= h->n_off + h->n_rows, "sn_shard_backward_update: bad arguments (B <= 16)");
  SN_REQUIRE(h->h_valid, "sn_shard_backward_update: no matching train-mode sn_shard_forward");
  SN_REQUIRE(!X_next || (B_next > 0 && B_next <= kBT), "sn_shard_backward_update: bad next batch");
  if (h->hyper.type == SN_OPT_ADAM) return shard_backward_impl<2>(h, X, dH, ld_dh, B, X_next, B_next, (cudaStream_t)stream);
  return shard_backward_impl<1>(h, X, dH, ld_dh, B, X_next, B_next, (cudaStream_t)stream);
}

// One pass over the cached batches with the first layer sharded (tp.py's loop, in C so that the
// host enqueues ~14 launches per step instead of crossing the Python/ctypes boundary six times:
// at two GPUs the Python loop ran at 165 us/step, host-bound).
int sn_tp_server_epoch(sn_shard* shard, sn_bob* head, int rank, int world, const float* cache, const int64_t* labels,
                       const int64_t* batch_offsets, int n_batches, const uint8_t* const* drop_masks, uint64_t seed,
                       uint64_t offset, float* H1, float* dH1, void* h_ready_flags, void* dh_ready_flag,
                       float* const* dH1_table_dev, const void* dh_flag_table_dev, uint32_t seq0, float* loss_per_step,
                       void* stream) {
  SN_REQUIRE(shard || (rank == 0 && head && world > 1), "sn_tp_server_epoch: only the head rank of a multi-rank run may come without a shard");
  DeviceGuard device_guard_(shard ? shard->device : head->device);
  if (n_batches == 0) return 0;
  SN_REQUIRE(batch_offsets && n_batches > 0 && H1 && dH1 && world >= 1 && rank >= 0 && rank < world,
             "sn_tp_server_epoch: bad arguments");
  SN_REQUIRE(cache || !shard, "sn_tp_server_epoch: a rank with a shard needs the cached activations");
  SN_REQUIRE((rank == 0) == (head != nullptr), "sn_tp_server_epoch: rank 0, and only rank 0, passes the head");
  SN_REQUIRE(rank != 0 || labels, "sn_tp_server_epoch: the head needs the labels");
  SN_REQUIRE(world == 1 || (h_ready_flags && dh_ready_flag), "sn_tp_server_epoch: flags are required with more than one rank");
  SN_REQUIRE(world == 1 || rank != 0 || (dH1_table_dev && dh_flag_table_dev),
             "sn_tp_server_epoch: the head needs the per-rank gradient destinations and flags");
  SN_REQUIRE(!head || !shard || (head->dims[0] == shard->n_total && head->device == shard->device),
             "sn_tp_server_epoch: the head's input width must be the sharded layer's width");
  cudaStream_t st = (cudaStream_t)stream;
  const int n1 = shard ? shard->n_total : head->dims[0];
  const int K = shard ? shard->K : 0;
  const int nm = head ? head->L - 1 : 0;
  std::vector<const uint8_t*> hm(std::max(nm, 1), nullptr);
  char* flags = reinterpret_cast<char*>(h_ready_flags);
  // column ownership of H1 / dH1: rank 0 owns the first r0 columns (its shard's rows; none without a
  // shard), ranks 1..world-1 equal ranges after that
  const int r0 = (rank == 0) ? (shard ? shard->n_rows : 0) : -1;
  for (int i = 0; i < n_batches; ++i) {
    const int64_t b0 = batch_offsets[i], b1 = batch_offsets[i + 1];
    SN_REQUIRE(b1 > b0 && b1 - b0 <= kBT, "sn_tp_server_epoch: batches of 1..16 rows");
    const int n = (int)(b1 - b0);
    const float* X = shard ? cache + (size_t)b0 * K : nullptr;
    const float* Xn = nullptr;
    int Bn = 0;
    if (shard && i + 1 < n_batches) {
      Xn = cache + (size_t)batch_offsets[i + 1] * K;
      Bn = (int)(batch_offsets[i + 2] - batch_offsets[i + 1]);
      if (Bn <= 0 || Bn > kBT) { Xn = nullptr; Bn = 0; }
    }
    const uint32_t seq = seq0 + (uint32_t)i + 1;
    const uint8_t* m1 = drop_masks ? drop_masks[0] + (size_t)b0 * n1 : nullptr;
    int rc = 0;
    if (shard) {
      rc = sn_shard_forward(shard, X, n, 1, m1, seed, offset + (uint64_t)i, H1, n1, stream);
      if (rc) return rc;
    }
    if (rank != 0) {
      rc = sn_flag_signal(flags + 128 * rank, seq, nullptr, nullptr, 0, stream);
      if (rc) return rc;
      rc = sn_flag_wait(dh_ready_flag, seq, stream);
      if (rc) return rc;
    } else {
      if (world > 1) {
        rc = sn_flag_wait_many(flags + 128, world - 1, 128, seq, stream);
        if (rc) return rc;
      }
      if (drop_masks)
        for (int l = 0; l < nm; ++l) hm[l] = drop_masks[l + 1] + (size_t)b0 * head->dims[l + 1];
      head->pf_valid = false;
      // the head's cut-gradient reduction pushes every rank's column range into that rank's own
      // memory (reading it from here over NVLink made the remote traversals 19 us slower), and
      // it does so BEFORE the head's own first-layer update (bob_backward_impl, after_cut)
      head->cut_table = world > 1 ? dH1_table_dev : nullptr;
      head->cut_first = r0;
      head->cut_cols = world > 1 ? ceil_div(n1 - r0, world - 1) : n1;
      // SN_TP_EARLY_CUT=1: hand the cut gradient out BEFORE the head's first-layer update (a read-only
      // pass, then the update on its own).  Measured (profiles/r2h_tp_trace_2gpu.txt, r2i_tp_trace_8gpu.txt):
      // it takes 6 us off the shards' wait, but the two passes cost 32 us where the fused one costs 21,
      // and from 4 GPUs up rank 0's own stream is the critical path - 8 GPUs: 106 us/step against 95.
      // Off by default.
      static const bool early_cut = getenv("SN_TP_EARLY_CUT") && atoi(getenv("SN_TP_EARLY_CUT")) != 0;
      if (world > 1 && early_cut)
        head->after_cut = [&](cudaStream_t) { return sn_flag_signal_many(dh_flag_table_dev, world - 1, seq, stream); };
      rc = bob_train_step_impl(head, 0, H1, labels + b0 * head->heads, n, drop_masks ? hm.data() : nullptr, seed + 1,
                               offset + (uint64_t)i, dH1, nullptr, loss_per_step ? loss_per_step + i : nullptr, st, nullptr, 0);
      const bool early = (bool)head->after_cut;
      head->after_cut = nullptr;
      head->cut_table = nullptr;
      head->cut_first = -1;
      if (rc) return rc;
      if (world > 1 && !early) {
        rc = sn_flag_signal_many(dh_flag_table_dev, world - 1, seq, stream);
        if (rc) return rc;
      }
    }
    if (shard) {
      rc = sn_shard_backward_update(shard, X, dH1, n1, n, Xn, Bn, stream);
      if (rc) return rc;
    }
  }
  return 0;
}

// ------------------------------------------------------------------ peer-memory mailboxes
int sn_ipc_alloc(size_t bytes, void** dptr, void* handle64) {
  SN_REQUIRE(bytes > 0 && dptr && handle64, "sn_ipc_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle is 64 bytes");
  void* p = nullptr;
  SN_CUDA(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail("sn_ipc_alloc", e, __FILE__, __LINE__);
  }
  memcpy(handle64, &h, 64);
  *dptr = p;
  return 0;
}
int sn_ipc_open(const void* handle64, void** dptr) {
  SN_REQUIRE(handle64 && dptr, "sn_ipc_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  SN_CUDA(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int sn_ipc_close(void* dptr) {
  SN_REQUIRE(dptr, "sn_ipc_close: bad arguments");
  SN_CUDA(cudaIpcCloseMemHandle(dptr));
  return 0;
}
int sn_ipc_free(void* dptr) {
  SN_REQUIRE(dptr, "sn_ipc_free: bad arguments");
  SN_CUDA(cudaFree(dptr));
  return 0;
}

int sn_flag_signal(void* flag, uint32_t value, void* copy_dst, const void* copy_src, size_t copy_bytes, void* stream) {
  SN_REQUIRE(flag && (reinterpret_cast<uintptr_t>(flag) & 3) == 0, "sn_flag_signal: flag must be a 4-byte aligned device address");
  SN_REQUIRE(copy_bytes == 0 || (copy_dst && copy_src && copy_bytes % 8 == 0 && copy_bytes <= 65536 &&
                                 (reinterpret_cast<uintptr_t>(copy_dst) & 7) == 0 && (reinterpret_cast<uintptr_t>(copy_src) & 7) == 0),
             "sn_flag_signal: the optional copy must be 8-byte aligned, a multiple of 8 and at most 64 KiB");
  launch_k(flag_signal_kernel, dim3(1), dim3(128), 0, (cudaStream_t)stream, reinterpret_cast<unsigned*>(flag), (unsigned)value,
           reinterpret_cast<unsigned long long*>(copy_dst), reinterpret_cast<const unsigned long long*>(copy_src),
           (int)(copy_bytes / 8), (unsigned* const*)nullptr, 0);
  SN_LAUNCH_CHECK();
  return 0;
}
int sn_flag_signal_many(const void* flag_table_dev, int count, uint32_t value, void* stream) {
  SN_REQUIRE(flag_table_dev && count >= 1 && count <= 128, "sn_flag_signal_many: a device array of 1..128 flag addresses");
  launch_k(flag_signal_kernel, dim3(1), dim3(128), 0, (cudaStream_t)stream, (unsigned*)nullptr, (unsigned)value,
           (unsigned long long*)nullptr, (const unsigned long long*)nullptr, 0,
           reinterpret_cast<unsigned* const*>(flag_table_dev), count);
  SN_LAUNCH_CHECK();
  return 0;
}
// Error word of the gate kernels: 4 words of pinned host memory, mapped into every device.
static unsigned* g_flag_err = nullptr;
static std::mutex g_flag_err_mu;
static unsigned* flag_err_word() {
  std::lock_guard<std::mutex> lock(g_flag_err_mu);
  if (!g_flag_err) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    memset(p, 0, 64);
    g_flag_err = static_cast<unsigned*>(p);
  }
  return g_flag_err;
}
static unsigned long long flag_timeout_ns() {
  // SN_FLAG_TIMEOUT_MS: how long a gate kernel polls before it gives up (default 20 s; 0 = for ever)
  static const unsigned long long ns = [] {
    const char* e = getenv("SN_FLAG_TIMEOUT_MS");
    const double ms = e ? atof(e) : 20000.0;
    return ms <= 0 ? 0ull : (unsigned long long)(ms * 1e6);
  }();
  return ns;
}
static int flag_check() {
  unsigned* e = g_flag_err;
  if (e && *reinterpret_cast<volatile unsigned*>(e)) {
    const unsigned want = e[1], seen = e[2], idx = e[3];
    return fail_msg("libsplitnn: a flag wait timed out on the GPU (waiting for value " + std::to_string(want) + ", flag " +
                    std::to_string(idx) + " still held " + std::to_string(seen) +
                    "): the peer that should have signalled is gone or stuck; results after that point are invalid");
  }
  return 0;
}
int sn_flag_status(void) { return flag_check(); }
int sn_flag_status_clear(void) {
  if (g_flag_err) memset(g_flag_err, 0, 64);
  return 0;
}

int sn_flag_wait(const void* flag, uint32_t value, void* stream) {
  SN_REQUIRE(flag && (reinterpret_cast<uintptr_t>(flag) & 3) == 0, "sn_flag_wait: flag must be a 4-byte aligned device address");
  if (flag_check()) return 2;
  launch_k(flag_wait_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, reinterpret_cast<const unsigned*>(flag), (unsigned)value, 1, 0,
           flag_timeout_ns(), flag_err_word());
  SN_LAUNCH_CHECK();
  return 0;
}
int sn_flag_wait_many(const void* flags, int count, int stride_bytes, uint32_t value, void* stream) {
  SN_REQUIRE(flags && (reinterpret_cast<uintptr_t>(flags) & 3) == 0 && count >= 1 && count <= 32 && stride_bytes >= 4 &&
                 (stride_bytes & 3) == 0,
             "sn_flag_wait_many: 1..32 flags, 4-byte aligned, stride a multiple of 4 bytes");
  if (flag_check()) return 2;
  launch_k(flag_wait_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, reinterpret_cast<const unsigned*>(flags), (unsigned)value,
           count, stride_bytes / 4, flag_timeout_ns(), flag_err_word());
  SN_LAUNCH_CHECK();
  return 0;
}

typedef CUresult (*StreamValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static StreamValue32Fn stream_value_fn(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
  return reinterpret_cast<StreamValue32Fn>(p);
}
int sn_stream_signal(void* flag, uint32_t value, void* stream) {
  SN_REQUIRE(flag && (reinterpret_cast<uintptr_t>(flag) & 3) == 0, "sn_stream_signal: flag must be a 4-byte aligned device address");
  static StreamValue32Fn fn = stream_value_fn("cuStreamWriteValue32");
  SN_REQUIRE(fn, "sn_stream_signal: cuStreamWriteValue32 is not available from this driver");
  CUresult r = fn((CUstream)stream, (CUdeviceptr)flag, value, CU_STREAM_WRITE_VALUE_DEFAULT);
  if (r != CUDA_SUCCESS) return fail_msg("libsplitnn: cuStreamWriteValue32 failed (" + std::to_string((int)r) + ")");
  return 0;
}
int sn_stream_wait(void* flag, uint32_t value, void* stream) {
  SN_REQUIRE(flag && (reinterpret_cast<uintptr_t>(flag) & 3) == 0, "sn_stream_wait: flag must be a 4-byte aligned device address");
  static StreamValue32Fn fn = stream_value_fn("cuStreamWaitValue32");
  SN_REQUIRE(fn, "sn_stream_wait: cuStreamWaitValue32 is not available from this driver");
  CUresult r = fn((CUstream)stream, (CUdeviceptr)flag, value, CU_STREAM_WAIT_VALUE_GEQ);
  if (r != CUDA_SUCCESS) return fail_msg("libsplitnn: cuStreamWaitValue32 failed (" + std::to_string((int)r) + ")");
  return 0;
}

}  // extern "C"
