// tcgen05 (5th-gen tensor core) GEMM for the tensor-bound regime of the split-NN step: large
// batches (BASELINE.json config 5: B 64..65536, H 256..8192), where the MLP layers stop being a
// weight stream and become real GEMMs.
//
//   C[M,N] = epilogue( sum_r A(m,r) * B(n,r) ),  fp32 operands read as TF32, fp32 accumulate.
//   Each operand is either K-major (reduction index contiguous in memory) or MN-major (its M/N
//   index contiguous); the three products of a training step need all combinations:
//     forward   Y[b,n]  = sum_k X[b,k]  W[n,k]    A = X  K-major,  B = W  K-major,  bias/ReLU/dropout epilogue
//     dX        dA[b,k] = sum_n dZ[b,n] W[n,k]    A = dZ K-major,  B = W  MN-major, ReLU/dropout-gate epilogue
//     dW+update g[n,k]  = sum_b dZ[b,n] A[b,k]    A = dZ MN-major, B = A  MN-major, optimizer epilogue (W, state
//                                                 updated in place; dW is never written to memory)
//   so no operand is ever transposed or copied: nn.Linear's [out,in] weights serve all three.
//
// One 128x128 output tile per CTA (or, on CTA pairs, one 256x256 tile per pair: see CG below),
// warp-specialised:
//   warp 0      TMA producer: cp.async.bulk.tensor 2-D loads of 128x32-float tiles (128-byte rows,
//               SWIZZLE_128B) of A and B into a kStages-deep shared-memory ring, mbarrier complete_tx
//   warp 1      TMEM allocation + single-thread tcgen05.mma.kind::tf32 issue (K=8 per instruction,
//               fp32 accumulation in TMEM), tcgen05.commit releases ring slots
//   warps 2..5  3xTF32: operand warps during the main loop (A rows and their lo twins -> tensor
//               memory, B's lo twin -> shared memory); then the epilogue: tcgen05.ld the accumulator
//               (32 lanes x 32 columns per instruction), through shared memory so every global access
//               is a contiguous row segment, then bias/ReLU/dropout, the ReLU/dropout gate, or the
//               optimizer update streamed through the idle operand ring
// Out-of-range rows/columns/k are zero-filled by TMA and masked in the epilogue, so any M, N and
// any K that is a multiple of 4 work.  Split-K (grid.z) writes raw partial tiles for the host to reduce.
//
// Precision: kind::tf32 reads the fp32 bit patterns and uses the top 19 bits (10-bit mantissa).
// One pass is ~5e-4 relative on the logits of this model - outside the 1e-4 parity
// bound - so one-pass TF32 is the PERF mode (sn_set_matmul_precision(1)); the default is 3xTF32
// (X3 below), which is fp32-accurate.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace sn {
namespace tc {

constexpr int BM = 128, BK = 32;                    // BK floats = 128 bytes = one swizzle row
// accumulator width BN is a template parameter: 128 (one-pass: 3 stages, two CTAs per SM, one's epilogue
// overlaps the other's main loop; 3xTF32: 4 stages, or 2 with two CTAs per SM) or 256 (one-pass, one CTA
// per SM, half the A traffic per flop; and the CTA-pair forms, whose CTAs stage 128 columns each)
__host__ __device__ constexpr int stages_for(int bn, bool x3) { return (x3 || bn == 256) ? 4 : 3; }
constexpr int kTileBytes = BM * BK * 4;             // 16 KB per A tile per stage
constexpr int kThreads = 192;

// X3 = 3xTF32: every operand also has a "lo" twin holding x - trunc_tf32(x); D = lo.hi + hi.lo + hi.hi
// recovers fp32 accuracy (~1e-5 relative: the tensor core's own fp32 accumulation is the floor) at 3x
// the MMAs.  Shared-memory bandwidth decides how that runs: a 128x128x8 TF32 MMA with both operands
// in shared memory reads 8 KB in ~67 clocks, i.e. all of the SM's 128 B/clk by itself; with hi and lo
// tiles of both operands in shared memory (the first version) a k block of 32 cost 96 KB of MMA
// reads + 64 KB of twin traffic = 1250 clocks against 800 clocks of tensor time, and measured 1 us.
// So the M operand goes through TENSOR MEMORY: the four operand warps read each row of the A tile once
// (thread = row), and write it and its lo twin with tcgen05.st; the MMAs then read only B (and its lo
// twin) from shared memory - 48 + 48 KB per k block.  The TMEM operand buffer of stage s needs no
// barrier of its own: the TMA refill of stage s is issued after the MMAs that read it have committed.
template <bool X3, int BN, int STG = stages_for(BN, X3)>
struct __align__(1024) SmemT {
  static constexpr int kStages = STG;
  float a[kStages][BM * BK];
  float b[kStages][BN * BK];
  float b_lo[X3 ? kStages : 1][X3 ? BN * BK : 4];
  unsigned long long full_bar[kStages];
  unsigned long long split_bar[kStages];
  unsigned long long empty_bar[kStages];
  unsigned long long tmem_full_bar;
  unsigned tmem_base;
  // floats of the operand ring (a, b and b_lo are contiguous): the epilogues reuse it as staging space
  static constexpr int kRingFloats = kStages * (BM * BK + BN * BK * (X3 ? 2 : 1));
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, unsigned long long* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor):
//   [0,14) start address >> 4 | [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1)
//   [32,46) stride byte offset >> 4 = 1024 B between 8-row groups | [46,48) version = 1 (sm_100)
//   [61,64) layout type = 2 (SWIZZLE_128B)
__device__ __forceinline__ unsigned long long make_desc_kmajor_sw128(const void* tile) {
  unsigned long long d = 0;
  d |= (unsigned long long)((smem_u32(tile) >> 4) & 0x3FFF);
  d |= (unsigned long long)1 << 16;
  d |= (unsigned long long)(1024 >> 4) << 32;
  d |= (unsigned long long)1 << 46;
  d |= (unsigned long long)2 << 61;
  return d;
}

// MN-major TF32 operands have exactly one legal shared-memory layout (CUTLASS sm100_common.inl:
// "for mn-major tf32 operands, SW128_32B is the only available smem layout"): layout type 1,
// SWIZZLE_128B_BASE32B = Swizzle<2,5,2>, i.e. 128-byte rows along M/N whose 32-byte chunks are
// XOR-ed with (reduction row mod 4); TMA writes it with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
// The tile is stored as 32-float (128 B) wide column blocks, each block one TMA box of 32
// reduction rows x 128 B.  Leading byte offset = distance between column blocks (4096 B), stride
// byte offset = distance between groups of 4 reduction rows (512 B).
__device__ __forceinline__ unsigned long long make_desc_mnmajor_sw128(const void* tile) {
  unsigned long long d = 0;
  d |= (unsigned long long)((smem_u32(tile) >> 4) & 0x3FFF);
  d |= (unsigned long long)(4096 >> 4) << 16;
  d |= (unsigned long long)(512 >> 4) << 32;
  d |= (unsigned long long)1 << 46;
  d |= (unsigned long long)1 << 61;
  return d;
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 [4,6)=1, a/b_format TF32
// [7,10)=[10,13)=2, a/b K-major [15],[16]=0, N>>3 at [17,23), M>>4 at [24,29).
__device__ __forceinline__ unsigned make_idesc_tf32(int m, int n, bool a_mn, bool b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
         ((unsigned)(n >> 3) << 17) | ((unsigned)(m >> 4) << 24);
}

__device__ __forceinline__ void mma_tf32(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc,
                                         unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

enum { EPI_FWD = 0, EPI_GATE = 1, EPI_UPDATE = 2 };

__device__ __forceinline__ void cp16(float* smem, const float* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

#define SN_TMEM_LD32(r, taddr)                                                                                          \
  asm volatile(                                                                                                         \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"  \
      "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"                                                        \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),     \
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),          \
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),         \
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])                       \
      : "r"(taddr))

struct Epilogue {
  // EPI_FWD: C = act(acc + bias) * dropout
  const float* bias;       // [N], nullable
  int relu;
  const uint8_t* mask;     // [M,N] keep bytes (row stride N), nullable
  float keep_prob, scale;  // Philox dropout when mask == nullptr and keep_prob < 1
  uint64_t seed, offset;
  int layer;
  float* out2;             // optional dense copy [M,N]
  // EPI_GATE: C = gate ? (gate[m,n] > 0 ? acc * scale : 0) : acc
  const float* gate;
  int ldgate;
  // EPI_UPDATE: g = acc; optimizer step on C (= the weights), s1, s2 in place
  float* s1;
  float* s2;
  OptParams op;
};

// D[tmem] (+)= A[tmem] * B[smem descriptor]: the M operand comes from tensor memory (lane = row,
// one 32-bit column per k), so the MMA reads no shared memory for it
__device__ __forceinline__ void mma_tf32_ts(unsigned tmem_d, unsigned tmem_a, unsigned long long bdesc, unsigned idesc,
                                            unsigned accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void tmem_st32(unsigned taddr, const unsigned (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
      "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
      "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]),
      "r"(r[31])
      : "memory");
}
// ---- CTA pairs (cta_group::2): two CTAs of a 2-cluster on one TPC run ONE 256-row MMA; each holds its
// 128 rows of A (here: in its tensor memory) and HALF of the B tile in its shared memory, and each
// half is read once for both tensor cores.  Only the pair's rank-0 CTA issues MMAs and commits.
__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(unsigned long long* bar, unsigned rank) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(rank)
      : "memory");
}
__device__ __forceinline__ void mma_tf32_ts_pair(unsigned tmem_d, unsigned tmem_a, unsigned long long bdesc, unsigned idesc,
                                                 unsigned accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void mma_tf32_ss_pair(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc, unsigned idesc,
                                                 unsigned accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
// TMA load whose completion bytes are counted on the barrier at this offset in the pair's rank-0 CTA
// (bit 24 of a shared::cluster address selects the CTA of the pair): the MMA issuer there waits for
// the operands of both CTAs on one barrier.  Source: this GPU's memory only.
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, unsigned long long* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
// completion of the pair's MMAs so far -> one arrival on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void mma_commit_pair(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"((unsigned short)3)
               : "memory");
}

__device__ __forceinline__ float tf32_lo(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

// STG = ring depth.  The 3xTF32 dW + optimizer form runs with STG = 2: half the shared memory and half
// the tensor memory per CTA, so two CTAs share an SM and one's read-modify-write epilogue (most of
// its time: the reduction is only `batch` deep) overlaps the other's main loop.
// CG = 2 (3xTF32 only): CTA pairs, 256 x 256 output tile per pair; BN is the accumulator width (256),
// each CTA stages BN / 2 columns of B.  Launched as clusters of (2,1,1): blockIdx.x = row tile of 128.
template <bool AMN, bool BMN, int EPI, int OPT, bool X3, int BN, int STG, int CG = 1>
__global__ void __launch_bounds__(kThreads, ((X3 && STG > 2) || BN == 256) ? 1 : 2)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, int M, int N,
                 int K, float* __restrict__ C, int ldc, Epilogue ep, int kb_per_split, size_t split_stride) {
  extern __shared__ unsigned char smem_raw[];
  static_assert(CG == 1 || (CG == 2 && BN == 256), "CTA pairs: 256-wide accumulator");
  constexpr int BNS = BN / CG;                      // columns of B staged by this CTA
  using Smem = SmemT<X3, BNS, STG>;
  constexpr int kStages = Smem::kStages;
  // fp32 accumulator: 128 lanes x BN columns; X3: + per stage 32 hi and 32 lo columns of the A operand
  constexpr int kTmemCols = !X3 ? BN : (BN + kStages * 64 <= 256 ? 256 : 512);
  static_assert(!X3 || BN + kStages * 64 <= 512, "A operand buffers do not fit tensor memory");
  // 1 KB alignment by pointer arithmetic on the __shared__ array (an integer round trip would turn every
  // access below into a generic load/store)
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned rank = CG == 2 ? cluster_ctarank() : 0u;
  const int m0 = (CG == 2 ? blockIdx.x : blockIdx.y) * BM, n0 = (CG == 2 ? blockIdx.y : blockIdx.x) * BN;
  const int nb0 = n0 + (int)rank * BNS;             // first B column this CTA stages
  // split-K: grid.z slices of kb_per_split k blocks each, slice z writing its partial tile to
  // C + z * split_stride (the host then launches the reduction as a separate element-wise kernel)
  const int kb0 = (int)blockIdx.z * kb_per_split;
  const int kblocks = min((K + BK - 1) / BK - kb0, kb_per_split);
  C += (size_t)blockIdx.z * split_stride;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sm.full_bar[s], 1);
      mbar_init(&sm.split_bar[s], 4 * CG);          // one arrival per operand warp (of both CTAs of a pair)
      mbar_init(&sm.empty_bar[s], 1);
    }
    mbar_init(&sm.tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // one warp allocates TMEM and later frees it (pairs: the same warp of both CTAs)
    if (CG == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                   "r"(kTmemCols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                   "r"(kTmemCols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned tmem_d = sm.tmem_base;

  pdl_wait();          // A is the previous kernel's output
  pdl_launch();

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % kStages;
        const unsigned ph = (kb / kStages) & 1;
        mbar_wait(&sm.empty_bar[s], ph ^ 1);
        // one-pass pairs: the operands of both CTAs are counted on the rank-0 CTA's barrier (its MMA
        // thread is their only consumer); 3xTF32: each CTA's operand warps consume their own loads
        constexpr bool kLeaderBar = CG == 2 && !X3;
        if (!kLeaderBar) mbar_expect_tx(&sm.full_bar[s], kTileBytes + BNS * BK * 4);
        else if (rank == 0) mbar_expect_tx(&sm.full_bar[s], 2 * (kTileBytes + BNS * BK * 4));
        auto load = [&](float* dst, const CUtensorMap* map, int c0, int c1) {
          if (kLeaderBar) tma_load_2d_pair(dst, map, &sm.full_bar[s], c0, c1);
          else tma_load_2d(dst, map, &sm.full_bar[s], c0, c1);
        };
        if (!AMN) load(sm.a[s], &map_a, (kb0 + kb) * BK, m0);
        else
#pragma unroll
          for (int j = 0; j < BM / 32; ++j) load(sm.a[s] + j * 1024, &map_a, m0 + 32 * j, (kb0 + kb) * BK);
        if (!BMN) load(sm.b[s], &map_b, (kb0 + kb) * BK, nb0);
        else
#pragma unroll
          for (int j = 0; j < BNS / 32; ++j) load(sm.b[s] + j * 1024, &map_b, nb0 + 32 * j, (kb0 + kb) * BK);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      const unsigned idesc = make_idesc_tf32(BM * CG, BN, X3 ? false : AMN, BMN);
      for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % kStages;
        const unsigned ph = (kb / kStages) & 1;
        mbar_wait(X3 ? &sm.split_bar[s] : &sm.full_bar[s], ph);     // X3: the operand warps have done their part
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const unsigned long long bd = BMN ? make_desc_mnmajor_sw128(sm.b[s]) : make_desc_kmajor_sw128(sm.b[s]);
        // one instruction consumes 8 reduction steps: K-major +32 B (2 x 16 B), MN-major +8 rows of
        // 128 B (64 x 16 B); an operand in tensor memory +8 columns
        constexpr unsigned long long astep = AMN ? 64 : 2, bstep = BMN ? 64 : 2;
        if (!X3) {
          const unsigned long long ad = AMN ? make_desc_mnmajor_sw128(sm.a[s]) : make_desc_kmajor_sw128(sm.a[s]);
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            if (CG == 2) mma_tf32_ss_pair(tmem_d, ad + astep * k, bd + bstep * k, idesc, (kb | k) != 0);
            else mma_tf32(tmem_d, ad + astep * k, bd + bstep * k, idesc, (kb | k) != 0);
          }
        } else {
          const unsigned long long bl = BMN ? make_desc_mnmajor_sw128(sm.b_lo[s]) : make_desc_kmajor_sw128(sm.b_lo[s]);
          const unsigned a_hi = tmem_d + (unsigned)(BN + s * 64), a_lo = a_hi + 32;
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            if (CG == 2) {
              mma_tf32_ts_pair(tmem_d, a_lo + 8 * k, bd + bstep * k, idesc, (kb | k) != 0);
              mma_tf32_ts_pair(tmem_d, a_hi + 8 * k, bl + bstep * k, idesc, 1);
              mma_tf32_ts_pair(tmem_d, a_hi + 8 * k, bd + bstep * k, idesc, 1);
            } else {
              mma_tf32_ts(tmem_d, a_lo + 8 * k, bd + bstep * k, idesc, (kb | k) != 0);   // small terms first
              mma_tf32_ts(tmem_d, a_hi + 8 * k, bl + bstep * k, idesc, 1);
              mma_tf32_ts(tmem_d, a_hi + 8 * k, bd + bstep * k, idesc, 1);
            }
          }
        }
        if (CG == 2) mma_commit_pair(&sm.empty_bar[s]);
        else mma_commit(&sm.empty_bar[s]);          // slot free once these MMAs have read it
      }
      if (CG == 2) mma_commit_pair(&sm.tmem_full_bar);
      else mma_commit(&sm.tmem_full_bar);           // accumulator complete
    }
  } else {
    // epilogue warps 2..5: TMEM lane quarter = warp % 4
    const int q = warp & 3;
    if (X3) {
      // operand warps: thread = row of the A tile.  It reads its row's 32 reduction values of the stage
      // (K-major: 8 LDS.128 through the 128-byte swizzle; MN-major: 32 LDS.32, a warp reading one
      // 128-byte line per value) and writes them and their lo twins into the stage's tensor-memory
      // operand buffer; the 128 threads together write B's lo twin (same swizzled offsets as the hi
      // tile, so the layout needs no decoding).
      const int t = threadIdx.x - 64;
      const int r = q * 32 + lane;
      const unsigned lane_addr = (unsigned)(q * 32) << 16;
      for (int kb = 0; kb < kblocks; ++kb) {
        const int s = kb % kStages;
        const unsigned ph = (kb / kStages) & 1;
        mbar_wait(&sm.full_bar[s], ph);
        unsigned hi[32], lo[32];
        if (!AMN) {
          const float4* rowp = reinterpret_cast<const float4*>(sm.a[s]) + r * 8;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const float4 v = rowp[c ^ (r & 7)];
            hi[4 * c + 0] = __float_as_uint(v.x); hi[4 * c + 1] = __float_as_uint(v.y);
            hi[4 * c + 2] = __float_as_uint(v.z); hi[4 * c + 3] = __float_as_uint(v.w);
          }
        } else {
          // column block q (32 rows of the tile), line kk = reduction index, 32-byte chunks XOR-ed with kk mod 4
          const float* blk = sm.a[s] + q * 1024;
#pragma unroll
          for (int kk = 0; kk < 32; ++kk)
            hi[kk] = __float_as_uint(blk[kk * 32 + ((((lane >> 3) ^ (kk & 3)) << 3) | (lane & 7))]);
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) lo[e] = __float_as_uint(tf32_lo(__uint_as_float(hi[e])));
        const float4* bh = reinterpret_cast<const float4*>(sm.b[s]);
        float4* blo = reinterpret_cast<float4*>(sm.b_lo[s]);
#pragma unroll
        for (int i = 0; i < BNS * BK / 4 / 128; ++i) {
          const float4 w = bh[t + 128 * i];
          blo[t + 128 * i] = make_float4(tf32_lo(w.x), tf32_lo(w.y), tf32_lo(w.z), tf32_lo(w.w));
        }
        // the operand buffer of stage s was last read by the MMAs of k block kb - kStages; the TMA
        // load this thread has just seen land was issued after their commit
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const unsigned a_hi = tmem_d + lane_addr + (unsigned)(BN + s * 64);
        tmem_st32(a_hi, hi);
        tmem_st32(a_hi + 32, lo);
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // b_lo: generic-proxy writes -> tensor core reads
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) {                       // the MMA issuer lives in the pair's rank-0 CTA
          if (CG == 2 && rank != 0) mbar_arrive_remote(&sm.split_bar[s], 0);
          else mbar_arrive(&sm.split_bar[s]);
        }
      }
    }
    mbar_wait(&sm.tmem_full_bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if constexpr (EPI == EPI_UPDATE) {
      // dW + optimizer.  The tile's weights and optimizer state are a pure read-modify-write stream
      // (16-24 bytes per element against 2*B flops), so what matters is bytes in flight, and registers
      // cannot hold enough of them (the register-staged form: 16-24 KB per SM, 1.5-1.9 TB/s).  The
      // operand ring is idle once the accumulator is complete, so each warp streams its 32 rows through
      // its quarter of it: chunks of 32 columns x {W, s1, (s2)} arrive by cp.async (rows padded to 36
      // floats: float4 accesses by row are conflict-free), D chunks in flight; the lane that owns the
      // accumulator row updates the chunk in place, then the warp copies it out row-contiguously.
      constexpr int NA = (OPT == 2) ? 3 : 2;
      constexpr int kRowF = 36, kArrF = 32 * kRowF, kChunkF = NA * kArrF;
      constexpr int kRingF = Smem::kRingFloats;
      constexpr int NC = BN / 32;
      constexpr int D = (kRingF / 4 / kChunkF) < NC ? (kRingF / 4 / kChunkF) : NC;
      static_assert(D >= 1, "operand ring too small for one staged chunk");
      float* stage = reinterpret_cast<float*>(&sm.a[0][0]) + q * (kRingF / 4);
      const int rows = min(32, M - (m0 + q * 32));
      const int lrow = lane >> 3, lc4 = (lane & 7) * 4;
      const size_t row0 = (size_t)(m0 + q * 32);
      auto issue = [&](int c) {
        float* dst = stage + (c % D) * kChunkF + lrow * kRowF + lc4;
        const int n = n0 + c * 32 + lc4;
        if (c < NC && n < N) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (i * 4 + lrow < rows) {
              const size_t o = (row0 + i * 4 + lrow) * ldc + n;
              cp16(dst + i * 4 * kRowF, C + o);
              cp16(dst + kArrF + i * 4 * kRowF, ep.s1 + o);
              if (OPT == 2) cp16(dst + 2 * kArrF + i * 4 * kRowF, ep.s2 + o);
            }
          }
        }
        cp_commit();
      };
#pragma unroll
      for (int c = 0; c < D; ++c) issue(c);
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        unsigned r[32];
        const unsigned taddr = tmem_d + ((unsigned)(q * 32) << 16) + (unsigned)(c * 32);
        SN_TMEM_LD32(r, taddr);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        cp_wait<D - 1>();
        __syncwarp();
        float* slot = stage + (c % D) * kChunkF;
        if (lane < rows) {
          float* pw = slot + lane * kRowF;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (n0 + c * 32 + j * 4 < N) {
              float4 p = *reinterpret_cast<float4*>(pw + j * 4);
              float4 a = *reinterpret_cast<float4*>(pw + kArrF + j * 4);
              float4 b = (OPT == 2) ? *reinterpret_cast<float4*>(pw + 2 * kArrF + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
              opt_update<OPT>(ep.op, p.x, __uint_as_float(r[j * 4 + 0]), a.x, b.x);
              opt_update<OPT>(ep.op, p.y, __uint_as_float(r[j * 4 + 1]), a.y, b.y);
              opt_update<OPT>(ep.op, p.z, __uint_as_float(r[j * 4 + 2]), a.z, b.z);
              opt_update<OPT>(ep.op, p.w, __uint_as_float(r[j * 4 + 3]), a.w, b.w);
              *reinterpret_cast<float4*>(pw + j * 4) = p;
              *reinterpret_cast<float4*>(pw + kArrF + j * 4) = a;
              if (OPT == 2) *reinterpret_cast<float4*>(pw + 2 * kArrF + j * 4) = b;
            }
          }
        }
        __syncwarp();
        const int n = n0 + c * 32 + lc4;
        if (n < N) {
          const float* src = slot + lrow * kRowF + lc4;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (i * 4 + lrow < rows) {
              const size_t o = (row0 + i * 4 + lrow) * ldc + n;
              *reinterpret_cast<float4*>(C + o) = *reinterpret_cast<const float4*>(src + i * 4 * kRowF);
              *reinterpret_cast<float4*>(ep.s1 + o) = *reinterpret_cast<const float4*>(src + kArrF + i * 4 * kRowF);
              if (OPT == 2) *reinterpret_cast<float4*>(ep.s2 + o) = *reinterpret_cast<const float4*>(src + 2 * kArrF + i * 4 * kRowF);
            }
          }
        }
        __syncwarp();
        issue(c + D);                          // refill the slot just drained (an empty group past the last chunk)
      }
      cp_wait<0>();
    } else {
    // The operand ring is idle once the accumulator is complete: reuse it as a per-warp 32x33
    // transpose tile, so that global accesses are row-contiguous (lane = column) instead of one
    // row per lane.  What the element-wise step READS (the ReLU gate of dX, the keep bytes of a forward
    // under teacher-forced masks) is requested for all 32 rows of a chunk at once, before the chunk's
    // accumulator is fetched: row-at-a-time it was eight dependent round trips per chunk with nothing
    // to hide behind (one CTA per SM).
    float* tile = reinterpret_cast<float*>(&sm.a[0][0]) + q * (32 * 33);
    constexpr int NC = BN / 32;
    const bool has_side = (EPI == EPI_GATE) ? ep.gate != nullptr : ep.mask != nullptr;
    auto load_side = [&](int c, float (&g)[32]) {
      const int n = n0 + c * 32 + lane;
      if (!has_side || c >= NC || n >= N) return;
#pragma unroll
      for (int rr = 0; rr < 32; ++rr) {
        const int row = m0 + q * 32 + rr;
        if (row < M) {
          if (EPI == EPI_GATE) g[rr] = ep.gate[(size_t)row * ep.ldgate + n];
          else g[rr] = (float)ep.mask[(size_t)row * N + n];
        }
      }
    };
    // one chunk: accumulator columns [32c, 32c+32) -> transpose tile -> element-wise step, row-contiguous stores
    auto chunk = [&](int c, const float (&g)[32]) {
      unsigned r[32];
      const unsigned taddr = tmem_d + ((unsigned)(q * 32) << 16) + (unsigned)(c * 32);
      SN_TMEM_LD32(r, taddr);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 32; ++j) tile[lane * 33 + j] = __uint_as_float(r[j]);     // lane holds row q*32+lane
      __syncwarp();
      const int n = n0 + c * 32 + lane;                                            // now lane = column
      if (n < N) {
        const float bias_v = (EPI == EPI_FWD && ep.bias) ? ep.bias[n] : 0.f;
        if (has_side) {
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) {
            const int row = m0 + q * 32 + rr;
            if (row < M) {
              float v = tile[rr * 33 + lane];
              if (EPI == EPI_FWD) {
                v += bias_v;
                if (ep.relu) v = fmaxf(v, 0.f);
                v = g[rr] != 0.f ? v * ep.scale : 0.f;
                C[(size_t)row * ldc + n] = v;
                if (ep.out2) ep.out2[(size_t)row * N + n] = v;
              } else {
                C[(size_t)row * ldc + n] = g[rr] > 0.f ? v * ep.scale : 0.f;
              }
            }
          }
        } else if (EPI == EPI_FWD && ep.keep_prob < 1.f && (N & 3) == 0) {
          // handled below: every lane takes part (the Philox words are shared inside quads of lanes)
        } else {
#pragma unroll 4
          for (int rr = 0; rr < 32; ++rr) {
            const int row = m0 + q * 32 + rr;
            if (row >= M) break;
            float v = tile[rr * 33 + lane];
            if (EPI == EPI_FWD) {
              v += bias_v;
              if (ep.relu) v = fmaxf(v, 0.f);
              const size_t idx = (size_t)row * N + n;
              if (ep.keep_prob < 1.f) v = philox_keep(ep.seed, ep.offset, ep.layer, idx, ep.keep_prob) ? v * ep.scale : 0.f;
              C[(size_t)row * ldc + n] = v;
              if (ep.out2) ep.out2[idx] = v;
            } else {
              C[(size_t)row * ldc + n] = v;
            }
          }
        }
      }
      if (EPI == EPI_FWD && !has_side && ep.keep_prob < 1.f && (N & 3) == 0) {
        // Philox dropout: one Philox4x32-10 decides four consecutive elements of a row, i.e. the four
        // columns of a quad of lanes.  Lane j of a quad evaluates it for row 4t + j, and four xor-shuffles
        // hand every lane the word of its own column for each of the four rows (per element the Philox
        // cost the epilogue 27 us per 256 x 256 tile against 47 us of main loop in one-pass TF32).
        const float bias_v = (ep.bias && n < N) ? ep.bias[n] : 0.f;
        const int j = lane & 3;
        uint4 w[8];                                                  // eight independent Philox chains in flight
#pragma unroll
        for (int t = 0; t < 8; ++t)
          w[t] = philox_words(ep.seed, ep.offset, ep.layer, ((size_t)(m0 + q * 32 + 4 * t + j) * N + (n & ~3)) >> 2);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
#pragma unroll
          for (int sft = 0; sft < 4; ++sft) {
            const int k = j ^ sft;                                   // word this lane gives away = column of its partner
            const uint32_t lo = (k & 1) ? w[t].y : w[t].x, hi = (k & 1) ? w[t].w : w[t].z;
            const uint32_t got = __shfl_xor_sync(0xffffffffu, (k & 2) ? hi : lo, sft);   // partner's word for MY column, row 4t + k
            const int rr = 4 * t + k, row = m0 + q * 32 + rr;
            float v = tile[rr * 33 + lane] + bias_v;
            if (ep.relu) v = fmaxf(v, 0.f);
            v = philox_word_keeps(got, ep.keep_prob) ? v * ep.scale : 0.f;
            if (n < N && row < M) {
              C[(size_t)row * ldc + n] = v;
              if (ep.out2) ep.out2[(size_t)row * N + n] = v;
            }
          }
        }
      }
      __syncwarp();
    };
    float side[32];
#pragma unroll 1
    for (int c = 0; c < NC; ++c) {
      load_side(c, side);          // in flight while the accumulator chunk is fetched and transposed
      chunk(c, side);
    }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(kTmemCols) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(kTmemCols) : "memory");
  }
}

// Bias gradient of a large batch in two fixed-order stages (the one-thread-per-column loop over all rows
// costs 3.7 ms at 65536 rows when the layer is 256 wide: one CTA): part[c][n] = sum of rows
// [c*rpc, (c+1)*rpc) of column n - block (32 columns, 8 row lanes), lanes combined in order - then one
// thread per column adds the chunks in order and applies the optimizer.
__global__ void bias_grad_partial_kernel(const float* __restrict__ dZ, int ldz, int B, int N, int rpc, float* __restrict__ part) {
  pdl_wait();
  pdl_launch();
  __shared__ float red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + tx, r0 = blockIdx.y * rpc, r1 = min(B, r0 + rpc);
  float acc = 0.f;
  if (n < N)
    for (int b = r0 + ty; b < r1; b += 8) acc += __ldg(dZ + (size_t)b * ldz + n);
  red[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && n < N) {
    float s = red[0][tx];
#pragma unroll
    for (int j = 1; j < 8; ++j) s += red[j][tx];
    part[(size_t)blockIdx.y * N + n] = s;
  }
}
template <int OPT>
__global__ void bias_opt_from_partials_kernel(const float* __restrict__ part, int chunks, int N, float* __restrict__ bias,
                                              float* __restrict__ s1, float* __restrict__ s2, OptParams op) {
  pdl_wait();
  pdl_launch();
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float db = 0.f;
  for (int c = 0; c < chunks; ++c) db += __ldcg(part + (size_t)c * N + n);
  float pv = bias[n], a = s1[n], q = (OPT == 2) ? s2[n] : 0.f;
  opt_update<OPT>(op, pv, db, a, q);
  bias[n] = pv;
  s1[n] = a;
  if (OPT == 2) s2[n] = q;
}

// W, state <- optimizer(W, sum_s P[s]) - the reduction behind a split-K dW GEMM (n4 = elements / 4; slices
// are dense slabs `slab` floats apart; fixed add order)
template <int OPT>
__global__ void splitk_opt_kernel(float* __restrict__ W, const float* __restrict__ P, int nsplit, size_t slab, float* __restrict__ s1,
                                  float* __restrict__ s2, size_t n4, OptParams op) {
  pdl_wait();
  pdl_launch();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 g = __ldcg(reinterpret_cast<const float4*>(P) + i);
    for (int s = 1; s < nsplit; ++s) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(P + s * slab) + i);
      g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
    }
    float4 w = reinterpret_cast<float4*>(W)[i], a = reinterpret_cast<float4*>(s1)[i];
    float4 b = (OPT == 2) ? reinterpret_cast<float4*>(s2)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    opt_update<OPT>(op, w.x, g.x, a.x, b.x);
    opt_update<OPT>(op, w.y, g.y, a.y, b.y);
    opt_update<OPT>(op, w.z, g.z, a.z, b.z);
    opt_update<OPT>(op, w.w, g.w, a.w, b.w);
    reinterpret_cast<float4*>(W)[i] = w;
    reinterpret_cast<float4*>(s1)[i] = a;
    if (OPT == 2) reinterpret_cast<float4*>(s2)[i] = b;
  }
}

// =====================================================================================
// Skinny forward on tensor cores: P[q][b][n] = partial sums of X[b][:] . W[n][:] for B <= 16.
//
// At batch 16 the first layer is a 108 MB weight stream with 16 FMAs per weight: on the CUDA
// cores that is FP32-pipe-bound (27 us against a 17 us HBM floor).  Here the weights are
// the M operand: D[128 weight rows x 16 batch] += W_tile[128 x 8] * X^T[8 x 16], kind::tf32,
// M=128 N=16 K=8 per instruction, as 3xTF32 (lo.hi + hi.lo + hi.hi) so the result keeps fp32
// accuracy.  The SIMT work per weight drops from 16 FMAs to one AND and one SUB (the lo twin),
// and the kernel becomes the HBM stream it should be.
//
// Work unit = (row tile of 128 weight rows, k block of 32); units are ordered row-tile-major and cut
// into equal contiguous ranges, one per CTA (grid = #SMs, whatever N and K are).  A range touches
// at most two row tiles: two sets of accumulators in TMEM.  A row tile's CTAs are consecutive;
// CTA c writes its partial into slot c - c_first(tile), and the last CTA of a tile zero-fills the
// slots nobody wrote, so fwd_epilogue_kernel adds a fixed `qmax` slots per element.
//   warp 0      TMA producer (8-stage ring of 16 KB weight tiles + 2 KB activation blocks; the
//               weight loads of the first stages are issued BEFORE the dependency wait when the
//               host allows, and rows below l2_keep_w carry the evict_last L2 policy)
//   warp 1      TMEM alloc + MMA issue (M operand from tensor memory, N operand from shared memory)
//   warps 2..5  per stage: weight rows -> hi and lo operands in tensor memory (tcgen05.st), lo twin
//               of the activation block; at the end accumulators -> partial stores (lane = weight
//               row, so each of the 16 stores of a warp is one contiguous 128-byte segment)
// =====================================================================================
constexpr int kSkStages = 8;
constexpr int kSkNB = 16;                           // batch columns of one accumulator
constexpr int kSkAccs = 12;                         // independent accumulators per segment
constexpr int kSkAccCols = 2 * kSkAccs * kSkNB;     // 384 TMEM columns of accumulators (two segments)
constexpr int kSkTmemCols = 512;                    // + 2 x (32 hi + 32 lo) columns of weight operand

struct __align__(1024) SkSmem {
  float a[kSkStages][BM * BK];                      // weight tiles as TMA delivers them (SWIZZLE_128B)
  float b[kSkStages][kSkNB * BK];                   // activation blocks, hi (as loaded) ...
  float b_lo[kSkStages][kSkNB * BK];                // ... and lo twins
  unsigned long long full_bar[kSkStages];           // TMA landed
  unsigned long long split_bar[kSkStages];          // operands of the stage are ready for the MMAs
  unsigned long long empty_bar[kSkStages];          // MMAs of the stage are done (smem slot free)
  unsigned long long a_empty_bar[2];                // MMAs reading TMEM operand buffer j are done
  unsigned long long tmem_full_bar[2];
  unsigned tmem_base;
};
constexpr size_t sk_smem_bytes() { return sizeof(SkSmem) + 1024; }

__device__ __forceinline__ void tma_load_2d_hint(void* dst, const CUtensorMap* map, unsigned long long* bar, int c0, int c1,
                                                 unsigned long long policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}

__device__ __forceinline__ void tmem_ld16(unsigned taddr, unsigned (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}

// first / last CTA whose unit range [floor(c*U/G), floor((c+1)*U/G)) contains unit x
__host__ __device__ inline int sk_cta_of_unit(long long x, long long U, long long G) { return (int)(((x + 1) * G - 1) / U); }

__global__ void __launch_bounds__(kThreads, 1)
skinny_fwd_x3_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x, int N, int K,
                     int B, float* __restrict__ P, int Bpad, int qmax, int prefetch_early, int l2_keep_w) {
  extern __shared__ unsigned char smem_raw[];
  SkSmem& sm = *reinterpret_cast<SkSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long KB = (K + BK - 1) / BK, RT = (N + BM - 1) / BM, U = KB * RT, G = gridDim.x;
  const long long u0 = (long long)blockIdx.x * U / G, u1 = (long long)(blockIdx.x + 1) * U / G;
  const int nunits = (int)(u1 - u0);
  const int rt_first = (int)(u0 / KB);
  constexpr unsigned kStageBytes = (BM * BK + kSkNB * BK) * 4;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    for (int s = 0; s < kSkStages; ++s) {
      mbar_init(&sm.full_bar[s], 1);
      mbar_init(&sm.split_bar[s], 4);              // one arrival per splitter warp
      mbar_init(&sm.empty_bar[s], 1);
    }
    for (int j = 0; j < 2; ++j) {
      mbar_init(&sm.a_empty_bar[j], 1);
      mbar_init(&sm.tmem_full_bar[j], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                 "r"(kSkTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned tmem_d = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      unsigned long long pol_keep, pol_norm;
      asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
      asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol_norm));
      auto load_w = [&](int i) {
        const long long u = u0 + i;
        const int rt = (int)(u / KB), kb = (int)(u - (long long)rt * KB), s = i % kSkStages;
        tma_load_2d_hint(sm.a[s], &map_w, &sm.full_bar[s], kb * BK, rt * BM, rt * BM < l2_keep_w ? pol_keep : pol_norm);
      };
      auto load_x = [&](int i) {
        const long long u = u0 + i;
        const int kb = (int)(u % KB), s = i % kSkStages;
        tma_load_2d(sm.b[s], &map_x, &sm.full_bar[s], kb * BK, 0);
      };
      // the weights were last written at least two launches back (the host says so): start
      // their stream before the dependency wait; the activations are the predecessor's output
      const int npre = prefetch_early ? min(kSkStages, nunits) : 0;
      for (int i = 0; i < npre; ++i) {
        mbar_expect_tx(&sm.full_bar[i % kSkStages], kStageBytes);
        load_w(i);
      }
      pdl_wait();
      pdl_launch();
      for (int i = 0; i < npre; ++i) load_x(i);
      for (int i = npre; i < nunits; ++i) {
        const int s = i % kSkStages;
        const unsigned ph = (i / kSkStages) & 1;
        mbar_wait(&sm.empty_bar[s], ph ^ 1);
        mbar_expect_tx(&sm.full_bar[s], kStageBytes);
        load_w(i);
        load_x(i);
      }
    } else {
      pdl_wait();
      pdl_launch();
    }
  } else if (warp == 1) {
    pdl_wait();
    pdl_launch();
    if (lane == 0) {
      const unsigned idesc = make_idesc_tf32(BM, kSkNB, false, false);
      for (int i = 0; i < nunits; ++i) {
        const int s = i % kSkStages, j = i & 1;
        const unsigned ph = (i / kSkStages) & 1;
        const long long u = u0 + i;
        const int rt = (int)(u / KB), kb = (int)(u - (long long)rt * KB);
        const int seg = rt != rt_first;
        const bool seg_start = (i == 0) || (kb == 0);
        const bool seg_end = (i == nunits - 1) || (kb == KB - 1);
        mbar_wait(&sm.split_bar[s], ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // Twelve independent accumulators per segment (3 products x 4 k-steps of the block), 16
        // TMEM columns each: an N=16 MMA is far shorter than the pipeline's accumulate latency, so
        // one accumulator would serialise the issue stream; and each chain adds once per k block
        // instead of twelve times, which cuts the accumulator's truncation bias by an order of
        // magnitude.  The epilogue adds the twelve in fp32.
        const unsigned d = tmem_d + (unsigned)(seg * kSkAccs * kSkNB);
        const unsigned a_hi = tmem_d + (unsigned)(kSkAccCols + j * 64), a_lo = a_hi + 32;
        const unsigned long long bd = make_desc_kmajor_sw128(sm.b[s]), bl = make_desc_kmajor_sw128(sm.b_lo[s]);
        const unsigned acc = seg_start ? 0u : 1u;
#pragma unroll
        for (int k = 0; k < BK / 8; ++k) {
          mma_tf32_ts(d + (unsigned)((0 * 4 + k) * kSkNB), a_lo + 8 * k, bd + 2 * k, idesc, acc);
          mma_tf32_ts(d + (unsigned)((1 * 4 + k) * kSkNB), a_hi + 8 * k, bl + 2 * k, idesc, acc);
          mma_tf32_ts(d + (unsigned)((2 * 4 + k) * kSkNB), a_hi + 8 * k, bd + 2 * k, idesc, acc);
        }
        mma_commit(&sm.empty_bar[s]);
        mma_commit(&sm.a_empty_bar[j]);
        if (seg_end) mma_commit(&sm.tmem_full_bar[seg]);
      }
    }
  } else {
    // Operand warps 2..5.  Thread = weight row (TMEM lane quarter = warp % 4): it reads its row's
    // 32 floats of the stage from the swizzled tile (8 conflict-free LDS.128), and writes them and
    // their lo twins straight into tensor memory, from where the MMAs take the M operand.  The
    // first version kept hi and lo tiles in shared memory: 6.5 bytes of shared-memory traffic per
    // weight byte (TMA write, twin read + write, three MMA reads) - more than the SMs have at HBM
    // rate - and ran slower than the CUDA-core kernel.  Now: TMA write + one read.
    pdl_wait();
    pdl_launch();
    const int q = warp & 3;
    const int t = threadIdx.x - 64;
    const int r = q * 32 + lane;
    const unsigned lane_addr = (unsigned)(q * 32) << 16;
    for (int i = 0; i < nunits; ++i) {
      const int s = i % kSkStages, j = i & 1;
      const unsigned ph = (i / kSkStages) & 1;
      mbar_wait(&sm.full_bar[s], ph);
      const float4* rowp = reinterpret_cast<const float4*>(sm.a[s]) + r * 8;
      unsigned hi[32], lo[32];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float4 v = rowp[c ^ (r & 7)];
        hi[4 * c + 0] = __float_as_uint(v.x); hi[4 * c + 1] = __float_as_uint(v.y);
        hi[4 * c + 2] = __float_as_uint(v.z); hi[4 * c + 3] = __float_as_uint(v.w);
      }
      const float4 xb = reinterpret_cast<const float4*>(sm.b[s])[t];
#pragma unroll
      for (int e = 0; e < 32; ++e) lo[e] = __float_as_uint(tf32_lo(__uint_as_float(hi[e])));
      reinterpret_cast<float4*>(sm.b_lo[s])[t] = make_float4(tf32_lo(xb.x), tf32_lo(xb.y), tf32_lo(xb.z), tf32_lo(xb.w));
      // operand buffer j was last read by the MMAs of unit i-2
      if (i >= 2) mbar_wait(&sm.a_empty_bar[j], (unsigned)(((i >> 1) - 1) & 1));
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const unsigned a_hi = tmem_d + lane_addr + (unsigned)(kSkAccCols + j * 64);
      tmem_st32(a_hi, hi);
      tmem_st32(a_hi + 32, lo);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // b_lo: generic-proxy write -> tensor core read
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.split_bar[s]);
    }
    // accumulators -> partial slots: sum of the twelve, small products first (lo.hi, hi.lo, then
    // hi.hi), fixed order
    const int nseg = nunits == 0 ? 0 : ((int)((u1 - 1) / KB) != rt_first ? 2 : 1);
    for (int seg = 0; seg < nseg; ++seg) {
      const int rt = rt_first + seg;
      mbar_wait(&sm.tmem_full_bar[seg], 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      float acc[kSkNB];
#pragma unroll
      for (int b = 0; b < kSkNB; ++b) acc[b] = 0.f;
#pragma unroll
      for (int a0 = 0; a0 < kSkAccs; a0 += 4) {
        unsigned v[4][kSkNB];
#pragma unroll
        for (int a = 0; a < 4; ++a) tmem_ld16(tmem_d + lane_addr + (unsigned)((seg * kSkAccs + a0 + a) * kSkNB), v[a]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < kSkNB; ++b) acc[b] += __uint_as_float(v[a][b]);
      }
      const int c_first = sk_cta_of_unit((long long)rt * KB, U, G);
      const int c_last = sk_cta_of_unit((long long)(rt + 1) * KB - 1, U, G);
      const int slot = (int)blockIdx.x - c_first;
      const int n = rt * BM + r;
      if (n < N) {
        float* dst = P + (size_t)slot * Bpad * N + n;
#pragma unroll
        for (int b = 0; b < kSkNB; ++b)
          if (b < B) dst[(size_t)b * N] = acc[b];
        if ((int)blockIdx.x == c_last) {
          for (int z = slot + 1; z < qmax; ++z) {
            float* zd = P + (size_t)z * Bpad * N + n;
#pragma unroll
            for (int b = 0; b < kSkNB; ++b)
              if (b < B) zd[(size_t)b * N] = 0.f;
          }
        }
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(kSkTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// [rows, cols] fp32 row-major with leading dimension ld (floats); box = 32 floats x box_rows, 128B swizzle.
// K-major operand: rows = its M/N extent, cols = reduction, box_rows = 128.
// MN-major operand: rows = reduction, cols = its M/N extent, box_rows = 32 (four boxes per tile).
inline int make_map(CUtensorMap* map, const float* base, int rows, int cols, int ld, int box_rows, bool mn_major = false) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return fail_msg("libsplitnn: cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail_msg("libsplitnn: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  return 0;
}

inline bool usable(const float* A, int lda, const float* B, int ldb, int K) {
  return (K % 4) == 0 && (lda % 4) == 0 && (ldb % 4) == 0 && ((uintptr_t)A % 16) == 0 && ((uintptr_t)B % 16) == 0;
}

template <bool X3, int BN, int STG = stages_for(BN, X3)>
constexpr size_t smem_bytes() { return sizeof(SmemT<X3, BN, STG>) + 1024; }

}  // namespace tc
}  // namespace sn
