// Bob's server MLP at the reference's batch sizes (B <= 16 per tile): every kernel here is one
// streaming traversal of a weight matrix and is bounded by HBM bandwidth, not by math
// ("roofline").  fp32 CUDA-core FMA gives the reference's fp32 results without the
// 3xTF32 operand split a tensor-core path would need, and at 16 rows the FMA pipe keeps up with
// HBM.  The large-batch (tensor-bound) regime is served by the tcgen05 kernels in gemm_tc.cuh.
// Reference: models.py:33-82 (model2 / model2_sisa / model2_sisa_concat), autograd's Linear
// backward, torch optim/sgd.py and optim/adam.py single-tensor paths.
#pragma once
#include "common.cuh"

namespace sn {

// How a kernel body meets the work it depends on.  As its own launch: programmatic dependent launch
// (wait for the previous kernel, then let the next one start its prologue).  As one phase of the
// persistent tail kernel (tail_kernel.cuh): a grid-wide barrier.
struct PdlSync {
  static constexpr bool kCoherent = false;
  __device__ __forceinline__ void operator()() const {
    pdl_wait();
    pdl_launch();
  }
};
// Load of a value that an EARLIER PHASE may have produced.  Across kernel launches the read-only path
// (ld.global.nc) is fine: the producer is another kernel.  Inside the persistent tail kernel the
// producer is another CTA of the same launch, the non-coherent path and L1 are off limits, and the
// load goes to L2 (ld.global.cg).
template <class Sync, class T>
__device__ __forceinline__ T ld_dep(const T* p) {
  if constexpr (Sync::kCoherent) return __ldcg(p);
  else return __ldg(p);
}

// ------------------------------------------------------------------------------------------
// Forward, skinny: P[ks][b][n] = sum_{k in chunk ks} X[b][k] * W[n][k]
//   grid (nx, ksplit, batch tiles), 256 threads.  The CTA stages its X chunk [16][kc] in
//   shared memory once; each warp owns a contiguous range of W rows and streams them with
//   coalesced 128-bit loads (lanes stride k), R rows at a time so each X read from smem feeds
//   R*4 FMAs; per-lane partial sums are transposed-and-reduced with shuffles at the end of the
//   chunk.  Split-K partials are summed in a fixed order by fwd_epilogue_kernel: deterministic.
// ------------------------------------------------------------------------------------------
// cp.async helpers: 16-byte global->shared copies that bypass L1 (.cg) and complete asynchronously
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// L2 eviction-priority policies (createpolicy) and the hinted forms of the copies/stores that use
// them.  The first layer's weights (108 MB) are touched twice per step and fit in the 126 MB L2
// if nothing pushes them out: their lines are tagged evict_last, the optimizer state (touched once
// per step, never re-read before its lines would have been evicted anyway) evict_first.
__device__ __forceinline__ uint64_t l2_policy(int kind) {   // 0 normal, 1 evict_last, 2 evict_first
  uint64_t p;
  if (kind == 1) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  else if (kind == 2) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void cp_async16_hint(void* smem, const void* gmem, uint64_t policy) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "l"(policy) : "memory");
}
__device__ __forceinline__ void st_hint(float4* p, const float4& v, uint64_t policy) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;\n" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(policy) : "memory");
}

template <int V>
__device__ __forceinline__ void butterfly_reduce(float* v, int lane) {
  int cnt = V;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    if (cnt > 1) {
      const int half = cnt >> 1;
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < V / 2; ++i) {
        if (i < half) {
          const float send = up ? v[i] : v[i + half];
          const float keep = up ? v[i + half] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      cnt = half;
    } else {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
    }
  }
}

#ifdef SN_TIMING
__device__ long long g_fwd_timing[148 * 8 * 8];
#define SN_T(i) do { if (lane == 0) tstamp[i] = clock64(); } while (0)
#else
#define SN_T(i) do {} while (0)
#endif

constexpr int kFwdRing = 4;      // W tiles (8 rows x 128 k = 4 KB) in flight per warp
constexpr int kFwdSlots = 2;     // a row group's k-range is covered by at most two warps

__device__ __forceinline__ void cp_async16_zfill(void* smem, const void* gmem, bool valid) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  const int n = valid ? 16 : 0;                                  // src-size 0 => 16 zero bytes
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(n) : "memory");
}

// grid (nx, ksplit, batch tiles), 256 threads.  Work unit: a TILE = 8 rows x 128 k of W.  The
// tiles of a k-chunk, ordered (row group, k step), are cut into equal contiguous ranges, one per
// warp, so every warp streams the same number of bytes whatever N is; a warp's cp.async ring
// runs continuously across row-group boundaries.  A row group whose k steps straddle two warps
// gets two partial sums (slot 0 from the warp that holds its first k step, slot 1 from the
// other); complete groups write zeros to slot 1.  fwd_epilogue_kernel adds the 2*ksplit partials
// in a fixed order.
//
// Math: packed FFMA2 (fma.rn.f32x2, sm_100): one instruction does acc[b],acc[b+1] += w * x[b],x[b+1]
// with w as the scalar-broadcast operand.  Scalar FFMA made this kernel issue-bound (ncu:
// 68 % issue-active, "selected" the top stall, at 45 % of HBM); FFMA2 halves the FMA issue slots.
// For the batch pairs to be register pairs the X chunk is staged TRANSPOSED: row = k, 16 batch
// values per row (64 B).  Lane l of a tile owns k = 4l..4l+3 (one float4 of each W row); its four
// X rows are stored at physical rows kk*32 + l and the 16-byte batch quads are XOR-swizzled with
// (row>>1)&3, which makes every LDS.128 of a quarter-warp hit 8 distinct bank groups.
// R = weight rows per tile (8 or 4), WPC = 64 / R warps per CTA: the same 128 KB of W rings and the
// same FMAs per SM either way; R = 4 halves the accumulators per thread (64 instead of 128
// registers), which lets 16 warps instead of 8 be resident to hide the FMA / LDS latencies.
template <int R, class Sync>
__device__ __forceinline__ void
fwd_skinny_body(float4* smem4, const dim3 bid, const int gdx, Sync sync,
                const float* __restrict__ X, int ldx, int B, const float* __restrict__ W, int ldw, int N,
                int K, int kc, float* __restrict__ P, int Bpad, int ksplit, int prefetch_early, int l2_keep_w) {
  // rows below l2_keep_w are tagged evict_last (see l2_policy), the others stream
  const uint64_t pol_keep = l2_policy(1);
  constexpr int WPC = 64 / R, NT = 32 * WPC;        // warps, threads per CTA
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ks = bid.y, k0 = ks * kc;
  const int klen = min(kc, K - k0);
  const int iters = (klen + 127) >> 7;
  const int b0 = bid.z * kBT, nb = min(kBT, B - b0);
  const int kc4 = kc >> 2;
  float4* wring = smem4 + kBT * kc4 + warp * (kFwdRing * R * 32);   // after the X chunk
#ifdef SN_TIMING
  long long tstamp[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long t_red = 0;
#endif
  SN_T(0);

  const int groups = (N + R - 1) / R;
  const long long T = (long long)groups * iters;
  // a warp's range must be at least one row group long, or a group could be cut into three or more
  // pieces and two of them would share partial slot 1: with fewer groups than warps (a narrow layer)
  // only the first `groups` warps take work
  const int gw = bid.x * WPC + warp, nw = min(gdx * WPC, groups);
  const int t0 = gw < nw ? (int)(T * gw / nw) : (int)T, t1 = gw < nw ? (int)(T * (gw + 1) / nw) : (int)T;

  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool ktail = (klen & 127) != 0;      // only then can a lane sit beyond the end of K

  // W stream: (gi, iti) = coordinates of the next tile to issue
  int gi = t0 / iters, iti = t0 - gi * iters, ti = t0;
  auto issue = [&](int slot) {
    if (ti < t1) {
      const int k = iti * 128 + lane * 4;
      if (k < klen) {
        const float* src = W + (size_t)(gi * R) * ldw + k0 + k;
        if (gi * R + R <= N) {
          if (gi * R < l2_keep_w) {
#pragma unroll
            for (int i = 0; i < R; ++i) cp_async16_hint(wring + (slot * R + i) * 32 + lane, src + (size_t)i * ldw, pol_keep);
          } else {
#pragma unroll
            for (int i = 0; i < R; ++i) cp_async16(wring + (slot * R + i) * 32 + lane, src + (size_t)i * ldw);
          }
        } else {
#pragma unroll
          for (int i = 0; i < R; ++i)
            cp_async16(wring + (slot * R + i) * 32 + lane, src + (size_t)min(i, N - 1 - gi * R) * ldw);
        }
      }
      ++ti;
      if (++iti == iters) { iti = 0; ++gi; }
    }
    cp_async_commit();
  };
  // PDL: the W stream may start before the previous kernel finishes when the host says nothing
  // in between writes W (prefetch_early); X is the previous kernel's output and must wait.
  if (prefetch_early) {
#pragma unroll
    for (int d = 0; d < kFwdRing - 1; ++d) issue(d);
  }
  sync();
  if (!prefetch_early) {
#pragma unroll
    for (int d = 0; d < kFwdRing - 1; ++d) issue(d);
  }
  SN_T(1);

  // X chunk -> smem, transposed and swizzled (see above); zero outside the batch / the chunk.
  // A thread takes one (batch quad, float4 of k) item: 4 coalesced LDG.128 (one per batch row),
  // a 4x4 register transpose, 4 conflict-free STS.128 (one per k).  All loads of a thread are
  // issued before its stores.  (clock64 phase timing: the first version - one LDG.128 then four
  // STS.32 per item, with a divide - cost 12.5k of the kernel's 56k cycles.)
  {
    constexpr int XI = R == 8 ? 2 : 1;                      // k4 items per thread per pass (16 float4 of temporaries each)
    for (int kbase = 0; kbase < kc4; kbase += NT * XI) {
      float4 v[XI][4][4];
#pragma unroll
      for (int u = 0; u < XI; ++u) {
        const int k4 = kbase + u * NT + tid;
#pragma unroll
        for (int bq = 0; bq < 4; ++bq)
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const int b = bq * 4 + r;
            v[u][bq][r] = zero;
            if (k4 < kc4 && k4 * 4 < klen && b < nb)
              v[u][bq][r] = ld_dep<Sync>(reinterpret_cast<const float4*>(X + (size_t)(b0 + b) * ldx + k0) + k4);
          }
      }
#pragma unroll
      for (int u = 0; u < XI; ++u) {
        const int k4 = kbase + u * NT + tid;
        if (k4 < kc4) {
          const int m = k4 & 31, base = (k4 >> 5) * 128 + m, swz = (m >> 1) & 3;
#pragma unroll
          for (int bq = 0; bq < 4; ++bq) {
            float4* dst = smem4 + (size_t)base * 4 + (bq ^ swz);
            dst[0 * 32 * 4] = make_float4(v[u][bq][0].x, v[u][bq][1].x, v[u][bq][2].x, v[u][bq][3].x);
            dst[1 * 32 * 4] = make_float4(v[u][bq][0].y, v[u][bq][1].y, v[u][bq][2].y, v[u][bq][3].y);
            dst[2 * 32 * 4] = make_float4(v[u][bq][0].z, v[u][bq][1].z, v[u][bq][2].z, v[u][bq][3].z);
            dst[3 * 32 * 4] = make_float4(v[u][bq][0].w, v[u][bq][1].w, v[u][bq][2].w, v[u][bq][3].w);
          }
        }
      }
    }
  }
  SN_T(2);
  __syncthreads();
  SN_T(3);

  float acc[R * kBT];
#pragma unroll
  for (int i = 0; i < R * kBT; ++i) acc[i] = 0.f;
  int g = t0 / iters, it = t0 - g * iters;
  bool has_first = (it == 0);
  int slot = 0;
  const int sw = (lane >> 1) & 3;
  const size_t slot_stride = (size_t)ksplit * Bpad * N;
  float* Pbase = P + ((size_t)ks * Bpad + b0) * N;

  for (int t = t0; t < t1; ++t) {
    issue((slot + kFwdRing - 1) & (kFwdRing - 1));
    cp_async_wait<kFwdRing - 1>();
    float4 wc[R];
#pragma unroll
    for (int i = 0; i < R; ++i) wc[i] = wring[(slot * R + i) * 32 + lane];
    if (ktail && it == iters - 1 && it * 128 + lane * 4 >= klen) {   // nothing was copied for this lane
#pragma unroll
      for (int i = 0; i < R; ++i) wc[i] = zero;
    }
    slot = (slot + 1) & (kFwdRing - 1);
    const float4* xrow = smem4 + (size_t)(it * 128 + lane) * 4;     // physical row kk*32 + lane
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int bq = 0; bq < 4; ++bq) {
        const float4 x = xrow[kk * 32 * 4 + (bq ^ sw)];
        const float2 xlo = make_float2(x.x, x.y), xhi = make_float2(x.z, x.w);
#pragma unroll
        for (int i = 0; i < R; ++i) {
          const float w = kk == 0 ? wc[i].x : kk == 1 ? wc[i].y : kk == 2 ? wc[i].z : wc[i].w;
          const float2 ww = make_float2(w, w);
          float* a = acc + i * kBT + bq * 4;
          const float2 r0 = __ffma2_rn(ww, xlo, make_float2(a[0], a[1]));
          const float2 r1 = __ffma2_rn(ww, xhi, make_float2(a[2], a[3]));
          a[0] = r0.x; a[1] = r0.y; a[2] = r1.x; a[3] = r1.y;
        }
      }
    }
    const bool group_done = (it == iters - 1);
    if (group_done || t == t1 - 1) {
#ifdef SN_TIMING
      const long long tr0 = clock64();
#endif
      // after the butterfly lane l holds the flat indices (row*16 + b) l*R/2 .. l*R/2 + R/2 - 1
      butterfly_reduce<R * kBT>(acc, lane);
      constexpr int PL = R / 2;                    // values per lane
      const int row = g * R + (lane * PL) / kBT;
      if (row < N) {
        float* dst = Pbase + (has_first ? 0 : slot_stride) + row;
#pragma unroll
        for (int j = 0; j < PL; ++j) {
          const int b = (lane * PL) % kBT + j;
          if (b < nb) {
            dst[(size_t)b * N] = acc[j];
            if (has_first && group_done) dst[slot_stride + (size_t)b * N] = 0.f;
          }
        }
      }
#pragma unroll
      for (int i = 0; i < R * kBT; ++i) acc[i] = 0.f;
#ifdef SN_TIMING
      t_red += clock64() - tr0;
#endif
    }
    if (++it == iters) { it = 0; ++g; has_first = true; }
  }
  cp_async_wait<0>();
  SN_T(4);
#ifdef SN_TIMING
  if (lane == 0 && bid.z == 0) {
    long long* o = g_fwd_timing + ((bid.y * gdx + bid.x) * WPC + warp) * 8;
    o[0] = tstamp[1] - tstamp[0]; o[1] = tstamp[2] - tstamp[1]; o[2] = tstamp[3] - tstamp[2];
    o[3] = tstamp[4] - tstamp[3]; o[4] = t_red; o[5] = t1 - t0; o[6] = tstamp[0]; o[7] = tstamp[4];
  }
#endif
}

template <int R>
__global__ void __launch_bounds__(32 * (64 / R), 1)
fwd_skinny_kernel(const float* __restrict__ X, int ldx, int B, const float* __restrict__ W, int ldw, int N,
                  int K, int kc, float* __restrict__ P, int Bpad, int ksplit, int prefetch_early, int l2_keep_w) {
  extern __shared__ float4 smem4[];
  fwd_skinny_body<R>(smem4, blockIdx, (int)gridDim.x, PdlSync{}, X, ldx, B, W, ldw, N, K, kc, P, Bpad, ksplit, prefetch_early,
                     l2_keep_w);
}

// H[b][n] = act(sum_ks P[ks][b][n] + bias[n]) * dropout.   One thread per element.
// mask (nullable) is [B][N] keep-bytes with row stride N; otherwise, when keep_prob < 1, Philox.
template <class Sync>
__device__ __forceinline__ void
fwd_epilogue_body(const unsigned bx, const unsigned gdx, Sync sync, const float* __restrict__ P, int ksplit, int Bpad, int B,
                  int N, const float* __restrict__ bias, int relu, const uint8_t* __restrict__ mask, float keep_prob,
                  float scale, uint64_t seed, uint64_t offset, int layer, float* __restrict__ H, int ldh,
                  float* __restrict__ out2, int idx_ld, int idx_off) {
  // idx_ld > 0: this launch computes rows [idx_off, idx_off+N) of a layer that is idx_ld rows wide
  // (a tensor-parallel shard): the dropout decision of an element must not depend on the sharding,
  // so the mask / Philox index is the element's index in the FULL [B, idx_ld] activation.
  sync();
  const size_t total = (size_t)B * N;
  const size_t stride = (size_t)gdx * blockDim.x;
  size_t idx = bx * (size_t)blockDim.x + threadIdx.x;
  for (; idx < total; idx += stride) {
    const int b = (int)(idx / N), n = (int)(idx - (size_t)b * N);
    const size_t gidx = idx_ld > 0 ? (size_t)b * idx_ld + idx_off + n : idx;
    const float bn = bias[n];                    // requested before the partials, not after them
    float s = 0.f;
    const float* pp = P + (size_t)b * N + n;
    const size_t pstride = (size_t)Bpad * N;
    int ks = 0;
    if (ksplit > 16) {
      // many partials (the 43 k-slice partials of a fused traversal): sixteen loads in flight -
      // with four, this launch is eleven dependent round trips (9-13 us in the chain)
      for (; ks + 16 <= ksplit; ks += 16) {
        float t[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) t[q] = ld_dep<Sync>(pp + (size_t)(ks + q) * pstride);
#pragma unroll
        for (int q = 0; q < 16; ++q) s += t[q];
      }
      if (ks < ksplit) {
        float t[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) t[q] = ks + q < ksplit ? ld_dep<Sync>(pp + (size_t)(ks + q) * pstride) : 0.f;
#pragma unroll
        for (int q = 0; q < 16; ++q)
          if (ks + q < ksplit) s += t[q];
        ks = ksplit;
      }
    }
    for (; ks + 4 <= ksplit; ks += 4) {          // four independent loads in flight, fixed add order
      const float p0 = ld_dep<Sync>(pp + (size_t)ks * pstride), p1 = ld_dep<Sync>(pp + (size_t)(ks + 1) * pstride);
      const float p2 = ld_dep<Sync>(pp + (size_t)(ks + 2) * pstride), p3 = ld_dep<Sync>(pp + (size_t)(ks + 3) * pstride);
      s += p0; s += p1; s += p2; s += p3;
    }
    for (; ks < ksplit; ++ks) s += ld_dep<Sync>(pp + (size_t)ks * pstride);
    s += bn;
    if (relu) s = fmaxf(s, 0.f);
    if (mask) {
      s = mask[gidx] ? s * scale : 0.f;
    } else if (keep_prob < 1.f) {
      s = philox_keep(seed, offset, layer, gidx, keep_prob) ? s * scale : 0.f;
    }
    H[(size_t)b * ldh + n] = s;
    if (out2) out2[idx] = s;
  }
}

__global__ void fwd_epilogue_kernel(const float* __restrict__ P, int ksplit, int Bpad, int B, int N,
                                    const float* __restrict__ bias, int relu, const uint8_t* __restrict__ mask,
                                    float keep_prob, float scale, uint64_t seed, uint64_t offset, int layer,
                                    float* __restrict__ H, int ldh, float* __restrict__ out2, int idx_ld, int idx_off) {
  fwd_epilogue_body(blockIdx.x, gridDim.x, PdlSync{}, P, ksplit, Bpad, B, N, bias, relu, mask, keep_prob, scale, seed, offset,
                    layer, H, ldh, out2, idx_ld, idx_off);
}

// ------------------------------------------------------------------------------------------
// Cross-entropy of one row by one warp: loss = lse - z[y]; dlogits = (softmax - onehot)/B.
// Returns the row loss (same value in every lane).  Register form: the row has at most 32*NC
// classes and `zc` holds this lane's logits (class lane+32i, -inf beyond C), so the three passes run
// from registers - one L2 round trip per row instead of twelve dependent ones (clock64 phase
// timing: 15.8k cycles for two rows per warp before, at ~500 cycles per dependent load).  Memory
// form: CG reads the logits with ld.global.cg (L2), needed when other CTAs of the same launch wrote
// them.
// ------------------------------------------------------------------------------------------
// GW = lanes per row (32: one row per warp; 16: two rows per warp, in lock step - halves the
// dependent-latency chain when a few rows are all a CTA has to do)
template <int NC, int GW = 32>
__device__ __forceinline__ float ce_row_reg(const float (&zc)[NC], int64_t lab, int C, float invB,
                                            float* __restrict__ dzr, int gl) {
  float m = -INFINITY;
  int arg = 0x7fffffff;
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int c = gl + GW * i;
    if (c < C && zc[i] > m) { m = zc[i]; arg = c; }
  }
#pragma unroll
  for (int o = GW / 2; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  // sum of exp(z-m) WITHOUT the max term (which is exactly 1): log1p keeps full precision when
  // one class dominates, where log(1+eps) computed from a rounded sum would lose it
  float s1 = 0.f, zl = 0.f, ex[NC];
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int c = gl + GW * i;
    ex[i] = 1.f;
    if (c < C) {
      if (c != arg) { ex[i] = expf(zc[i] - m); s1 += ex[i]; }
      if (c == lab) zl = zc[i];
    }
  }
#pragma unroll
  for (int o = GW / 2; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    zl += __shfl_xor_sync(0xffffffffu, zl, o);     // exactly one lane holds z[label], the others add zeros
  }
  const bool ok = lab >= 0 && lab < C;
  // -log_softmax(z)[y] = log(sum exp(z-m)) - (z[y]-m): no cancellation against the row max
  const float loss = ok ? (log1pf(s1) - (zl - m)) : NAN;   // torch raises on a bad target
  if (dzr) {
    const float invs = 1.f / (1.f + s1);
#pragma unroll
    for (int i = 0; i < NC; ++i) {
      const int c = gl + GW * i;
      if (c < C) {
        float p = ex[i] * invs;
        if (c == lab) p -= 1.f;
        dzr[c] = p * invB;
      }
    }
  }
  return loss;
}

// The same with the gradient returned in registers (dzv[i] for class gl + GW*i) instead of stored.
template <int NC, int GW = 32>
__device__ __forceinline__ float ce_row_reg_vals(const float (&zc)[NC], int64_t lab, int C, float invB, float (&dzv)[NC],
                                                 int gl) {
  float m = -INFINITY;
  int arg = 0x7fffffff;
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int c = gl + GW * i;
    if (c < C && zc[i] > m) { m = zc[i]; arg = c; }
  }
#pragma unroll
  for (int o = GW / 2; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  float s1 = 0.f, zl = 0.f, ex[NC];
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int c = gl + GW * i;
    ex[i] = 1.f;
    if (c < C) {
      if (c != arg) { ex[i] = expf(zc[i] - m); s1 += ex[i]; }
      if (c == lab) zl = zc[i];
    }
  }
#pragma unroll
  for (int o = GW / 2; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    zl += __shfl_xor_sync(0xffffffffu, zl, o);
  }
  const bool ok = lab >= 0 && lab < C;
  const float loss = ok ? (log1pf(s1) - (zl - m)) : NAN;
  const float invs = 1.f / (1.f + s1);
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int c = gl + GW * i;
    float p = ex[i] * invs;
    if (c == lab) p -= 1.f;
    dzv[i] = c < C ? p * invB : 0.f;
  }
  return loss;
}

// the same from memory, for rows of any length (three streaming passes)
template <bool CG>
__device__ __forceinline__ float ce_row_mem(const float* __restrict__ zr, int64_t lab, int C, float invB,
                                            float* __restrict__ dzr, int lane) {
  auto ld = [&](int c) { return CG ? __ldcg(zr + c) : zr[c]; };
  float m = -INFINITY;
  int arg = 0x7fffffff;
  for (int c = lane; c < C; c += 32) {
    const float v = ld(c);
    if (v > m) { m = v; arg = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  float s1 = 0.f;
  for (int c = lane; c < C; c += 32)
    if (c != arg) s1 += expf(ld(c) - m);
  s1 = warp_sum(s1);
  const float s = 1.f + s1;
  const bool ok = lab >= 0 && lab < C;
  const float loss = ok ? (log1pf(s1) - (ld(ok ? (int)lab : 0) - m)) : NAN;
  if (dzr) {
    const float invs = 1.f / s;
    for (int c = lane; c < C; c += 32) {
      float p = expf(ld(c) - m) * invs;
      if (c == lab) p -= 1.f;
      dzr[c] = p * invB;
    }
  }
  return loss;
}

constexpr int kCeRegClasses = 4;     // rows of up to 128 classes are held in registers (4 per lane)

// one warp, one row; picks the register-resident form when the row is short enough
template <bool CG>
__device__ __forceinline__ float ce_row_auto(const float* __restrict__ zr, int64_t lab, int C, float invB,
                                             float* __restrict__ dzr, int lane) {
  if (C <= 32 * kCeRegClasses) {
    float zc[kCeRegClasses];
#pragma unroll
    for (int i = 0; i < kCeRegClasses; ++i) {
      const int c = lane + 32 * i;
      zc[i] = c < C ? (CG ? __ldcg(zr + c) : zr[c]) : -INFINITY;
    }
    return ce_row_reg<kCeRegClasses>(zc, lab, C, invB, dzr, lane);
  }
  return ce_row_mem<CG>(zr, lab, C, invB, dzr, lane);
}

// Cross-entropy fused behind the last layer (fwd_small_kernel): labels == nullptr turns it off.
// The logits H[B][N] are read as rows = B*heads rows of C = N/heads classes (multi-head CE).
struct CeTail {
  const int64_t* labels;
  float* dz;            // [rows][C]
  float* row_loss;      // [rows] scratch
  unsigned* ticket;     // zero before the launch; re-armed by the kernel
  float* loss;          // mean over rows
  int rows, C;
};

// Tiny layers (fc3 100x1000, the U-shape tail 10x100): the weight matrix is L2-resident and the
// launch is pure latency, so the split-K machinery above costs more than it saves.  One launch,
// bias / ReLU / dropout fused, no global partial sums.
//   grid (ceil(N/2), batch tiles), 256 threads.  The CTA owns kHeadRows = 2 output rows.  Thread
//   (warp w, lane l) owns the 4 k-columns of float4 w*32+l (+256 per pass): it holds X[16][4] and
//   W[2 rows][4] in registers - loaded straight from global, one round trip, no staging - and
//   does 128 FMAs per pass; a 32-value butterfly sums across the lanes, the 8 warps' sums meet in
//   shared memory and 32 threads finish (fixed order: deterministic).
//   (The first version gave each WARP a row and read the whole X tile from shared memory in
//   every warp: 8 x 64 KB of LDS per CTA, 3.4 us of shared-memory bandwidth for 0.2 us of math.)
// In a training step this is the last layer and the loss follows: the last CTA to finish (ticket)
// computes the cross-entropy of all rows and the logit gradient in the same launch.
constexpr int kHeadRows = 2;

template <class Sync>
__device__ __forceinline__ void
fwd_small_body(const dim3 bid, const unsigned ncta, Sync sync,
               const float* __restrict__ X, int ldx, int B, const float* __restrict__ W, int ldw, int N, int K,
               const float* __restrict__ bias, int relu, const uint8_t* __restrict__ mask, float keep_prob,
               float scale, uint64_t seed, uint64_t offset, int layer, float* __restrict__ H, int ldh,
               float* __restrict__ out2, int prefetch_early, CeTail ce) {
  __shared__ float part[8][kHeadRows * kBT];
  __shared__ float rl[256];                          // row losses of the tail (rows beyond 256: global scratch)
  __shared__ unsigned is_last;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = bid.x * kHeadRows;
  const int b0 = bid.y * kBT, nb = min(kBT, B - b0);
  const int K4 = (K + 3) >> 2;
  const bool wvec = (K & 3) == 0 && (ldw & 3) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0;
  const bool xvec = (K & 3) == 0 && (ldx & 3) == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  SN_TP_DECL;
  SN_TP(0);

  auto load4 = [&](const float* row, int k4, bool vec) {
    if (vec) return __ldg(reinterpret_cast<const float4*>(row) + k4);
    const int k = k4 * 4;
    float4 v;
    v.x = k < K ? __ldg(row + k) : 0.f;
    v.y = k + 1 < K ? __ldg(row + k + 1) : 0.f;
    v.z = k + 2 < K ? __ldg(row + k + 2) : 0.f;
    v.w = k + 3 < K ? __ldg(row + k + 3) : 0.f;
    return v;
  };
  auto load4x = [&](const float* row, int k4, bool vec) {        // X may come from an earlier phase
    if (vec) return ld_dep<Sync>(reinterpret_cast<const float4*>(row) + k4);
    const int k = k4 * 4;
    float4 v;
    v.x = k < K ? ld_dep<Sync>(row + k) : 0.f;
    v.y = k + 1 < K ? ld_dep<Sync>(row + k + 1) : 0.f;
    v.z = k + 2 < K ? ld_dep<Sync>(row + k + 2) : 0.f;
    v.w = k + 3 < K ? ld_dep<Sync>(row + k + 3) : 0.f;
    return v;
  };
  auto load_w = [&](int k4, float4* w) {
#pragma unroll
    for (int r = 0; r < kHeadRows; ++r)
      w[r] = (k4 < K4 && n0 + r < N) ? load4(W + (size_t)(n0 + r) * ldw, k4, wvec) : zero4;
  };

  // when the host says the previous kernel cannot be writing W or the bias (prefetch_early), the
  // first pass's W values are fetched before the PDL wait; otherwise right after it
  float4 w[kHeadRows];
  if (!prefetch_early) sync();
  load_w(warp * 32 + lane, w);
  float bv = 0.f;
  if (tid < kHeadRows * kBT && n0 + (tid >> 4) < N) bv = __ldg(bias + n0 + (tid >> 4));
  if (prefetch_early) sync();
  SN_TP(1);

  float acc[kHeadRows * kBT];
#pragma unroll
  for (int i = 0; i < kHeadRows * kBT; ++i) acc[i] = 0.f;
  for (int kb = 0; kb < K4; kb += 256) {
    const int k4 = kb + warp * 32 + lane;
    if (kb > 0) load_w(k4, w);
    float4 x[kBT];
#pragma unroll
    for (int b = 0; b < kBT; ++b) x[b] = (k4 < K4 && b < nb) ? load4x(X + (size_t)(b0 + b) * ldx, k4, xvec) : zero4;
#pragma unroll
    for (int r = 0; r < kHeadRows; ++r)
#pragma unroll
      for (int b = 0; b < kBT; ++b) {
        float a = acc[r * kBT + b];
        a = __fmaf_rn(w[r].x, x[b].x, a);
        a = __fmaf_rn(w[r].y, x[b].y, a);
        a = __fmaf_rn(w[r].z, x[b].z, a);
        acc[r * kBT + b] = __fmaf_rn(w[r].w, x[b].w, a);
      }
  }
  SN_TP(2);
  // after the butterfly lane l holds the flat indices (row*16 + b) l*V/32 .. l*V/32 + V/32 - 1
  butterfly_reduce<kHeadRows * kBT>(acc, lane);
  constexpr int kPerLane = kHeadRows * kBT / 32;
#pragma unroll
  for (int j = 0; j < kPerLane; ++j) part[warp][kPerLane * lane + j] = acc[j];
  __syncthreads();
  if (tid < kHeadRows * kBT) {
    const int r = tid >> 4, b = tid & 15, n = n0 + r;
    if (n < N && b < nb) {
      float s = part[0][tid];
#pragma unroll
      for (int q = 1; q < 8; ++q) s += part[q][tid];
      s += bv;
      if (relu) s = fmaxf(s, 0.f);
      const size_t idx = (size_t)(b0 + b) * N + n;
      if (mask) {
        s = mask[idx] ? s * scale : 0.f;
      } else if (keep_prob < 1.f) {
        s = philox_keep(seed, offset, layer, idx, keep_prob) ? s * scale : 0.f;
      }
      H[(size_t)(b0 + b) * ldh + n] = s;
      if (out2) out2[idx] = s;
    }
  }
  if (!ce.labels) return;
  // ---- loss tail: the last CTA of the grid sees every logit
  // labels of this warp's first two rows, fetched while the ticket round trip is in flight
  const int r_a = warp, r_b = warp + 8;
  const int64_t lab_a = r_a < ce.rows ? ce.labels[r_a] : 0, lab_b = r_b < ce.rows ? ce.labels[r_b] : 0;
  SN_TP(3);
  __threadfence();
  __syncthreads();
  SN_TP(4);
  if (tid == 0) is_last = (atomicAdd(ce.ticket, 1u) == ncta - 1);
  __syncthreads();
  SN_TP(5);
  if (!is_last) return;
  __threadfence();
  const float invB = 1.f / (float)ce.rows;
  if (ce.C <= 128) {
    // two rows per warp (16 lanes x 8 classes each), sixteen rows per pass of the CTA
    const int gl = lane & 15;
    for (int r0 = warp * 2; r0 < ce.rows; r0 += 16) {     // warp-uniform trip count: the shuffles use full masks
      const int r = r0 + (lane >> 4);
      const bool live = r < ce.rows;
      const float* zr = H + (size_t)(live ? r : 0) * ce.C;
      float zc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int c = gl + 16 * i;
        zc[i] = (live && c < ce.C) ? __ldcg(zr + c) : -INFINITY;
      }
      const int64_t lab = live ? ce.labels[r] : 0;
      const float l = ce_row_reg<8, 16>(zc, lab, live ? ce.C : 0, invB, (live && ce.dz) ? ce.dz + (size_t)r * ce.C : nullptr, gl);
      if (live && gl == 0) {
        if (r < 256) rl[r] = l; else ce.row_loss[r] = l;
      }
    }
  } else {
    for (int r = warp; r < ce.rows; r += 8) {
      const int64_t lab = r == r_a ? lab_a : r == r_b ? lab_b : ce.labels[r];
      const float l = ce_row_auto<true>(H + (size_t)r * ce.C, lab, ce.C, invB, ce.dz ? ce.dz + (size_t)r * ce.C : nullptr, lane);
      if (lane == 0) {
        if (r < 256) rl[r] = l; else ce.row_loss[r] = l;
      }
    }
  }
  __syncthreads();
  SN_TP(6);
  SN_TP_FLUSH();
  if (warp == 0) {
    // fixed-order sum: lanes take rows lane, lane+32, ...; then a shuffle tree (same as ce_kernel)
    float s = 0.f;
    for (int r = lane; r < ce.rows; r += 32) s += r < 256 ? rl[r] : ce.row_loss[r];
    s = warp_sum(s);
    if (lane == 0) {
      if (ce.loss) *ce.loss = s / (float)ce.rows;
      *ce.ticket = 0u;
    }
  }
}

__global__ void __launch_bounds__(256)
fwd_small_kernel(const float* __restrict__ X, int ldx, int B, const float* __restrict__ W, int ldw, int N, int K,
                 const float* __restrict__ bias, int relu, const uint8_t* __restrict__ mask, float keep_prob,
                 float scale, uint64_t seed, uint64_t offset, int layer, float* __restrict__ H, int ldh,
                 float* __restrict__ out2, int prefetch_early, CeTail ce) {
  fwd_small_body(blockIdx, gridDim.x * gridDim.y, PdlSync{}, X, ldx, B, W, ldw, N, K, bias, relu, mask, keep_prob, scale, seed,
                 offset, layer, H, ldh, out2, prefetch_early, ce);
}

// ------------------------------------------------------------------------------------------
// Cross-entropy as its own launch: one warp per row; loss = mean_b(lse - z[y]).
// Row losses go to row_loss[B]; the last CTA (ticket) sums them in row order.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ce_kernel(const float* __restrict__ z, int ldz, const int64_t* __restrict__ y,
                                                 int B, int C, float* __restrict__ dz, int lddz,
                                                 float* __restrict__ row_loss, unsigned* __restrict__ ticket,
                                                 float* __restrict__ loss_out, float* __restrict__ loss_out2) {
  pdl_wait();
  pdl_launch();
  __shared__ unsigned is_last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  if (b < B) {
    const float l = ce_row_auto<false>(z + (size_t)b * ldz, y[b], C, 1.f / (float)B, dz ? dz + (size_t)b * lddz : nullptr, lane);
    if (lane == 0) row_loss[b] = l;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (warp == 0) {
    // fixed-order sum: lanes take rows lane, lane+32, ...; then a shuffle tree
    float s = 0.f;
    for (int r = lane; r < B; r += 32) s += __ldcg(row_loss + r);
    s = warp_sum(s);
    if (lane == 0) {
      const float L = s / (float)B;
      if (loss_out) *loss_out = L;
      if (loss_out2) *loss_out2 = L;
      *ticket = 0u;
    }
  }
}

// Cross-entropy over LONG rows (an Alice's local loss treats her 5408 activations as logits): one
// CTA per row, the whole row in registers (NV values per thread, one coalesced read), CTA-wide
// argmax / sum through shared memory.  The warp-per-row kernel streams such a row three times from
// one warp: 21.5 us of the 31 us local step (profiles/micro/step_trace.py local); this form reads it once.
// Same first-argmax, log1p and gradient formulas; fixed reduction order (deterministic).
template <int NV>
__global__ void __launch_bounds__(256) ce_wide_kernel(const float* __restrict__ z, int ldz, const int64_t* __restrict__ y,
                                                      int B, int C, float* __restrict__ dz, int lddz,
                                                      float* __restrict__ row_loss, unsigned* __restrict__ ticket,
                                                      float* __restrict__ loss_out) {
  __shared__ float sm_m[8], sm_s[8], sm_zl;
  __shared__ int sm_a[8];
  __shared__ unsigned is_last;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x;
  const float* zr = z + (size_t)b * ldz;
  pdl_wait();
  pdl_launch();
  const int64_t lab = y[b];
  float v[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = tid + 256 * i;
    v[i] = c < C ? zr[c] : -INFINITY;
  }
  float m = -INFINITY;
  int arg = 0x7fffffff;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = tid + 256 * i;
    if (c < C && v[i] > m) { m = v[i]; arg = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  if (lane == 0) { sm_m[warp] = m; sm_a[warp] = arg; }
  if (tid == 0) sm_zl = 0.f;
  __syncthreads();
  m = sm_m[0];
  arg = sm_a[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) {
    const float om = sm_m[w];
    const int oa = sm_a[w];
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  float s1 = 0.f, ex[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = tid + 256 * i;
    ex[i] = 1.f;
    if (c < C) {
      if (c != arg) { ex[i] = expf(v[i] - m); s1 += ex[i]; }
      if (c == lab) sm_zl = v[i];                       // exactly one thread
    }
  }
  s1 = warp_sum(s1);
  if (lane == 0) sm_s[warp] = s1;
  __syncthreads();
  s1 = sm_s[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) s1 += sm_s[w];
  const bool ok = lab >= 0 && lab < C;
  if (tid == 0) row_loss[b] = ok ? (log1pf(s1) - (sm_zl - m)) : NAN;   // torch raises on a bad target
  if (dz) {
    const float invs = 1.f / (1.f + s1), invB = 1.f / (float)B;
    float* dzr = dz + (size_t)b * lddz;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = tid + 256 * i;
      if (c < C) {
        float p = ex[i] * invs;
        if (c == lab) p -= 1.f;
        dzr[c] = p * invB;
      }
    }
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (warp == 0) {
    float s = 0.f;
    for (int r = lane; r < B; r += 32) s += __ldcg(row_loss + r);
    s = warp_sum(s);
    if (lane == 0) {
      if (loss_out) *loss_out = s / (float)B;
      *ticket = 0u;
    }
  }
}

// counters6 += {hit, total, hit&omit, omit, hit&!omit, !omit}; prediction = first argmax.
__global__ void __launch_bounds__(256) eval_counts_kernel(const float* __restrict__ z, const int64_t* __restrict__ y,
                                                          int B, int C, int64_t omit,
                                                          unsigned long long* __restrict__ counters) {
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  if (b >= B) return;
  const float* zr = z + (size_t)b * C;
  float m = -INFINITY;
  int arg = 0x7fffffff;
  for (int c = lane; c < C; c += 32) {
    const float v = zr[c];
    if (v > m || (v == m && c < arg)) { m = v; arg = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  if (lane == 0) {
    const int64_t lab = y[b];
    const bool hit = (int64_t)arg == lab, unl = lab == omit;
    atomicAdd(counters + 1, 1ull);
    if (hit) atomicAdd(counters + 0, 1ull);
    if (unl) { atomicAdd(counters + 3, 1ull); if (hit) atomicAdd(counters + 2, 1ull); }
    else { atomicAdd(counters + 5, 1ull); if (hit) atomicAdd(counters + 4, 1ull); }
  }
}

// ------------------------------------------------------------------------------------------
// Backward + optimizer, skinny: ONE traversal of W (and its optimizer state) per layer.
//   For each owned element: dW = sum_b dZ[b][n]*A[b][k] is formed in registers from the 16-row
//   dZ and A (never written to HBM), dA[b][k] += dZ[b][n]*W_old[n][k] is accumulated, and the
//   optimizer update is applied in place.  grid (k groups, n splits), 256 threads: warp
//   (wk, wr) owns the 128-wide k slice (blockIdx.x*kw + wk) — one float4 per lane — and rows
//   nb0+wr, nb0+wr+rw, ... of the CTA's row range.  Lanes keep A[16][4] and the dA accumulators
//   [16][4] in registers; dZ rows are broadcast from shared memory.  dA partials are reduced
//   across the CTA's row-warps in shared memory and written to PA[blockIdx.y][b][k]; the n-split
//   partials are summed in fixed order by bwd_epilogue_kernel.
// ------------------------------------------------------------------------------------------
#ifndef SN_KRING
#define SN_KRING 8
#endif
constexpr int kRing = SN_KRING;   // rows in flight per thread (W + optimizer state): ~56-84 KB per SM (power of two)

// FUSE_FWD (SISA server epochs): the NEXT batch's activations are already known, so the same
// traversal also produces the next step's forward through this layer with the freshly updated
// weights: per row, each lane dots its 4 new weights with Xn[16][4], a 16-value butterfly sums
// across the warp's 128-wide k slice, and the per-slice partials PF[slice][b][n] are summed by
// fwd_epilogue_kernel at the start of the next step.  W is then read once per step instead of
// twice (28*P -> 24*P bytes with Adam).
template <int OPT, bool NEED_DA, bool FUSE_FWD, bool PFS, class Sync>
__device__ __forceinline__ void
bwd_update_body(float4* smem4, const dim3 bid, Sync sync,
                float* __restrict__ W, int ldw, int N, int K, float* __restrict__ bias, float* __restrict__ s1,
                float* __restrict__ s2, float* __restrict__ bs1, float* __restrict__ bs2,
                const float* __restrict__ dZ, int ldz, const float* __restrict__ A, int lda, int B,
                float* __restrict__ PA, int kw, int rows_per_cta, OptParams op,
                const float* __restrict__ Xn, int Bn, float* __restrict__ PF, int l2_keep_w,
                const float* __restrict__ gate, float gscale) {
  constexpr bool pf_smem = PFS;
  // gate != nullptr (tensor-parallel shard): dZ arrives UNGATED (it is the head GPU's cut gradient
  // w.r.t. this layer's output); the ReLU/dropout gate is this shard's own dense activation
  // gate[b][N]: dZ <- dZ * (gate > 0 ? gscale : 0), applied while the tile is staged.
  // rows below l2_keep_w: weights tagged evict_last, their optimizer state evict_first
  const uint64_t pol_keep = l2_policy(1), pol_norm = l2_policy(0), pol_s = l2_policy(l2_keep_w ? 2 : 0);
  float* dzs = reinterpret_cast<float*>(smem4);                  // [rows_per_cta][16]
  // per-thread ring of the next kRing rows' float4s: [array][slot][thread]; a thread only ever
  // reads what it copied itself, so cp.async.wait_group is the only synchronisation needed
  float4* ring = smem4 + ((rows_per_cta * kBT + 3) >> 2);
  // FUSE_FWD with pf_smem: the next-forward partials of this CTA's rows are collected in shared memory
  // ([k slice of the CTA][row][16 batch values, padded to 17]) and written out once at the end, summed
  // over the CTA's kw slices: kgroups partials per output instead of one per 128-wide k slice, and
  // coalesced stores instead of sixteen scattered 4-byte stores per row and warp
  float* pfs = reinterpret_cast<float*>(ring + (OPT == 2 ? 3 : 2) * kRing * 256);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wk = warp % kw, wr = warp / kw, rw = 8 / kw;
  const int nb0 = bid.y * rows_per_cta, nb1 = min(N, nb0 + rows_per_cta);
  const int nrows = nb1 - nb0;
  const int k0 = (bid.x * kw + wk) * 128 + lane * 4;
  const bool kok = k0 < K;

  auto issue = [&](int slot, int row) {
    if (kok && row < nb1) {
      const size_t o = (size_t)row * ldw + k0;
      if (OPT != 3 || NEED_DA) cp_async16_hint(ring + (0 * kRing + slot) * 256 + tid, W + o, row < l2_keep_w ? pol_keep : pol_norm);
      if (OPT != 4 && (OPT != 3 || !op.acc_first)) cp_async16_hint(ring + (1 * kRing + slot) * 256 + tid, s1 + o, pol_s);
      if (OPT == 2) cp_async16_hint(ring + (2 * kRing + slot) * 256 + tid, s2 + o, pol_s);
    }
    cp_async_commit();
  };
  int n = nb0 + wr;
  if (OPT != 3) {
#pragma unroll
    for (int d = 0; d < kRing - 1; ++d) issue(d, n + d * rw);    // start streaming before anything else
  }
  sync();            // W / optimizer state are never written by the kernel (phase) just before this one;
                     // dZ (and A) are, so everything below waits for it
  if (OPT == 3) {    // ... except the gradient buffer of an accumulation launch (the previous tile's launch wrote it)
#pragma unroll
    for (int d = 0; d < kRing - 1; ++d) issue(d, n + d * rw);
  }

  {  // dZ tile, transposed to [row][16]; loads batched ahead of their stores (latency, not bandwidth)
    constexpr int SB = 8;
    const int total = nrows * kBT;
    for (int base_idx = tid; base_idx < total; base_idx += 256 * SB) {
      float v[SB];
#pragma unroll
      for (int q = 0; q < SB; ++q) {
        const int idx = base_idx + q * 256;
        const int b = idx / nrows, j = idx - b * nrows;
        v[q] = (idx < total && b < B) ? ld_dep<Sync>(dZ + (size_t)b * ldz + nb0 + j) : 0.f;
        if (gate && idx < total && b < B) v[q] = ld_dep<Sync>(gate + (size_t)b * N + nb0 + j) > 0.f ? v[q] * gscale : 0.f;
      }
#pragma unroll
      for (int q = 0; q < SB; ++q) {
        const int idx = base_idx + q * 256;
        if (idx < total) {
          const int b = idx / nrows, j = idx - b * nrows;
          dzs[j * kBT + b] = v[q];
        }
      }
    }
  }
  // Packed math (fma.rn.f32x2): every product of the loop pairs batch rows (2p, 2p+1), so A, the next
  // batch and the dA accumulators are held as float2 register pairs [pair][column].  This kernel was
  // 57-61 % issue-active in ncu at 4.6 TB/s; packing halves its FMA issue slots (128 -> 64 per row).
  constexpr int kBP = kBT / 2;
  auto load_pairs = [&](const float* base, int ld, int rows, float2 (&dst)[kBP][4]) {
#pragma unroll
    for (int p = 0; p < kBP; ++p) {
      const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 e = (kok && 2 * p < rows) ? ld_dep<Sync>(reinterpret_cast<const float4*>(base + (size_t)(2 * p) * ld + k0)) : zero;
      const float4 o = (kok && 2 * p + 1 < rows) ? ld_dep<Sync>(reinterpret_cast<const float4*>(base + (size_t)(2 * p + 1) * ld + k0)) : zero;
      dst[p][0] = make_float2(e.x, o.x);
      dst[p][1] = make_float2(e.y, o.y);
      dst[p][2] = make_float2(e.z, o.z);
      dst[p][3] = make_float2(e.w, o.w);
    }
  };
  float2 a2[kBP][4];
  load_pairs(A, lda, B, a2);
  float2 acc2[NEED_DA ? kBP : 1][4];
  if (NEED_DA) {
#pragma unroll
    for (int p = 0; p < kBP; ++p)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc2[p][c] = make_float2(0.f, 0.f);
  }
  float2 xn2[kBP][4];                               // dead (and removed) unless FUSE_FWD
  if (FUSE_FWD) load_pairs(Xn, lda, Bn, xn2);
  const int kslice = bid.x * kw + wk;
  const bool slice_ok = kslice * 128 < K;
  __syncthreads();

  int slot = 0;
  for (; n < nb1; n += rw) {
    issue((slot + kRing - 1) & (kRing - 1), n + (kRing - 1) * rw);
    cp_async_wait<kRing - 1>();
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f), m1 = w, m2 = w;
    if (kok) {
      if (OPT != 3 || NEED_DA) w = ring[(0 * kRing + slot) * 256 + tid];
      if (OPT != 4 && (OPT != 3 || !op.acc_first)) m1 = ring[(1 * kRing + slot) * 256 + tid];
      if (OPT == 2) m2 = ring[(2 * kRing + slot) * 256 + tid];
    }
    slot = (slot + 1) & (kRing - 1);
    const float4* dzr = reinterpret_cast<const float4*>(dzs + (n - nb0) * kBT);
    float2 dz2[kBP];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 t = dzr[q];
      dz2[2 * q] = make_float2(t.x, t.y);
      dz2[2 * q + 1] = make_float2(t.z, t.w);
    }
    // dW: even and odd batch rows accumulate in the two halves, added at the end
    float2 g2[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
    const float2 wb[4] = {make_float2(w.x, w.x), make_float2(w.y, w.y), make_float2(w.z, w.z), make_float2(w.w, w.w)};
#pragma unroll
    for (int p = 0; p < kBP; ++p) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        g2[c] = __ffma2_rn(dz2[p], a2[p][c], g2[c]);
        if (NEED_DA) acc2[p][c] = __ffma2_rn(dz2[p], wb[c], acc2[p][c]);     // dA uses the OLD weights
      }
    }
    float4 g = make_float4(g2[0].x + g2[0].y, g2[1].x + g2[1].y, g2[2].x + g2[2].y, g2[3].x + g2[3].y);
    if (OPT == 4) {
      // input-gradient-only launch (tensor-parallel head: dA goes out to the shards before the update
      // traversal starts): the weights are read, nothing is written per row
    } else if (OPT == 3) {
      // accumulation launch (batches of 17..64 rows go through in tiles of 16): the gradient of this
      // tile is added into the buffer behind `s1`; the weights are only read (for dA)
      if (kok) {
        m1 = make_float4(m1.x + g.x, m1.y + g.y, m1.z + g.z, m1.w + g.w);     // m1 is zero on the first tile
        *reinterpret_cast<float4*>(s1 + (size_t)n * ldw + k0) = m1;
      }
    } else if (kok) {
      opt_update<OPT>(op, w.x, g.x, m1.x, m2.x);
      opt_update<OPT>(op, w.y, g.y, m1.y, m2.y);
      opt_update<OPT>(op, w.z, g.z, m1.z, m2.z);
      opt_update<OPT>(op, w.w, g.w, m1.w, m2.w);
      const size_t o = (size_t)n * ldw + k0;
      if (l2_keep_w) {
        st_hint(reinterpret_cast<float4*>(W + o), w, n < l2_keep_w ? pol_keep : pol_norm);
        st_hint(reinterpret_cast<float4*>(s1 + o), m1, pol_s);
        if (OPT == 2) st_hint(reinterpret_cast<float4*>(s2 + o), m2, pol_s);
      } else {
        *reinterpret_cast<float4*>(W + o) = w;
        __stcs(reinterpret_cast<float4*>(s1 + o), m1);
        if (OPT == 2) __stcs(reinterpret_cast<float4*>(s2 + o), m2);
      }
    }
    if (FUSE_FWD) {                                   // next step's forward with the updated row
      float y[kBT];
#pragma unroll
      for (int p = 0; p < kBP; ++p) {
        float2 t = __fmul2_rn(make_float2(w.x, w.x), xn2[p][0]);
        t = __ffma2_rn(make_float2(w.y, w.y), xn2[p][1], t);
        t = __ffma2_rn(make_float2(w.z, w.z), xn2[p][2], t);
        t = __ffma2_rn(make_float2(w.w, w.w), xn2[p][3], t);
        y[2 * p] = t.x;
        y[2 * p + 1] = t.y;
      }
      // (Sharing one 32-value butterfly between two consecutive rows - same five shuffle stages, twice
      // the independent work per stage - was built and measured SLOWER: 122.8 -> 134 us for this
      // traversal, 228 registers instead of 185.)
      butterfly_reduce<kBT>(y, lane);               // lanes 2b, 2b+1 hold batch row b
      if (pf_smem) {
        if (!(lane & 1)) pfs[((size_t)wk * rows_per_cta + (n - nb0)) * (kBT + 1) + (lane >> 1)] = slice_ok ? y[0] : 0.f;
      } else if (slice_ok && !(lane & 1)) {
        PF[((size_t)kslice * kBT + (lane >> 1)) * N + n] = y[0];
      }
    }
  }
  cp_async_wait<0>();
  if (FUSE_FWD && pf_smem) {
    __syncthreads();
    for (int idx = tid; idx < nrows * kBT; idx += 256) {
      const int b = idx / nrows, j = idx - b * nrows;
      float t = pfs[(size_t)j * (kBT + 1) + b];
      for (int q = 1; q < kw; ++q) t += pfs[((size_t)q * rows_per_cta + j) * (kBT + 1) + b];   // fixed order
      PF[((size_t)bid.x * kBT + b) * N + nb0 + j] = t;
    }
  }

  // bias: db[n] = sum_b dZ[b][n].  One thread per row AFTER the streaming loop (a per-row
  // load->update->store chain inside the loop made the k-group-0 CTAs latency-bound and the
  // whole grid wait for them).
  if (OPT != 3 && OPT != 4 && bid.x == 0) {
    for (int j = tid; j < nrows; j += 256) {
      float db = 0.f;
#pragma unroll
      for (int b = 0; b < kBT; ++b) db += dzs[j * kBT + b];
      const int nn = nb0 + j;
      float p = bias[nn], q1 = bs1[nn], q2 = (OPT == 2) ? bs2[nn] : 0.f;
      opt_update<OPT>(op, p, db, q1, q2);
      bias[nn] = p;
      bs1[nn] = q1;
      if (OPT == 2) bs2[nn] = q2;
    }
  }

  if (NEED_DA) {
    // cross-warp reduction over the row-warps, fixed order; the ring is dead, reuse it
    float4* red = ring;                                          // [8 warps][16][32 lanes] float4 = 64 KB
    __syncthreads();
#pragma unroll
    for (int b = 0; b < kBT; ++b) {
      const int p = b >> 1;
      red[(warp * kBT + b) * 32 + lane] = (b & 1) ? make_float4(acc2[p][0].y, acc2[p][1].y, acc2[p][2].y, acc2[p][3].y)
                                                  : make_float4(acc2[p][0].x, acc2[p][1].x, acc2[p][2].x, acc2[p][3].x);
    }
    __syncthreads();
    for (int o = tid; o < kw * kBT * 32; o += 256) {
      const int l = o & 31, b = (o >> 5) % kBT, s = o / (32 * kBT);
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < rw; ++r) {
        const float4 u = red[((r * kw + s) * kBT + b) * 32 + l];
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      const int k = (bid.x * kw + s) * 128 + l * 4;
      if (k < K && b < B) *reinterpret_cast<float4*>(PA + ((size_t)bid.y * kBT + b) * K + k) = t;
    }
  }
}

template <int OPT, bool NEED_DA, bool FUSE_FWD = false, bool PFS = false>
__global__ void __launch_bounds__(256, 1)
bwd_update_kernel(float* __restrict__ W, int ldw, int N, int K, float* __restrict__ bias, float* __restrict__ s1,
                  float* __restrict__ s2, float* __restrict__ bs1, float* __restrict__ bs2,
                  const float* __restrict__ dZ, int ldz, const float* __restrict__ A, int lda, int B,
                  float* __restrict__ PA, int kw, int rows_per_cta, OptParams op,
                  const float* __restrict__ Xn, int Bn, float* __restrict__ PF, int l2_keep_w,
                  const float* __restrict__ gate, float gscale) {
  extern __shared__ float4 smem4[];
  bwd_update_body<OPT, NEED_DA, FUSE_FWD, PFS>(smem4, blockIdx, PdlSync{}, W, ldw, N, K, bias, s1, s2, bs1, bs2, dZ, ldz, A, lda, B,
                                               PA, kw, rows_per_cta, op, Xn, Bn, PF, l2_keep_w, gate, gscale);
}

// Backward + optimizer for tiny layers (fc3 100x1000: 400 KB, L2-resident, pure latency).  The
// streaming kernel above needs an n split to fill the GPU and then a second launch to add the
// splits; here a CTA sees EVERY row, so its input-gradient sum is final and the gate is applied
// in place: one launch.  To still spread the 3.2 MFMA over enough SMs the k slices are narrow:
//   grid ceil(K/32), 256 threads.  Lane l owns the single column k = 32*blockIdx.x + l; warp w
//   owns rows w, w+8, ... (N <= 128: at most 16 per warp, their W / optimizer-state values
//   prefetched into registers BEFORE the dependency wait).  Per row: dW = sum_b dZ[b][n]*A[b][k]
//   (b order, as everywhere), dA[b] += dZ[b][n]*W_old, optimizer update in place.  The 8 warps' dA
//   sums meet in shared memory (fixed order), are gated with A itself (A is the gate activation of
//   a hidden layer) and written to dst[b][k].
constexpr int kSmallBwdRows = 16;   // rows per warp held in registers => N <= 128

// `ce` (tail kernel only, labels != nullptr): dZ is not read - the CTA first computes the cross-entropy
// of the logits ce.dz-shaped [rows][C] held in `logits` itself (every CTA does, redundantly: 16 rows of
// 100 classes cost less than a grid barrier) and block 0 also stores the loss and the logit gradient.
template <int OPT, bool NEED_DA, class Sync>
__device__ __forceinline__ void
bwd_small_body(const unsigned bx, Sync sync,
               float* __restrict__ W, int ldw, int N, int K, float* __restrict__ bias, float* __restrict__ s1,
               float* __restrict__ s2, float* __restrict__ bs1, float* __restrict__ bs2,
               const float* __restrict__ dZ, int ldz, const float* __restrict__ A, int lda, int B, OptParams op,
               int gate, float gscale, float* __restrict__ dst, int ldd, const float* __restrict__ logits = nullptr,
               CeTail ce = CeTail{}) {
  __shared__ float dzs[128 * kBT];                 // [row][16]
  __shared__ float red[8][kBT][32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int k = bx * 32 + lane;
  const bool kok = k < K;
  float w[kSmallBwdRows], m1[kSmallBwdRows], m2[OPT == 2 ? kSmallBwdRows : 1];
#pragma unroll
  for (int i = 0; i < kSmallBwdRows; ++i) {
    const int n = warp + 8 * i;
    const bool ok = kok && n < N;
    const size_t o = (size_t)(ok ? n : 0) * ldw + (ok ? k : 0);
    w[i] = ok ? W[o] : 0.f;
    m1[i] = ok ? s1[o] : 0.f;
    if (OPT == 2) m2[i] = ok ? s2[o] : 0.f;
  }
  sync();            // W / optimizer state are never written by the kernel (phase) just before this one;
                     // dZ (and A) are
  if (ce.labels && logits) {
    // cross-entropy in place: warp w takes rows 2w, 2w+1 (16 lanes each), the logit gradient lands in
    // dzs[class][row]; rows beyond B stay zero
    for (int idx = tid; idx < N * kBT; idx += 256) dzs[idx] = 0.f;
    __syncthreads();
    __shared__ float rl_s[kBT];
    const int gl = lane & 15, r = warp * 2 + (lane >> 4);
    const bool live = r < ce.rows;
    float zc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = gl + 16 * i;
      zc[i] = (live && c < ce.C) ? __ldcg(logits + (size_t)r * ce.C + c) : -INFINITY;
    }
    const int64_t lab = live ? ce.labels[r] : 0;
    // ce_row_reg writes (softmax - onehot)/rows for class c to dzr[c]: aim it at column r of dzs
    float dzr[8];
    const float l = ce_row_reg_vals<8, 16>(zc, lab, live ? ce.C : 0, 1.f / (float)ce.rows, dzr, gl);
    if (live) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int c = gl + 16 * i;
        if (c < ce.C) {
          dzs[c * kBT + r] = dzr[i];
          if (bx == 0 && ce.dz) ce.dz[(size_t)r * ce.C + c] = dzr[i];
        }
      }
      if (gl == 0) rl_s[r] = l;
    }
    __syncthreads();
    if (bx == 0 && warp == 0) {
      float sl = 0.f;
      for (int q = lane; q < ce.rows; q += 32) sl += rl_s[q];
      sl = warp_sum(sl);
      if (lane == 0 && ce.loss) *ce.loss = sl / (float)ce.rows;
    }
  } else {
    for (int idx = tid; idx < N * kBT; idx += 256) {
      const int b = idx / N, j = idx - b * N;
      dzs[j * kBT + b] = b < B ? ld_dep<Sync>(dZ + (size_t)b * ldz + j) : 0.f;
    }
  }
  float a[kBT];
#pragma unroll
  for (int b = 0; b < kBT; ++b) a[b] = (kok && b < B) ? ld_dep<Sync>(A + (size_t)b * lda + k) : 0.f;
  float acc[NEED_DA ? kBT : 1];
  if (NEED_DA) {
#pragma unroll
    for (int b = 0; b < kBT; ++b) acc[b] = 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kSmallBwdRows; ++i) {
    const int n = warp + 8 * i;
    if (n < N) {
      const float4* dzr = reinterpret_cast<const float4*>(dzs + n * kBT);
      float dz[kBT];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 t = dzr[q];
        dz[q * 4 + 0] = t.x; dz[q * 4 + 1] = t.y; dz[q * 4 + 2] = t.z; dz[q * 4 + 3] = t.w;
      }
      float g = 0.f;
#pragma unroll
      for (int b = 0; b < kBT; ++b) {
        g = __fmaf_rn(dz[b], a[b], g);
        if (NEED_DA) acc[b] = __fmaf_rn(dz[b], w[i], acc[b]);
      }
      if (kok) {
        float pw = w[i], q1 = m1[i], q2 = OPT == 2 ? m2[OPT == 2 ? i : 0] : 0.f;
        opt_update<OPT>(op, pw, g, q1, q2);
        const size_t o = (size_t)n * ldw + k;
        W[o] = pw;
        s1[o] = q1;
        if (OPT == 2) s2[o] = q2;
      }
    }
  }
  if (bx == 0) {                                    // bias: db[n] = sum_b dZ[b][n]
    for (int j = tid; j < N; j += 256) {
      float db = 0.f;
#pragma unroll
      for (int b = 0; b < kBT; ++b) db += dzs[j * kBT + b];
      float p = bias[j], q1 = bs1[j], q2 = (OPT == 2) ? bs2[j] : 0.f;
      opt_update<OPT>(op, p, db, q1, q2);
      bias[j] = p;
      bs1[j] = q1;
      if (OPT == 2) bs2[j] = q2;
    }
  }
  if (NEED_DA) {
#pragma unroll
    for (int b = 0; b < kBT; ++b) red[warp][b][lane] = acc[b];
    __syncthreads();
    for (int o = tid; o < kBT * 32; o += 256) {
      const int l = o & 31, b = o >> 5;
      float t = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) t += red[r][b][l];
      const int kk = bx * 32 + l;
      if (kk < K && b < B) {
        if (gate) t = ld_dep<Sync>(A + (size_t)b * lda + kk) > 0.f ? t * gscale : 0.f;
        dst[(size_t)b * ldd + kk] = t;
      }
    }
  }
}

template <int OPT, bool NEED_DA>
__global__ void __launch_bounds__(256)
bwd_small_kernel(float* __restrict__ W, int ldw, int N, int K, float* __restrict__ bias, float* __restrict__ s1,
                 float* __restrict__ s2, float* __restrict__ bs1, float* __restrict__ bs2,
                 const float* __restrict__ dZ, int ldz, const float* __restrict__ A, int lda, int B, OptParams op,
                 int gate, float gscale, float* __restrict__ dst, int ldd) {
  bwd_small_body<OPT, NEED_DA>(blockIdx.x, PdlSync{}, W, ldw, N, K, bias, s1, s2, bs1, bs2, dZ, ldz, A, lda, B, op, gate, gscale,
                               dst, ldd);
}

// dst[b][k] = (sum_s PA[s][b][k]) * gate, gate = (H[b][k] > 0 ? scale : 0) for hidden layers
// (H is the post-ReLU, post-dropout activation so H>0 <=> ReLU active and kept), 1 at the cut.
// dst_table != nullptr (tensor-parallel head): column k goes to dst_table[owner of column k] instead of
// dst - each consumer GPU gets its column range pushed into its own memory over NVLink.
template <class Sync>
__device__ __forceinline__ void
bwd_epilogue_body(const unsigned bx, const unsigned gdx, Sync sync, const float* __restrict__ PA, int nsplit, int B, int K,
                  const float* __restrict__ H, int ldh, float scale, float* __restrict__ dst, int ldd,
                  float* const* __restrict__ dst_table, int cols_per_dst, int first_cols = -1) {
  sync();
  const size_t total = (size_t)B * K;
  const size_t stride = (size_t)gdx * blockDim.x;
  size_t idx = bx * (size_t)blockDim.x + threadIdx.x;
  for (; idx < total; idx += stride) {
    const int b = (int)(idx / K), k = (int)(idx - (size_t)b * K);
    const float hv = H ? ld_dep<Sync>(H + (size_t)b * ldh + k) : 1.f;   // requested before the partials
    float s = 0.f;
    const float* pp = PA + (size_t)b * K + k;
    const size_t pstride = (size_t)kBT * K;
    int p = 0;
    for (; p + 4 <= nsplit; p += 4) {            // four independent loads in flight, fixed add order
      const float p0 = ld_dep<Sync>(pp + (size_t)p * pstride), p1 = ld_dep<Sync>(pp + (size_t)(p + 1) * pstride);
      const float p2 = ld_dep<Sync>(pp + (size_t)(p + 2) * pstride), p3 = ld_dep<Sync>(pp + (size_t)(p + 3) * pstride);
      s += p0; s += p1; s += p2; s += p3;
    }
    for (; p < nsplit; ++p) s += ld_dep<Sync>(pp + (size_t)p * pstride);
    if (H) s = hv > 0.f ? s * scale : 0.f;
    // first_cols >= 0: consumer 0 owns the first `first_cols` columns (possibly none), the others
    // cols_per_dst each after that (the tensor-parallel head keeps a smaller slice for its own GPU)
    float* d = dst_table ? dst_table[first_cols < 0 ? k / cols_per_dst : (k < first_cols ? 0 : 1 + (k - first_cols) / cols_per_dst)]
                         : dst;
    d[(size_t)b * ldd + k] = s;
  }
}

__global__ void bwd_epilogue_kernel(const float* __restrict__ PA, int nsplit, int B, int K,
                                    const float* __restrict__ H, int ldh, float scale, float* __restrict__ dst,
                                    int ldd, float* const* __restrict__ dst_table, int cols_per_dst, int first_cols) {
  bwd_epilogue_body(blockIdx.x, gridDim.x, PdlSync{}, PA, nsplit, B, K, H, ldh, scale, dst, ldd, dst_table, cols_per_dst,
                    first_cols);
}

// dst[b][k..k+3] = (sum_s P[s][b][k..]) gated by H (nullable) - the reduction behind a split-K dX GEMM
// (K % 4 == 0; slices are dense [B][K] slabs `slab` floats apart; fixed add order)
__global__ void splitk_gate_kernel(const float* __restrict__ P, int nsplit, size_t slab, int B, int K,
                                   const float* __restrict__ H, int ldh, float scale, float* __restrict__ dst, int ldd) {
  pdl_wait();
  pdl_launch();
  const size_t total4 = (size_t)B * K / 4;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x) {
    const size_t e = i * 4;
    const int b = (int)(e / K), k = (int)(e - (size_t)b * K);
    float4 hv = make_float4(1.f, 1.f, 1.f, 1.f);
    if (H) hv = *reinterpret_cast<const float4*>(H + (size_t)b * ldh + k);
    float4 acc = __ldcg(reinterpret_cast<const float4*>(P + e));
    for (int s = 1; s < nsplit; ++s) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(P + s * slab + e));
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    if (H) {
      acc.x = hv.x > 0.f ? acc.x * scale : 0.f; acc.y = hv.y > 0.f ? acc.y * scale : 0.f;
      acc.z = hv.z > 0.f ? acc.z * scale : 0.f; acc.w = hv.w > 0.f ? acc.w * scale : 0.f;
    }
    *reinterpret_cast<float4*>(dst + (size_t)b * ldd + k) = acc;
  }
}

// dst = src * (out > 0)  (ReLU on the last layer, U-shape model2)
__global__ void relu_gate_kernel(const float* __restrict__ src, const float* __restrict__ out, size_t n,
                                 float* __restrict__ dst) {
  pdl_wait();
  pdl_launch();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = out[i] > 0.f ? src[i] : 0.f;
}

// ------------------------------------------------------------------------------------------
// General-batch fallback (B > 16 in the backward, B > 64 in the forward) until gemm_tc.cuh
// covers those shapes: a plain 64x64x16 register-tiled SGEMM with arbitrary element strides,
// C[m][n] = sum_k A(m,k) * Bm(k,n), plus elementwise optimizer / bias-gradient kernels.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sgemm_strided_kernel(const float* __restrict__ A, long long sam, long long sak,
                                                            const float* __restrict__ Bm, long long sbk, long long sbn,
                                                            float* __restrict__ C, long long ldc, int M, int N, int K) {
  pdl_wait();
  pdl_launch();
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      int mm, kk;
      if (sak == 1) { kk = i & 15; mm = i >> 4; } else { mm = i & 63; kk = i >> 6; }
      const int gm = m0 + mm, gk = k0 + kk;
      As[kk][mm] = (gm < M && gk < K) ? A[gm * sam + gk * sak] : 0.f;
      int nn, k2;
      if (sbk == 1) { k2 = i & 15; nn = i >> 4; } else { nn = i & 63; k2 = i >> 6; }
      const int gn = n0 + nn, gk2 = k0 + k2;
      Bs[k2][nn] = (gn < N && gk2 < K) ? Bm[gk2 * sbk + gn * sbn] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = As[kk][ty * 4 + i]; bv[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = __fmaf_rn(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gm = m0 + ty * 4 + i, gn = n0 + tx * 4 + j;
      if (gm < M && gn < N) C[gm * ldc + gn] = acc[i][j];
    }
}

template <int OPT>
__global__ void opt_elementwise_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ s1,
                                       float* __restrict__ s2, size_t n, OptParams op) {
  pdl_wait();
  pdl_launch();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float pv = p[i], a = s1[i], b = (OPT == 2) ? s2[i] : 0.f;
    opt_update<OPT>(op, pv, g[i], a, b);
    p[i] = pv;
    s1[i] = a;
    if (OPT == 2) s2[i] = b;
  }
}

// db[n] = sum_b dZ[b][n] (row order), then the optimizer on the bias.
template <int OPT>
__global__ void bias_grad_update_kernel(const float* __restrict__ dZ, int ldz, int B, int N, float* __restrict__ bias,
                                        float* __restrict__ s1, float* __restrict__ s2, OptParams op) {
  pdl_wait();
  pdl_launch();
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float db = 0.f;
  for (int b = 0; b < B; ++b) db += dZ[(size_t)b * ldz + n];
  float pv = bias[n], a = s1[n], q = (OPT == 2) ? s2[n] : 0.f;
  opt_update<OPT>(op, pv, db, a, q);
  bias[n] = pv;
  s1[n] = a;
  if (OPT == 2) s2[n] = q;
}

}  // namespace sn
