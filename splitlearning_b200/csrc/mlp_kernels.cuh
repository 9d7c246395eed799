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
__global__ void __launch_bounds__(256, 1)
fwd_skinny_kernel(const float* __restrict__ X, int ldx, int B, const float* __restrict__ W, int ldw, int N,
                  int K, int kc, float* __restrict__ P, int Bpad, int ksplit, int prefetch_early) {
  extern __shared__ float4 smem4[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ks = blockIdx.y, k0 = ks * kc;
  const int klen = min(kc, K - k0);
  const int iters = (klen + 127) >> 7;
  const int b0 = blockIdx.z * kBT, nb = min(kBT, B - b0);
  const int kc4 = kc >> 2;
  float4* wring = smem4 + kBT * kc4 + warp * (kFwdRing * 8 * 32);   // after the X chunk
#ifdef SN_TIMING
  long long tstamp[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long t_red = 0;
#endif
  SN_T(0);

  const int groups = (N + 7) >> 3;
  const long long T = (long long)groups * iters;
  const int gw = blockIdx.x * 8 + warp, nw = gridDim.x * 8;
  const int t0 = (int)(T * gw / nw), t1 = (int)(T * (gw + 1) / nw);

  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool ktail = (klen & 127) != 0;      // only then can a lane sit beyond the end of K

  // W stream: (gi, iti) = coordinates of the next tile to issue
  int gi = t0 / iters, iti = t0 - gi * iters, ti = t0;
  auto issue = [&](int slot) {
    if (ti < t1) {
      const int k = iti * 128 + lane * 4;
      if (k < klen) {
        const float* src = W + (size_t)(gi * 8) * ldw + k0 + k;
        if (gi * 8 + 8 <= N) {
#pragma unroll
          for (int i = 0; i < 8; ++i) cp_async16(wring + (slot * 8 + i) * 32 + lane, src + (size_t)i * ldw);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            cp_async16(wring + (slot * 8 + i) * 32 + lane, src + (size_t)min(i, N - 1 - gi * 8) * ldw);
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
  pdl_wait();
  pdl_launch();
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
    constexpr int XI = 2;                                   // k4 items per thread per pass
    for (int kbase = 0; kbase < kc4; kbase += 256 * XI) {
      float4 v[XI][4][4];
#pragma unroll
      for (int u = 0; u < XI; ++u) {
        const int k4 = kbase + u * 256 + tid;
#pragma unroll
        for (int bq = 0; bq < 4; ++bq)
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const int b = bq * 4 + r;
            v[u][bq][r] = zero;
            if (k4 < kc4 && k4 * 4 < klen && b < nb)
              v[u][bq][r] = __ldg(reinterpret_cast<const float4*>(X + (size_t)(b0 + b) * ldx + k0) + k4);
          }
      }
#pragma unroll
      for (int u = 0; u < XI; ++u) {
        const int k4 = kbase + u * 256 + tid;
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

  float acc[8 * kBT];
#pragma unroll
  for (int i = 0; i < 8 * kBT; ++i) acc[i] = 0.f;
  int g = t0 / iters, it = t0 - g * iters;
  bool has_first = (it == 0);
  int slot = 0;
  const int sw = (lane >> 1) & 3;
  const size_t slot_stride = (size_t)ksplit * Bpad * N;
  float* Pbase = P + ((size_t)ks * Bpad + b0) * N;

  for (int t = t0; t < t1; ++t) {
    issue((slot + kFwdRing - 1) & (kFwdRing - 1));
    cp_async_wait<kFwdRing - 1>();
    float4 wc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) wc[i] = wring[(slot * 8 + i) * 32 + lane];
    if (ktail && it == iters - 1 && it * 128 + lane * 4 >= klen) {   // nothing was copied for this lane
#pragma unroll
      for (int i = 0; i < 8; ++i) wc[i] = zero;
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
        for (int i = 0; i < 8; ++i) {
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
      butterfly_reduce<8 * kBT>(acc, lane);        // lane ends with row lane>>2, batch (lane&3)*4+j
      const int row = g * 8 + (lane >> 2);
      if (row < N) {
        float* dst = Pbase + (has_first ? 0 : slot_stride) + row;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int b = (lane & 3) * 4 + j;
          if (b < nb) {
            dst[(size_t)b * N] = acc[j];
            if (has_first && group_done) dst[slot_stride + (size_t)b * N] = 0.f;
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 8 * kBT; ++i) acc[i] = 0.f;
#ifdef SN_TIMING
      t_red += clock64() - tr0;
#endif
    }
    if (++it == iters) { it = 0; ++g; has_first = true; }
  }
  cp_async_wait<0>();
  SN_T(4);
#ifdef SN_TIMING
  if (lane == 0 && blockIdx.z == 0) {
    long long* o = g_fwd_timing + ((blockIdx.y * gridDim.x + blockIdx.x) * 8 + warp) * 8;
    o[0] = tstamp[1] - tstamp[0]; o[1] = tstamp[2] - tstamp[1]; o[2] = tstamp[3] - tstamp[2];
    o[3] = tstamp[4] - tstamp[3]; o[4] = t_red; o[5] = t1 - t0; o[6] = tstamp[0]; o[7] = tstamp[4];
  }
#endif
}

// H[b][n] = act(sum_ks P[ks][b][n] + bias[n]) * dropout.   One thread per element.
// mask (nullable) is [B][N] keep-bytes with row stride N; otherwise, when keep_prob < 1, Philox.
__global__ void fwd_epilogue_kernel(const float* __restrict__ P, int ksplit, int Bpad, int B, int N,
                                    const float* __restrict__ bias, int relu, const uint8_t* __restrict__ mask,
                                    float keep_prob, float scale, uint64_t seed, uint64_t offset, int layer,
                                    float* __restrict__ H, int ldh, float* __restrict__ out2) {
  pdl_wait();
  pdl_launch();
  const size_t total = (size_t)B * N;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(idx / N), n = (int)(idx - (size_t)b * N);
    float s = 0.f;
    const float* pp = P + (size_t)b * N + n;
    const size_t pstride = (size_t)Bpad * N;
    int ks = 0;
    for (; ks + 4 <= ksplit; ks += 4) {          // four independent loads in flight, fixed add order
      const float p0 = pp[(size_t)ks * pstride], p1 = pp[(size_t)(ks + 1) * pstride];
      const float p2 = pp[(size_t)(ks + 2) * pstride], p3 = pp[(size_t)(ks + 3) * pstride];
      s += p0; s += p1; s += p2; s += p3;
    }
    for (; ks < ksplit; ++ks) s += pp[(size_t)ks * pstride];
    s += bias[n];
    if (relu) s = fmaxf(s, 0.f);
    if (mask) {
      s = mask[idx] ? s * scale : 0.f;
    } else if (keep_prob < 1.f) {
      s = philox_keep(seed, offset, layer, idx, keep_prob) ? s * scale : 0.f;
    }
    H[(size_t)b * ldh + n] = s;
    if (out2) out2[idx] = s;
  }
}

// Tiny layers (fc3 100x1000, the U-shape tail 10x100): the weight matrix is L2-resident and the
// launch is pure latency, so the split-K machinery above costs more than it saves.  One warp per
// output row, lanes stride k, 16 batch accumulators per lane, one 16-value butterfly, and the
// bias / ReLU / dropout epilogue fused: one launch instead of two, no partial sums.
__global__ void __launch_bounds__(256)
fwd_small_kernel(const float* __restrict__ X, int ldx, int B, const float* __restrict__ W, int ldw, int N, int K,
                 const float* __restrict__ bias, int relu, const uint8_t* __restrict__ mask, float keep_prob,
                 float scale, uint64_t seed, uint64_t offset, int layer, float* __restrict__ H, int ldh,
                 float* __restrict__ out2, int prefetch_early) {
  extern __shared__ float4 smem4[];                 // X tile [16][K4] float4 (K padded to 4)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = blockIdx.x * 8 + warp;
  const int b0 = blockIdx.y * kBT, nb = min(kBT, B - b0);
  const int K4 = (K + 3) >> 2;
  const bool vec = (K & 3) == 0;
  // when the host says the previous kernel cannot be writing W (prefetch_early), this warp's W
  // row is fetched before the PDL wait; otherwise right after it
  constexpr int WPF = 8;                            // float4 per lane held in registers (K <= 1024)
  float4 wreg[WPF];
  const bool row_ok = n < N;
  const float* wr = W + (size_t)(row_ok ? n : 0) * ldw;
  if (!prefetch_early) { pdl_wait(); pdl_launch(); }
  if (vec) {
#pragma unroll
    for (int q = 0; q < WPF; ++q) {
      const int k4 = q * 32 + lane;
      wreg[q] = (row_ok && k4 < K4) ? __ldg(reinterpret_cast<const float4*>(wr) + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  if (prefetch_early) { pdl_wait(); pdl_launch(); }
  // X -> smem, all loads of a thread in flight together (one round trip)
  for (int idx = tid; idx < kBT * K4; idx += 256) {
    const int b = idx / K4, k4 = idx - b * K4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b < nb) {
      const float* xr = X + (size_t)(b0 + b) * ldx;
      if (vec) v = __ldg(reinterpret_cast<const float4*>(xr) + k4);
      else {
        const int k = k4 * 4;
        v.x = k < K ? __ldg(xr + k) : 0.f;
        v.y = k + 1 < K ? __ldg(xr + k + 1) : 0.f;
        v.z = k + 2 < K ? __ldg(xr + k + 2) : 0.f;
        v.w = k + 3 < K ? __ldg(xr + k + 3) : 0.f;
      }
    }
    smem4[idx] = v;
  }
  __syncthreads();
  if (!row_ok) return;
  float acc[kBT];
#pragma unroll
  for (int b = 0; b < kBT; ++b) acc[b] = 0.f;
  for (int q = 0; q * 32 < K4; ++q) {
    const int k4 = q * 32 + lane;
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k4 < K4) {
      if (vec) w = q < WPF ? wreg[q < WPF ? q : 0] : __ldg(reinterpret_cast<const float4*>(wr) + k4);
      else {
        const int k = k4 * 4;
        w.x = k < K ? __ldg(wr + k) : 0.f;
        w.y = k + 1 < K ? __ldg(wr + k + 1) : 0.f;
        w.z = k + 2 < K ? __ldg(wr + k + 2) : 0.f;
        w.w = k + 3 < K ? __ldg(wr + k + 3) : 0.f;
      }
#pragma unroll
      for (int b = 0; b < kBT; ++b) {
        const float4 x = smem4[b * K4 + k4];
        acc[b] = __fmaf_rn(w.x, x.x, acc[b]);
        acc[b] = __fmaf_rn(w.y, x.y, acc[b]);
        acc[b] = __fmaf_rn(w.z, x.z, acc[b]);
        acc[b] = __fmaf_rn(w.w, x.w, acc[b]);
      }
    }
  }
  butterfly_reduce<kBT>(acc, lane);             // lanes 2b and 2b+1 both end with the sum for batch row b
  const int b = lane >> 1;
  if ((lane & 1) || b >= nb) return;
  float s = acc[0] + bias[n];
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

// ------------------------------------------------------------------------------------------
// Cross-entropy: one warp per row; loss = mean_b(lse - z[y]); dlogits = (softmax - onehot)/B.
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
    const float* zr = z + (size_t)b * ldz;
    float m = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = zr[c];
      if (v > m) { m = v; arg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
    }
    // sum of exp(z-m) WITHOUT the max term (which is exactly 1): log1p keeps full precision when
    // one class dominates, where log(1+eps) computed from a rounded sum would lose it
    float s1 = 0.f;
    for (int c = lane; c < C; c += 32)
      if (c != arg) s1 += expf(zr[c] - m);
    s1 = warp_sum(s1);
    const float s = 1.f + s1;
    const int64_t lab = y[b];
    const bool ok = lab >= 0 && lab < C;
    // -log_softmax(z)[y] = log(sum exp(z-m)) - (z[y]-m): no cancellation against the row max
    if (lane == 0) row_loss[b] = ok ? (log1pf(s1) - (zr[lab] - m)) : NAN;   // torch raises on a bad target
    if (dz) {
      const float invB = 1.f / (float)B, invs = 1.f / s;
      for (int c = lane; c < C; c += 32) {
        float p = expf(zr[c] - m) * invs;
        if (c == lab) p -= 1.f;
        dz[(size_t)b * lddz + c] = p * invB;
      }
    }
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
constexpr int kRing = 8;   // rows in flight per thread (W + optimizer state): ~56-84 KB per SM

// FUSE_FWD (SISA server epochs): the NEXT batch's activations are already known, so the same
// traversal also produces the next step's forward through this layer with the freshly updated
// weights: per row, each lane dots its 4 new weights with Xn[16][4], a 16-value butterfly sums
// across the warp's 128-wide k slice, and the per-slice partials PF[slice][b][n] are summed by
// fwd_epilogue_kernel at the start of the next step.  W is then read once per step instead of
// twice (28*P -> 24*P bytes with Adam).
template <int OPT, bool NEED_DA, bool FUSE_FWD = false>
__global__ void __launch_bounds__(256, 1)
bwd_update_kernel(float* __restrict__ W, int ldw, int N, int K, float* __restrict__ bias, float* __restrict__ s1,
                  float* __restrict__ s2, float* __restrict__ bs1, float* __restrict__ bs2,
                  const float* __restrict__ dZ, int ldz, const float* __restrict__ A, int lda, int B,
                  float* __restrict__ PA, int kw, int rows_per_cta, OptParams op,
                  const float* __restrict__ Xn = nullptr, int Bn = 0, float* __restrict__ PF = nullptr) {
  extern __shared__ float4 smem4[];
  float* dzs = reinterpret_cast<float*>(smem4);                  // [rows_per_cta][16]
  // per-thread ring of the next kRing rows' float4s: [array][slot][thread]; a thread only ever
  // reads what it copied itself, so cp.async.wait_group is the only synchronisation needed
  float4* ring = smem4 + ((rows_per_cta * kBT + 3) >> 2);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wk = warp % kw, wr = warp / kw, rw = 8 / kw;
  const int nb0 = blockIdx.y * rows_per_cta, nb1 = min(N, nb0 + rows_per_cta);
  const int nrows = nb1 - nb0;
  const int k0 = (blockIdx.x * kw + wk) * 128 + lane * 4;
  const bool kok = k0 < K;

  auto issue = [&](int slot, int row) {
    if (kok && row < nb1) {
      const size_t o = (size_t)row * ldw + k0;
      cp_async16(ring + (0 * kRing + slot) * 256 + tid, W + o);
      cp_async16(ring + (1 * kRing + slot) * 256 + tid, s1 + o);
      if (OPT == 2) cp_async16(ring + (2 * kRing + slot) * 256 + tid, s2 + o);
    }
    cp_async_commit();
  };
  int n = nb0 + wr;
#pragma unroll
  for (int d = 0; d < kRing - 1; ++d) issue(d, n + d * rw);      // start streaming before anything else
  pdl_wait();        // W / optimizer state are never written by the kernel just before this one;
  pdl_launch();      // dZ (and A) are, so everything below waits for it

  {  // dZ tile, transposed to [row][16]; loads batched ahead of their stores (latency, not bandwidth)
    constexpr int SB = 8;
    const int total = nrows * kBT;
    for (int base_idx = tid; base_idx < total; base_idx += 256 * SB) {
      float v[SB];
#pragma unroll
      for (int q = 0; q < SB; ++q) {
        const int idx = base_idx + q * 256;
        const int b = idx / nrows, j = idx - b * nrows;
        v[q] = (idx < total && b < B) ? __ldg(dZ + (size_t)b * ldz + nb0 + j) : 0.f;
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
  float4 a[kBT];
#pragma unroll
  for (int b = 0; b < kBT; ++b)
    a[b] = (kok && b < B) ? __ldg(reinterpret_cast<const float4*>(A + (size_t)b * lda + k0)) : make_float4(0.f, 0.f, 0.f, 0.f);
  float4 acc[NEED_DA ? kBT : 1];
  if (NEED_DA) {
#pragma unroll
    for (int b = 0; b < kBT; ++b) acc[b] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float4 xn[FUSE_FWD ? kBT : 1];
  if (FUSE_FWD) {
#pragma unroll
    for (int b = 0; b < kBT; ++b)
      xn[b] = (kok && b < Bn) ? __ldg(reinterpret_cast<const float4*>(Xn + (size_t)b * lda + k0)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int kslice = blockIdx.x * kw + wk;
  const bool slice_ok = kslice * 128 < K;
  __syncthreads();

  int slot = 0;
  for (; n < nb1; n += rw) {
    issue((slot + kRing - 1) & (kRing - 1), n + (kRing - 1) * rw);
    cp_async_wait<kRing - 1>();
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f), m1 = w, m2 = w;
    if (kok) {
      w = ring[(0 * kRing + slot) * 256 + tid];
      m1 = ring[(1 * kRing + slot) * 256 + tid];
      if (OPT == 2) m2 = ring[(2 * kRing + slot) * 256 + tid];
    }
    slot = (slot + 1) & (kRing - 1);
    const float4* dzr = reinterpret_cast<const float4*>(dzs + (n - nb0) * kBT);
    float dz[kBT];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 t = dzr[q];
      dz[q * 4 + 0] = t.x; dz[q * 4 + 1] = t.y; dz[q * 4 + 2] = t.z; dz[q * 4 + 3] = t.w;
    }
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int b = 0; b < kBT; ++b) {
      g.x = __fmaf_rn(dz[b], a[b].x, g.x);
      g.y = __fmaf_rn(dz[b], a[b].y, g.y);
      g.z = __fmaf_rn(dz[b], a[b].z, g.z);
      g.w = __fmaf_rn(dz[b], a[b].w, g.w);
      if (NEED_DA) {
        acc[b].x = __fmaf_rn(dz[b], w.x, acc[b].x);
        acc[b].y = __fmaf_rn(dz[b], w.y, acc[b].y);
        acc[b].z = __fmaf_rn(dz[b], w.z, acc[b].z);
        acc[b].w = __fmaf_rn(dz[b], w.w, acc[b].w);
      }
    }
    if (kok) {
      opt_update<OPT>(op, w.x, g.x, m1.x, m2.x);
      opt_update<OPT>(op, w.y, g.y, m1.y, m2.y);
      opt_update<OPT>(op, w.z, g.z, m1.z, m2.z);
      opt_update<OPT>(op, w.w, g.w, m1.w, m2.w);
      const size_t o = (size_t)n * ldw + k0;
      *reinterpret_cast<float4*>(W + o) = w;
      __stcs(reinterpret_cast<float4*>(s1 + o), m1);
      if (OPT == 2) __stcs(reinterpret_cast<float4*>(s2 + o), m2);
    }
    if (FUSE_FWD) {                                   // next step's forward with the updated row
      float y[kBT];
#pragma unroll
      for (int b = 0; b < kBT; ++b) {
        float t = w.x * xn[b].x;
        t = __fmaf_rn(w.y, xn[b].y, t);
        t = __fmaf_rn(w.z, xn[b].z, t);
        y[b] = __fmaf_rn(w.w, xn[b].w, t);
      }
      butterfly_reduce<kBT>(y, lane);               // lanes 2b, 2b+1 hold batch row b
      if (slice_ok && !(lane & 1)) PF[((size_t)kslice * kBT + (lane >> 1)) * N + n] = y[0];
    }
  }
  cp_async_wait<0>();

  // bias: db[n] = sum_b dZ[b][n].  One thread per row AFTER the streaming loop (a per-row
  // load->update->store chain inside the loop made the k-group-0 CTAs latency-bound and the
  // whole grid wait for them).
  if (blockIdx.x == 0) {
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
    for (int b = 0; b < kBT; ++b) red[(warp * kBT + b) * 32 + lane] = acc[b];
    __syncthreads();
    for (int o = tid; o < kw * kBT * 32; o += 256) {
      const int l = o & 31, b = (o >> 5) % kBT, s = o / (32 * kBT);
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < rw; ++r) {
        const float4 u = red[((r * kw + s) * kBT + b) * 32 + l];
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      const int k = (blockIdx.x * kw + s) * 128 + l * 4;
      if (k < K && b < B) *reinterpret_cast<float4*>(PA + ((size_t)blockIdx.y * kBT + b) * K + k) = t;
    }
  }
}

// dst[b][k] = (sum_s PA[s][b][k]) * gate, gate = (H[b][k] > 0 ? scale : 0) for hidden layers
// (H is the post-ReLU, post-dropout activation so H>0 <=> ReLU active and kept), 1 at the cut.
__global__ void bwd_epilogue_kernel(const float* __restrict__ PA, int nsplit, int B, int K,
                                    const float* __restrict__ H, int ldh, float scale, float* __restrict__ dst,
                                    int ldd) {
  pdl_wait();
  pdl_launch();
  const size_t total = (size_t)B * K;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(idx / K), k = (int)(idx - (size_t)b * K);
    float s = 0.f;
    const float* pp = PA + (size_t)b * K + k;
    const size_t pstride = (size_t)kBT * K;
    int p = 0;
    for (; p + 4 <= nsplit; p += 4) {
      const float p0 = pp[(size_t)p * pstride], p1 = pp[(size_t)(p + 1) * pstride];
      const float p2 = pp[(size_t)(p + 2) * pstride], p3 = pp[(size_t)(p + 3) * pstride];
      s += p0; s += p1; s += p2; s += p3;
    }
    for (; p < nsplit; ++p) s += pp[(size_t)p * pstride];
    if (H) s = H[(size_t)b * ldh + k] > 0.f ? s * scale : 0.f;
    dst[(size_t)b * ldd + k] = s;
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
