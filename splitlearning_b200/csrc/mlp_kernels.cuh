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

constexpr int kFwdRing = 3;   // W iterations (8 rows x 128 k each) in flight per warp

// One group of R rows of W against the staged X chunk.  W is streamed global -> shared with
// cp.async into this warp's private ring (no registers held while in flight), X is read from the
// CTA's staged chunk one batch row ahead of the FMAs that use it.
template <int R>
__device__ __forceinline__ void fwd_rows(const float* __restrict__ Wr, int ldw, int klen,
                                         const float4* __restrict__ Xs4, int kc4, int lane, int nb,
                                         float* __restrict__ Pout, int N, float4* __restrict__ wring) {
  float acc[R * kBT];
#pragma unroll
  for (int i = 0; i < R * kBT; ++i) acc[i] = 0.f;
  const int iters = (klen + 127) >> 7;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  auto issue = [&](int it, int slot) {
    const int k = it * 128 + lane * 4;
    if (it < iters && k < klen) {
#pragma unroll
      for (int i = 0; i < R; ++i) cp_async16(wring + (slot * 8 + i) * 32 + lane, Wr + (size_t)i * ldw + k);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int d = 0; d < kFwdRing - 1; ++d) issue(d, d);
  int slot = 0;
  for (int it = 0; it < iters; ++it) {
    int wslot = slot + kFwdRing - 1;
    if (wslot >= kFwdRing) wslot -= kFwdRing;
    issue(it + kFwdRing - 1, wslot);
    cp_async_wait<kFwdRing - 1>();
    const bool ok = it * 128 + lane * 4 < klen;
    float4 wc[R];
#pragma unroll
    for (int i = 0; i < R; ++i) wc[i] = ok ? wring[(slot * 8 + i) * 32 + lane] : zero;
    if (++slot == kFwdRing) slot = 0;
    const float4* xs = Xs4 + it * 32 + lane;
    float4 xn = xs[0];
#pragma unroll
    for (int b = 0; b < kBT; ++b) {
      const float4 x = xn;
      if (b + 1 < kBT) xn = xs[(b + 1) * kc4];
#pragma unroll
      for (int i = 0; i < R; ++i) {
        float a = acc[i * kBT + b];
        a = __fmaf_rn(wc[i].x, x.x, a);
        a = __fmaf_rn(wc[i].y, x.y, a);
        a = __fmaf_rn(wc[i].z, x.z, a);
        a = __fmaf_rn(wc[i].w, x.w, a);
        acc[i * kBT + b] = a;
      }
    }
  }
  cp_async_wait<0>();
  constexpr int V = R * kBT;
  butterfly_reduce<V>(acc, lane);
  constexpr int CNT = V >= 32 ? V / 32 : 1;
  const int prefix = V >= 32 ? lane : (lane >> 1);
  if (V < 32 && (lane & 1)) return;
#pragma unroll
  for (int j = 0; j < CNT; ++j) {
    const int idx = prefix * CNT + j;
    const int i = idx >> 4, b = idx & 15;
    if (b < nb) Pout[(size_t)b * N + i] = acc[j];
  }
}

__global__ void __launch_bounds__(256, 1)
fwd_skinny_kernel(const float* __restrict__ X, int ldx, int B, const float* __restrict__ W, int ldw, int N,
                  int K, int kc, float* __restrict__ P, int Bpad) {
  extern __shared__ float4 smem4[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ks = blockIdx.y, k0 = ks * kc;
  const int klen = min(kc, K - k0);
  const int b0 = blockIdx.z * kBT, nb = min(kBT, B - b0);
  const int kc4 = kc >> 2;
  float4* wring = smem4 + kBT * kc4 + warp * (kFwdRing * 8 * 32);   // after the X chunk
  for (int idx = tid; idx < kBT * kc4; idx += 256) {
    const int b = idx / kc4, k4 = idx - b * kc4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b < nb && k4 * 4 < klen) v = __ldg(reinterpret_cast<const float4*>(X + (size_t)(b0 + b) * ldx + k0) + k4);
    smem4[idx] = v;
  }
  __syncthreads();
  const int gw = blockIdx.x * 8 + warp, nw = gridDim.x * 8;
  int r = (int)((long long)N * gw / nw);
  const int r1 = (int)((long long)N * (gw + 1) / nw);
  float* Pbase = P + ((size_t)ks * Bpad + b0) * N;
  while (r < r1) {
    const int rem = r1 - r;
    const float* Wr = W + (size_t)r * ldw + k0;
    if (rem >= 8) { fwd_rows<8>(Wr, ldw, klen, smem4, kc4, lane, nb, Pbase + r, N, wring); r += 8; }
    else if (rem >= 4) { fwd_rows<4>(Wr, ldw, klen, smem4, kc4, lane, nb, Pbase + r, N, wring); r += 4; }
    else if (rem >= 2) { fwd_rows<2>(Wr, ldw, klen, smem4, kc4, lane, nb, Pbase + r, N, wring); r += 2; }
    else { fwd_rows<1>(Wr, ldw, klen, smem4, kc4, lane, nb, Pbase + r, N, wring); r += 1; }
  }
}

// H[b][n] = act(sum_ks P[ks][b][n] + bias[n]) * dropout.   One thread per element.
// mask (nullable) is [B][N] keep-bytes with row stride N; otherwise, when keep_prob < 1, Philox.
__global__ void fwd_epilogue_kernel(const float* __restrict__ P, int ksplit, int Bpad, int B, int N,
                                    const float* __restrict__ bias, int relu, const uint8_t* __restrict__ mask,
                                    float keep_prob, float scale, uint64_t seed, uint64_t offset, int layer,
                                    float* __restrict__ H, int ldh, float* __restrict__ out2) {
  const size_t total = (size_t)B * N;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(idx / N), n = (int)(idx - (size_t)b * N);
    float s = 0.f;
    for (int ks = 0; ks < ksplit; ++ks) s += P[((size_t)ks * Bpad + b) * N + n];
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

// ------------------------------------------------------------------------------------------
// Cross-entropy: one warp per row; loss = mean_b(lse - z[y]); dlogits = (softmax - onehot)/B.
// Row losses go to row_loss[B]; the last CTA (ticket) sums them in row order.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ce_kernel(const float* __restrict__ z, int ldz, const int64_t* __restrict__ y,
                                                 int B, int C, float* __restrict__ dz, int lddz,
                                                 float* __restrict__ row_loss, unsigned* __restrict__ ticket,
                                                 float* __restrict__ loss_out, float* __restrict__ loss_out2) {
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

template <int OPT, bool NEED_DA>
__global__ void __launch_bounds__(256, 1)
bwd_update_kernel(float* __restrict__ W, int ldw, int N, int K, float* __restrict__ bias, float* __restrict__ s1,
                  float* __restrict__ s2, float* __restrict__ bs1, float* __restrict__ bs2,
                  const float* __restrict__ dZ, int ldz, const float* __restrict__ A, int lda, int B,
                  float* __restrict__ PA, int kw, int rows_per_cta, OptParams op) {
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

  for (int idx = tid; idx < nrows * kBT; idx += 256) {
    const int b = idx / nrows, j = idx - b * nrows;
    dzs[j * kBT + b] = (b < B) ? __ldg(dZ + (size_t)b * ldz + nb0 + j) : 0.f;
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
  const size_t total = (size_t)B * K;
  for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(idx / K), k = (int)(idx - (size_t)b * K);
    float s = 0.f;
    for (int p = 0; p < nsplit; ++p) s += PA[((size_t)p * kBT + b) * K + k];
    if (H) s = H[(size_t)b * ldh + k] > 0.f ? s * scale : 0.f;
    dst[(size_t)b * ldd + k] = s;
  }
}

// dst = src * (out > 0)  (ReLU on the last layer, U-shape model2)
__global__ void relu_gate_kernel(const float* __restrict__ src, const float* __restrict__ out, size_t n,
                                 float* __restrict__ dst) {
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
