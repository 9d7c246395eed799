// Alice's client block: conv3x3(1->32)+bias -> ReLU -> maxpool2x2 -> flatten, and its weight
// gradient fused with the 320-parameter optimizer update.  K=9 per output is far too thin for
// tensor cores: these are CUDA-core FMA kernels bounded by launch latency at the reference's
// batch sizes and by HBM (21.6 KB written per sample) at large ones.
// Reference: models.py:5-30 (model1 / model1_sisa).
#pragma once
#include "common.cuh"

namespace sn {

constexpr int kConvCG = 8;    // channels per CTA (4 channel groups)
constexpr int kPooled = 169;  // 13*13

// The four pre-pool conv values of pooled position (i,j), channel weights w[9], in a FIXED
// FMA order so that forward and backward recompute bit-identical numbers.
__device__ __forceinline__ void conv4(const float* __restrict__ img, const float* __restrict__ w, float bias,
                                      int i, int j, float v[4]) {
  float p[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) p[r][c] = img[(2 * i + r) * 28 + 2 * j + c];
#pragma unroll
  for (int d = 0; d < 4; ++d) {
    const int di = d >> 1, dj = d & 1;
    float a = bias;
#pragma unroll
    for (int u = 0; u < 3; ++u)
#pragma unroll
      for (int x = 0; x < 3; ++x) a = __fmaf_rn(w[u * 3 + x], p[di + u][dj + x], a);
    v[d] = a;
  }
}

// grid (samples, 4 channel groups), 256 threads.  act index = b*ld_act + c*169 + i*13 + j
// (ld_act = 5408 for a dense [B,5408] block; 5408*K when the rows are one Alice's slot of the
// --concat server's [rows, 5408*K] cache).
// x_mirror / y_src / y_dst (nullable; host-input form, sn_alice_forward_host): x and y_src are PINNED HOST
// memory that the kernel reads over PCIe itself - in its prologue, before the dependency wait, so the
// transfer overlaps the previous kernel - and it leaves device copies (x_mirror [B,784], y_dst [B]) for
// the kernels that need the batch again (the conv backward, the server's loss).  No side stream, no
// events, no break in the launch chain.
__global__ void __launch_bounds__(256) conv_pool_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                            const float* __restrict__ bias, int B,
                                                            float* __restrict__ act, size_t ld_act,
                                                            float* __restrict__ x_mirror,
                                                            const long long* __restrict__ y_src,
                                                            long long* __restrict__ y_dst) {
  __shared__ float img[28 * 28];
  __shared__ float ws[kConvCG * 9];
  __shared__ float bs[kConvCG];
  const int cg = blockIdx.y;
  // x is an input (never written by a kernel of the chain): the first image is staged before the
  // dependency wait.  The conv weights are NOT safe to touch early - the previous step's
  // conv_pool_bwd_kernel, which updates them, can be the launch right before this one.
  if (blockIdx.x < B)
    for (int t = threadIdx.x; t < 196; t += 256) {
      const float4 v = reinterpret_cast<const float4*>(x + (size_t)blockIdx.x * 784)[t];
      reinterpret_cast<float4*>(img)[t] = v;
      if (x_mirror && cg == 0) reinterpret_cast<float4*>(x_mirror + (size_t)blockIdx.x * 784)[t] = v;
    }
  if (y_dst && blockIdx.x == 0 && cg == 0)
    for (int t = threadIdx.x; t < B; t += 256) y_dst[t] = y_src[t];
  pdl_wait();
  pdl_launch();
  if (threadIdx.x < kConvCG * 9) ws[threadIdx.x] = w[cg * kConvCG * 9 + threadIdx.x];
  if (threadIdx.x < kConvCG) bs[threadIdx.x] = bias[cg * kConvCG + threadIdx.x];
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    if (b != blockIdx.x) {
      __syncthreads();
      for (int t = threadIdx.x; t < 196; t += 256) {
        const float4 v = reinterpret_cast<const float4*>(x + (size_t)b * 784)[t];
        reinterpret_cast<float4*>(img)[t] = v;
        if (x_mirror && cg == 0) reinterpret_cast<float4*>(x_mirror + (size_t)b * 784)[t] = v;
      }
    }
    __syncthreads();
    for (int o = threadIdx.x; o < kConvCG * kPooled; o += 256) {
      const int c = o / kPooled, p = o - c * kPooled;
      const int i = p / 13, j = p - i * 13;
      float v[4];
      conv4(img, ws + c * 9, bs[c], i, j, v);
      float m = fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
      act[(size_t)b * ld_act + (size_t)cg * kConvCG * kPooled + o] = fmaxf(m, 0.f);
    }
  }
}

// The same block over a WHOLE SHARD (SISA's activation-cache fill: one frozen Alice over all her
// samples, data_entities_vanilla_sisa.py:196-211): a streaming kernel, bounded by the FP32 pipe
// (36 FMA per output) and by the 21.6 KB it writes per sample, where the kernel above is a latency
// kernel for 16 samples.  Persistent grid; a CTA takes one image per iteration (double-buffered
// cp.async staging).  Work item = (pooled position, channel half): 338 of the CTA's 352 threads.  A
// thread keeps the 4x4 input patch of its position in registers as broadcast pairs (p,p) and walks
// its 8 CHANNEL PAIRS: the weights of channels (2q, 2q+1) lie interleaved in shared memory, so five
// broadcast LDS.128 deliver the nine (w_2q, w_2q+1) pairs and the bias pair as ready-made 64-bit
// operands, and every FMA is a packed fma.rn.f32x2 with no register moves around it - the same
// operand order per output as conv4, so the results are bit-identical to conv_pool_fwd_kernel's.
// Lanes are consecutive pooled positions: a warp's store is one contiguous run of a channel row.
constexpr int kStreamThreads = 352;

template <int CTAS_PER_SM>
__global__ void __launch_bounds__(kStreamThreads, CTAS_PER_SM) conv_pool_stream_kernel(const float* __restrict__ x,
                                                                             const float* __restrict__ w,
                                                                             const float* __restrict__ bias, int B,
                                                                             float* __restrict__ act, size_t ld_act) {
  __shared__ __align__(16) float img[2][784];
  __shared__ __align__(16) float2 wsm[16][10];         // [channel pair][9 weights, bias] as (even, odd) channel
  const int tid = threadIdx.x;
  const bool active = tid < 2 * kPooled;
  const int p = tid % kPooled, half = tid / kPooled;
  const int i = p / 13, j = p - i * 13;

  auto stage = [&](int buf, int b) {
    if (tid < 196) {
      const unsigned dst = (unsigned)__cvta_generic_to_shared(&img[buf][tid * 4]);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(x + (size_t)b * 784 + tid * 4) : "memory");
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  };
  int b = blockIdx.x;
  if (b < B) stage(0, b);                              // x is an input: staged before the dependency wait
  pdl_wait();
  pdl_launch();
  for (int t = tid; t < 16 * 10 * 2; t += kStreamThreads) {
    const int odd = t & 1, q = (t >> 1) % 10, cp = t / 20;
    const int c = 2 * cp + odd;
    reinterpret_cast<float*>(&wsm[cp][q])[odd] = q < 9 ? w[c * 9 + q] : bias[c];
  }
  int buf = 0;
  for (; b < B; b += gridDim.x, buf ^= 1) {
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();                                   // this image has landed; the other buffer is free
    if (b + (int)gridDim.x < B) stage(buf ^ 1, b + gridDim.x);
    if (!active) continue;
    float2 pt[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const float2 a0 = *reinterpret_cast<const float2*>(&img[buf][(2 * i + r) * 28 + 2 * j]);
      const float2 a1 = *reinterpret_cast<const float2*>(&img[buf][(2 * i + r) * 28 + 2 * j + 2]);
      pt[r][0] = make_float2(a0.x, a0.x);
      pt[r][1] = make_float2(a0.y, a0.y);
      pt[r][2] = make_float2(a1.x, a1.x);
      pt[r][3] = make_float2(a1.y, a1.y);
    }
    float* out = act + (size_t)b * ld_act + p;
#pragma unroll 2
    for (int cc = 0; cc < 8; ++cc) {
      const int cp = half * 8 + cc;
      const float4* wq = reinterpret_cast<const float4*>(&wsm[cp][0]);
      float2 wv[10];
#pragma unroll
      for (int q = 0; q < 5; ++q) {
        const float4 t = wq[q];
        wv[2 * q] = make_float2(t.x, t.y);
        wv[2 * q + 1] = make_float2(t.z, t.w);
      }
      float2 v[4];
#pragma unroll
      for (int d = 0; d < 4; ++d) {
        const int di = d >> 1, dj = d & 1;
        float2 a = wv[9];
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
          for (int q = 0; q < 3; ++q) a = __ffma2_rn(wv[u * 3 + q], pt[di + u][dj + q], a);
        v[d] = a;
      }
      const float m0 = fmaxf(fmaxf(fmaxf(v[0].x, v[1].x), fmaxf(v[2].x, v[3].x)), 0.f);
      const float m1 = fmaxf(fmaxf(fmaxf(v[0].y, v[1].y), fmaxf(v[2].y, v[3].y)), 0.f);
      __stcs(out + (2 * cp) * kPooled, m0);
      __stcs(out + (2 * cp + 1) * kPooled, m1);
    }
  }
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

// The same work with WARP-autonomous images: no CTA-wide barrier in the loop.  In the CTA-per-image
// form above every image ends at a __syncthreads, all warps of a CTA hit the FP32 pipe in the same
// phase and then all wait (ncu: `math_pipe_throttle` and `barrier` the two top stalls, FMA pipe 58 %
// busy = half the FP32 peak, profiles/r2_ffma_rate.txt).  Here each warp stages ITS OWN image into a
// private double buffer (cp.async, 7 x 16 B per lane), and its lanes walk the 338 (position, channel
// half) items of that image in 11 rounds; warps drift apart and keep the pipe evenly fed.  Same
// per-output operand order: bit-identical results.
constexpr int kStreamWarps = 8;
constexpr size_t kStreamWarpSmem = (size_t)kStreamWarps * 2 * 784 * 4 + 16 * 10 * 8;

__global__ void __launch_bounds__(32 * kStreamWarps, 2) conv_pool_stream_warp_kernel(const float* __restrict__ x,
                                                                                     const float* __restrict__ w,
                                                                                     const float* __restrict__ bias, int B,
                                                                                     float* __restrict__ act, size_t ld_act) {
  extern __shared__ __align__(16) unsigned char conv_smem[];
  float2(*wsm)[10] = reinterpret_cast<float2(*)[10]>(conv_smem);                       // [16][10]
  float* imgs = reinterpret_cast<float*>(conv_smem + 16 * 10 * 8);                      // [warp][buf][784]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* mine = imgs + (size_t)warp * 2 * 784;
  const int gw = blockIdx.x * kStreamWarps + warp, nw = gridDim.x * kStreamWarps;

  auto stage = [&](int buf, int b) {
    for (int t = lane; t < 196; t += 32) {
      const unsigned dst = (unsigned)__cvta_generic_to_shared(mine + buf * 784 + t * 4);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(x + (size_t)b * 784 + t * 4) : "memory");
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  };
  int b = gw;
  if (b < B) stage(0, b);                              // x is an input: staged before the dependency wait
  pdl_wait();
  pdl_launch();
  for (int t = tid; t < 16 * 10 * 2; t += 32 * kStreamWarps) {
    const int odd = t & 1, q = (t >> 1) % 10, cp = t / 20;
    const int c = 2 * cp + odd;
    reinterpret_cast<float*>(&wsm[cp][q])[odd] = q < 9 ? w[c * 9 + q] : bias[c];
  }
  __syncthreads();                                     // the only CTA-wide barrier
  int buf = 0;
  for (; b < B; b += nw, buf ^= 1) {
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();                                      // this image has landed; the other buffer is free
    if (b + nw < B) stage(buf ^ 1, b + nw);
    const float* img = mine + buf * 784;
    float* outb = act + (size_t)b * ld_act;
#pragma unroll 1
    for (int item = lane; item < 2 * kPooled; item += 32) {
      const int half = item >= kPooled, p = item - half * kPooled;
      const int i = p / 13, j = p - i * 13;
      float2 pt[4][4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float2 a0 = *reinterpret_cast<const float2*>(&img[(2 * i + r) * 28 + 2 * j]);
        const float2 a1 = *reinterpret_cast<const float2*>(&img[(2 * i + r) * 28 + 2 * j + 2]);
        pt[r][0] = make_float2(a0.x, a0.x);
        pt[r][1] = make_float2(a0.y, a0.y);
        pt[r][2] = make_float2(a1.x, a1.x);
        pt[r][3] = make_float2(a1.y, a1.y);
      }
      float* out = outb + p;
#pragma unroll 2
      for (int cc = 0; cc < 8; ++cc) {
        const int cp = half * 8 + cc;
        const float4* wq = reinterpret_cast<const float4*>(&wsm[cp][0]);
        float2 wv[10];
#pragma unroll
        for (int q = 0; q < 5; ++q) {
          const float4 t = wq[q];
          wv[2 * q] = make_float2(t.x, t.y);
          wv[2 * q + 1] = make_float2(t.z, t.w);
        }
        float2 v[4];
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          const int di = d >> 1, dj = d & 1;
          float2 a = wv[9];
#pragma unroll
          for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int q = 0; q < 3; ++q) a = __ffma2_rn(wv[u * 3 + q], pt[di + u][dj + q], a);
          v[d] = a;
        }
        const float m0 = fmaxf(fmaxf(fmaxf(v[0].x, v[1].x), fmaxf(v[2].x, v[3].x)), 0.f);
        const float m1 = fmaxf(fmaxf(fmaxf(v[0].y, v[1].y), fmaxf(v[2].y, v[3].y)), 0.f);
        __stcs(out + (2 * cp) * kPooled, m0);
        __stcs(out + (2 * cp + 1) * kPooled, m1);
      }
    }
  }
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

// The Alice -> Alice weight hand-off: 288 + 32 floats, one small CTA in the step's launch chain (two
// cudaMemcpyAsync device copies in its place cost ~40 us per hand-off: copy-engine latency, and the
// chain of dependent launches is broken twice).
__global__ void __launch_bounds__(128) conv_weights_copy_kernel(float* __restrict__ dw, float* __restrict__ db,
                                                                const float* __restrict__ sw, const float* __restrict__ sb) {
  pdl_wait();
  pdl_launch();
  for (int t = threadIdx.x; t < 288; t += 128) dw[t] = sw[t];
  if (threadIdx.x < 32) db[threadIdx.x] = sb[threadIdx.x];
}

// Weight gradient.  grid (gx, 4): CTA (bx, cg) loops over samples bx, bx+gx, ...; warp = one
// channel, lanes stride over the 169 pooled positions; 10 register accumulators per lane
// (9 dW + db).  Per-CTA partial sums go to `partial[gx][320]`; the last CTA to finish
// (ticket) adds them in a fixed order, then either stores dw/db (OPT==0) or applies the
// optimizer in place.  Deterministic.
template <int OPT>
__global__ void __launch_bounds__(256) conv_pool_bwd_kernel(const float* __restrict__ x, float* __restrict__ w,
                                                            float* __restrict__ bias,
                                                            const float* __restrict__ dact, int B,
                                                            float* __restrict__ partial, unsigned* __restrict__ ticket,
                                                            float* __restrict__ dw_out, float* __restrict__ db_out,
                                                            float* __restrict__ s1, float* __restrict__ s2,
                                                            OptParams op, int prefetch_early) {
  // (Letting this kernel add the server's n-split input-gradient partials itself - one launch less in
  // sn_vanilla_step - was built, bit-identical, and measured SLOWER: 5.7 -> 12.7 us here against the
  // 4.5 us reduction launch it saved.  Removed.)
  __shared__ float img[28 * 28];
  __shared__ unsigned is_last;
  const int cg = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = cg * kConvCG + warp;
  SN_TP_DECL;
  SN_TP(0);
  // Before the dependency wait (prefetch_early: the host knows the launch right before this one
  // is not another weight update of the same Alice): the conv weights, last written at least two
  // launches back, and the first image (an input, never written by the chain).
  if (!prefetch_early) { pdl_wait(); pdl_launch(); }
  float wr[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) wr[t] = w[c * 9 + t];
  const float bc = bias[c];
  if (blockIdx.x < B)
    for (int t = threadIdx.x; t < 196; t += 256)
      reinterpret_cast<float4*>(img)[t] = reinterpret_cast<const float4*>(x + (size_t)blockIdx.x * 784)[t];
  if (prefetch_early) { pdl_wait(); pdl_launch(); }
  SN_TP(1);
  float g[10];
#pragma unroll
  for (int t = 0; t < 10; ++t) g[t] = 0.f;

  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    // this lane's (up to 6) cut-gradient values, all requested before anything waits: inside the
    // position loop each load sat behind the ReLU test and cost a full L2 round trip per position
    // (clock64 phase timing: 5.4k of the kernel's 8.5k compute cycles)
    float dv[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      const int p = lane + 32 * q;
      dv[q] = p < kPooled ? __ldg(dact + (size_t)b * kD0 + (size_t)c * kPooled + p) : 0.f;
    }
    if (b != blockIdx.x) {
      __syncthreads();
      for (int t = threadIdx.x; t < 196; t += 256)
        reinterpret_cast<float4*>(img)[t] = reinterpret_cast<const float4*>(x + (size_t)b * 784)[t];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      const int p = lane + 32 * q;
      if (p < kPooled) {
        const int i = p / 13, j = p - i * 13;
        float v[4];
        conv4(img, wr, bc, i, j, v);
        int best = 0;                       // first maximum wins (torch CPU max_pool2d)
        float m = v[0];
#pragma unroll
        for (int d = 1; d < 4; ++d)
          if (v[d] > m) { m = v[d]; best = d; }
        if (m > 0.f) {                      // ReLU gate
          const float d = dv[q];
          const int r0 = 2 * i + (best >> 1), c0 = 2 * j + (best & 1);
#pragma unroll
          for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int x2 = 0; x2 < 3; ++x2) g[u * 3 + x2] = __fmaf_rn(d, img[(r0 + u) * 28 + c0 + x2], g[u * 3 + x2]);
          g[9] += d;
        }
      }
    }
  }
  SN_TP(2);
#pragma unroll
  for (int t = 0; t < 10; ++t) g[t] = warp_sum(g[t]);
  if (lane == 0) {
    float* dst = partial + (size_t)blockIdx.x * kConvParams;
#pragma unroll
    for (int t = 0; t < 9; ++t) dst[c * 9 + t] = g[t];
    dst[288 + c] = g[9];
  }
  SN_TP(3);
  __threadfence();
  __syncthreads();
  SN_TP(4);
  if (threadIdx.x == 0) {
    unsigned t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x * gridDim.y - 1);
  }
  __syncthreads();
  SN_TP(5);
  if (!is_last) return;
  __threadfence();
  SN_TP(6);
  {
    // 320 outputs on 256 threads: threads 0..63 own two (idx, idx+256).  Everything a thread needs -
    // parameter, optimizer state and all partials of BOTH its outputs - is requested before the
    // first add, so the tail is one round trip; partials are added in CTA order.
    constexpr int kMaxGx = 32;                       // partial rows summed from registers in one go
    float pv[2] = {0.f, 0.f}, a[2] = {0.f, 0.f}, bq[2] = {0.f, 0.f}, s[2] = {0.f, 0.f};
    float* pp[2];
    bool own[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int idx = threadIdx.x + 256 * j;
      own[j] = idx < kConvParams;
      pp[j] = idx < 288 ? (w + idx) : (bias + (own[j] ? idx - 288 : 0));
      if (OPT != 0 && own[j]) {
        pv[j] = *pp[j];
        a[j] = s1[idx];
        if (OPT == 2) bq[j] = s2[idx];
      }
    }
    for (unsigned r0 = 0; r0 < gridDim.x; r0 += kMaxGx) {
      float t[2][kMaxGx];
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (unsigned q = 0; q < kMaxGx; ++q)
          t[j][q] = (own[j] && r0 + q < gridDim.x) ? __ldcg(partial + (size_t)(r0 + q) * kConvParams + threadIdx.x + 256 * j) : 0.f;
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (unsigned q = 0; q < kMaxGx; ++q)
          if (r0 + q < gridDim.x) s[j] += t[j][q];
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int idx = threadIdx.x + 256 * j;
      if (!own[j]) continue;
      if (OPT == 0) {
        if (idx < 288) dw_out[idx] = s[j]; else db_out[idx - 288] = s[j];
      } else {
        opt_update<OPT>(op, pv[j], s[j], a[j], bq[j]);
        *pp[j] = pv[j];
        s1[idx] = a[j];
        if (OPT == 2) s2[idx] = bq[j];
      }
    }
  }
  SN_TP(7);
  SN_TP_FLUSH();
  if (threadIdx.x == 0) *ticket = 0u;     // re-arm for the next launch
}

}  // namespace sn
