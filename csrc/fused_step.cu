// The training step as two kernels.
//
//   fwd_bwd_kernel  one CTA per image: the whole forward pass, the loss, the data
//                   gradients of every layer and the two conv weight gradients,
//                   with every activation of the image resident in shared memory.
//                   (reference: models/model.py:20-27 + autograd of
//                   data_parallelism_train.py:195-198; kernels K1-K17 of SURVEY §2.7.)
//   update_kernel   one CTA per 512-float slice of the flat parameter vector: the
//                   batch reduction of the weight gradient for that slice (FC
//                   layers: outer products over the batch; conv: sum of the
//                   per-image partials), the peer-memory all-reduce of the slice,
//                   the 1/W scale and the SGD-momentum update, all in registers.
//                   (reference: data_parallelism_train.py:187,199 SGD; :210,226-244
//                   the parameter exchange; K13/K18/K19/K22.)
//
// Exploited structure: after max-pooling, only one of the four conv outputs in a
// window carries gradient, so the conv weight gradients run over the pooled grid
// with the 2-bit argmax as an address offset (4x fewer MACs than the dense form),
// and ReLU'/pool' cost nothing (K14/K15 fused away).
#include <cuda_runtime.h>
#include <type_traits>
#include <stdint.h>

#include "common.cuh"
#include "exchange.cuh"
#include "layout.h"
#include "step_types.h"

namespace dnnb200 {

// ============================================================================
// fwd_bwd_kernel
// ============================================================================
constexpr int kFBThreads = 512;
constexpr int kRedFloats = 6144;

struct __align__(16) FBSmem {
  alignas(16) float x[kImgNumel];          // input image, CHW (bulk-copied by the TMA engine)
  alignas(16) float w1t[75 * 8];           // conv1 weights [(c,kh,kw)][half*4 + j]   (j<3: out channel half*3+j)
  alignas(16) float b1[8];
  alignas(16) float p1[1176 + 8];          // pooled relu(conv1)  [6][14][14]
  alignas(16) float w2t[150 * 16];         // conv2 weights [(c,kh,kw)][o]   (forward)
  alignas(16) float w2d[16 * 25 * 8];      // conv2 weights [(o,kh,kw)][c8]  (data gradient)
  alignas(16) float b2[16];
  alignas(16) float p2[400];               // pooled relu(conv2)  [16][5][5] == fc1 input
  alignas(16) float h1[120];
  alignas(16) float h2[88];
  alignas(16) float logits[16];
  alignas(16) float d3[16];
  alignas(16) float d2[88];
  alignas(16) float d1[120];
  alignas(16) float g2[400];               // dL/d p2, masked by p2 > 0
  alignas(16) float g1[1176 + 8];          // dL/d p1, masked by p1 > 0
  alignas(16) float dc2p[16 * 18 * 18];    // dL/d conv2-out, dense, zero border of 4
  alignas(16) float red[kRedFloats];       // scratch for split reductions
  alignas(8) unsigned long long mbar;     // mbarrier for the bulk copy
  alignas(16) unsigned char idx1[1176 + 8];
  alignas(16) unsigned char idx2[400];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(phase)
      : "memory");
}
// 1-D bulk copy global -> shared through the TMA engine (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes,
                                         unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// y[m] = act(b[m] + W[m,:] . in), W row-major [M][K] in global, K % 4 == 0, in/out in smem.
// One warp per row, four rows in flight per warp to cover the L2 latency.
template <int M, int K, bool kRelu>
__device__ __forceinline__ void fc_forward(const float* __restrict__ W, const float* __restrict__ b,
                                           const float* in, float* out) {
  constexpr int K4 = K / 4;
  constexpr int kWarps = kFBThreads / 32;
  constexpr int kIter = (K4 + 31) / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float4* in4 = reinterpret_cast<const float4*>(in);
  for (int r0 = warp; r0 < M; r0 += 4 * kWarps) {
    float4 w[4][kIter];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = r0 + q * kWarps;
#pragma unroll
      for (int it = 0; it < kIter; ++it) {
        const int k4 = lane + it * 32;
        w[q][it] = (r < M && k4 < K4) ? __ldg(reinterpret_cast<const float4*>(W + (size_t)r * K) + k4)
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = r0 + q * kWarps;
      float acc = 0.f;
#pragma unroll
      for (int it = 0; it < kIter; ++it) {
        const int k4 = lane + it * 32;
        if (k4 < K4) {
          const float4 v = in4[k4];
          acc += w[q][it].x * v.x + w[q][it].y * v.y + w[q][it].z * v.z + w[q][it].w * v.w;
        }
      }
      acc = warp_sum(acc);
      if (lane == 0 && r < M) {
        acc += __ldg(b + r);
        out[r] = kRelu ? fmaxf(acc, 0.f) : acc;
      }
    }
  }
}

// dx[k] = [act[k] > 0] * sum_m W[m,k] * dy[m];  W row-major [M][K] global, dy/act/dx in smem.
// Threads are (row group g, float4 column k4); the groups' partial sums meet in `red`.
template <int M, int K>
__device__ __forceinline__ void fc_backward_dx(const float* __restrict__ W, const float* dy,
                                               const float* act, float* dx, float* red) {
  constexpr int K4 = K / 4;
  constexpr int G0 = kFBThreads / K4;
  constexpr int G = G0 < M ? G0 : M;
  constexpr int kRows = (M + G - 1) / G;
  static_assert(G * K <= kRedFloats, "reduction scratch too small");
  const int tid = threadIdx.x;
  if (tid < G * K4) {
    const int g = tid / K4, k4 = tid - g * K4;
    const int m0 = g * kRows;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 12
    for (int q = 0; q < kRows; ++q) {
      const int m = m0 + q;
      if (m < M) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(W + (size_t)m * K) + k4);
        const float d = dy[m];
        acc.x += w.x * d; acc.y += w.y * d; acc.z += w.z * d; acc.w += w.w * d;
      }
    }
    reinterpret_cast<float4*>(red)[g * K4 + k4] = acc;
  }
  __syncthreads();
  if (tid < K) {
    float s = 0.f;
#pragma unroll
    for (int g = 0; g < G; ++g) s += red[g * K + tid];
    dx[tid] = act[tid] > 0.f ? s : 0.f;
  }
  __syncthreads();
}

template <bool kTrain>
__global__ void __launch_bounds__(kFBThreads, 1)
fwd_bwd_kernel(const float* __restrict__ params, const float* __restrict__ x,
               const long long* __restrict__ labels, float* __restrict__ saved,
               float* __restrict__ part1, float* __restrict__ part2, float* __restrict__ img_loss,
               int* __restrict__ img_correct, float* __restrict__ logits_out, int B,
               float inv_batch) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  FBSmem& s = *reinterpret_cast<FBSmem*>(smem_raw);
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    mbar_init(&s.mbar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t phase = 0;
  int img = blockIdx.x;
  if (tid == 0 && img < B) {
    mbar_expect_tx(&s.mbar, kImgNumel * 4);
    bulk_g2s(s.x, x + (size_t)img * kImgNumel, kImgNumel * 4, &s.mbar);
  }

  // ---- conv weights into shared memory, re-laid out for broadcast vector loads ----
  for (int i = tid; i < 75 * 8; i += kFBThreads) {
    const int t = i >> 3, q = i & 7, half = q >> 2, j = q & 3;
    s.w1t[i] = j < 3 ? __ldg(params + kOffW1 + (half * 3 + j) * 75 + t) : 0.f;
  }
  if (tid < 8) s.b1[tid] = tid < 6 ? __ldg(params + kOffB1 + tid) : 0.f;
  for (int i = tid; i < 150 * 16; i += kFBThreads) {
    const int t = i >> 4, o = i & 15;
    s.w2t[i] = __ldg(params + kOffW2 + o * 150 + t);
  }
  if (kTrain) {
    for (int i = tid; i < 16 * 25 * 8; i += kFBThreads) {
      const int c = i & 7, ot = i >> 3, o = ot / 25, t = ot - o * 25;
      s.w2d[i] = c < 6 ? __ldg(params + kOffW2 + o * 150 + c * 25 + t) : 0.f;
    }
  }
  if (tid < 16) s.b2[tid] = __ldg(params + kOffB2 + tid);

  for (; img < B; img += gridDim.x) {
    mbar_wait(&s.mbar, phase);
    phase ^= 1u;
    __syncthreads();  // weights (first round) and image are visible to everyone

    // ================= conv1 + bias + ReLU + maxpool (K1-K3) =================
    // thread = (pooled pixel, half of the output channels); 12 accumulators
    {
      const int half = tid / 224, pp = tid - half * 224;
      if (tid < 448 && pp < 196) {
        const int ph = pp / 14, pw = pp - ph * 14;
        float acc[3][4];
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
          for (int q = 0; q < 4; ++q) acc[j][q] = 0.f;
#pragma unroll 1
        for (int c = 0; c < 3; ++c) {
          const float* xb = s.x + c * 1024 + (2 * ph) * 32 + 2 * pw;
          float patch[6][6];
#pragma unroll
          for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int q = 0; q < 3; ++q) {
              const float2 v = *reinterpret_cast<const float2*>(xb + r * 32 + 2 * q);
              patch[r][2 * q] = v.x;
              patch[r][2 * q + 1] = v.y;
            }
          const float* wb = s.w1t + (c * 25) * 8 + half * 4;
#pragma unroll
          for (int kh = 0; kh < 5; ++kh)
#pragma unroll
            for (int kw = 0; kw < 5; ++kw) {
              const float4 w = *reinterpret_cast<const float4*>(wb + (kh * 5 + kw) * 8);
              const float a00 = patch[kh][kw], a01 = patch[kh][kw + 1];
              const float a10 = patch[kh + 1][kw], a11 = patch[kh + 1][kw + 1];
              acc[0][0] += w.x * a00; acc[0][1] += w.x * a01; acc[0][2] += w.x * a10; acc[0][3] += w.x * a11;
              acc[1][0] += w.y * a00; acc[1][1] += w.y * a01; acc[1][2] += w.y * a10; acc[1][3] += w.y * a11;
              acc[2][0] += w.z * a00; acc[2][1] += w.z * a01; acc[2][2] += w.z * a10; acc[2][3] += w.z * a11;
            }
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int o = half * 3 + j;
          const float bias = s.b1[o];
          float best = fmaxf(acc[j][0] + bias, 0.f);
          int bi = 0;
#pragma unroll
          for (int q = 1; q < 4; ++q) {
            const float v = fmaxf(acc[j][q] + bias, 0.f);
            if (v > best) { best = v; bi = q; }  // strict: first maximum wins, as ATen does
          }
          s.p1[o * 196 + pp] = best;
          s.idx1[o * 196 + pp] = (unsigned char)bi;
        }
      }
    }
    __syncthreads();

    // ================= conv2 + bias + ReLU + maxpool (K4-K6) =================
    // thread = (pair of input channels, quad of output channels, pooled pixel)
    if (tid < 384) {
      const int grp = tid >> 5, pp = tid & 31;
      const int cp = grp >> 2, oq = grp & 3;
      if (pp < 25) {
        const int ph = pp / 5, pw = pp - ph * 5;
        float acc[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int q = 0; q < 4; ++q) acc[j][q] = 0.f;
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          const int c = cp * 2 + cc;
          const float* xb = s.p1 + c * 196 + (2 * ph) * 14 + 2 * pw;
          float patch[6][6];
#pragma unroll
          for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int q = 0; q < 3; ++q) {
              const float2 v = *reinterpret_cast<const float2*>(xb + r * 14 + 2 * q);
              patch[r][2 * q] = v.x;
              patch[r][2 * q + 1] = v.y;
            }
          const float* wb = s.w2t + (c * 25) * 16 + oq * 4;
#pragma unroll
          for (int kh = 0; kh < 5; ++kh)
#pragma unroll
            for (int kw = 0; kw < 5; ++kw) {
              const float4 w = *reinterpret_cast<const float4*>(wb + (kh * 5 + kw) * 16);
              const float a00 = patch[kh][kw], a01 = patch[kh][kw + 1];
              const float a10 = patch[kh + 1][kw], a11 = patch[kh + 1][kw + 1];
              acc[0][0] += w.x * a00; acc[0][1] += w.x * a01; acc[0][2] += w.x * a10; acc[0][3] += w.x * a11;
              acc[1][0] += w.y * a00; acc[1][1] += w.y * a01; acc[1][2] += w.y * a10; acc[1][3] += w.y * a11;
              acc[2][0] += w.z * a00; acc[2][1] += w.z * a01; acc[2][2] += w.z * a10; acc[2][3] += w.z * a11;
              acc[3][0] += w.w * a00; acc[3][1] += w.w * a01; acc[3][2] += w.w * a10; acc[3][3] += w.w * a11;
            }
        }
        float4* dst = reinterpret_cast<float4*>(s.red + (grp * 32 + pp) * 16);
#pragma unroll
        for (int j = 0; j < 4; ++j) dst[j] = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
      }
    }
    __syncthreads();
    if (tid < 400) {
      const int o = tid / 25, pp = tid - o * 25;
      const int oq = o >> 2, j = o & 3;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int cp = 0; cp < 3; ++cp) {
        const float4 t = *reinterpret_cast<const float4*>(s.red + ((cp * 4 + oq) * 32 + pp) * 16 + j * 4);
        v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
      }
      const float bias = s.b2[o];
      float best = fmaxf(v.x + bias, 0.f);
      int bi = 0;
      float t = fmaxf(v.y + bias, 0.f);
      if (t > best) { best = t; bi = 1; }
      t = fmaxf(v.z + bias, 0.f);
      if (t > best) { best = t; bi = 2; }
      t = fmaxf(v.w + bias, 0.f);
      if (t > best) { best = t; bi = 3; }
      s.p2[tid] = best;  // (C,H,W) flatten order: o*25 + ph*5 + pw
      s.idx2[tid] = (unsigned char)bi;
    }
    __syncthreads();

    // the image buffer is free only in eval mode (training re-reads it for conv1 wgrad)
    if (!kTrain && tid == 0 && img + gridDim.x < B) {
      mbar_expect_tx(&s.mbar, kImgNumel * 4);
      bulk_g2s(s.x, x + (size_t)(img + gridDim.x) * kImgNumel, kImgNumel * 4, &s.mbar);
    }

    // ================= fc1, fc2, fc3 (K8-K10) =================
    fc_forward<kF1, kFlat, true>(params + kOffF1W, params + kOffF1B, s.p2, s.h1);
    __syncthreads();
    fc_forward<kF2, kF1, true>(params + kOffF2W, params + kOffF2B, s.h1, s.h2);
    __syncthreads();
    fc_forward<kF3, kF2, false>(params + kOffF3W, params + kOffF3B, s.h2, s.logits);
    __syncthreads();

    // ================= softmax cross-entropy, forward and backward (K11/K12) =================
    if (warp == 0) {
      const int label = (int)__ldg(labels + img);
      const float z = lane < 10 ? s.logits[lane] : -INFINITY;
      const float m = warp_max(z);
      const float e = lane < 10 ? __expf(z - m) : 0.f;
      const float sum = warp_sum(e);
      const float lse = m + __logf(sum);
      const float zl = __shfl_sync(0xffffffffu, z, label & 31);
      // first index attaining the maximum, as torch.argmax
      const unsigned int hit = __ballot_sync(0xffffffffu, lane < 10 && z == m);
      const int pred = __ffs(hit) - 1;
      if (lane < 10) {
        s.d3[lane] = (e / sum - (lane == label ? 1.f : 0.f)) * inv_batch;
        if (logits_out) logits_out[(size_t)img * 10 + lane] = z;
      }
      if (lane == 0) {
        img_loss[img] = lse - zl;
        if (img_correct) img_correct[img] = (pred == label) ? 1 : 0;
      }
    }
    if (!kTrain) {
      __syncthreads();
      continue;
    }
    __syncthreads();

    // ================= FC data gradients (K13 dX, K14) =================
    fc_backward_dx<kF3, kF2>(params + kOffF3W, s.d3, s.h2, s.d2, s.red);
    fc_backward_dx<kF2, kF1>(params + kOffF2W, s.d2, s.h1, s.d1, s.red);
    fc_backward_dx<kF1, kFlat>(params + kOffF1W, s.d1, s.p2, s.g2, s.red);

    // record for the update kernel: FC inputs and deltas of this image
    {
      float* sv = saved + (size_t)img * kSavStride;
      if (tid < 400) sv[kSavP2 + tid] = s.p2[tid];
      if (tid < 120) { sv[kSavH1 + tid] = s.h1[tid]; sv[kSavD1 + tid] = s.d1[tid]; }
      if (tid < 84) { sv[kSavH2 + tid] = s.h2[tid]; sv[kSavD2 + tid] = s.d2[tid]; }
      if (tid < 10) sv[kSavD3 + tid] = s.d3[tid];
    }
    // conv gradients of this image: conv1 block [0,512) and conv2 block [512,2944) of the flat prefix
    float* part = part2 + (size_t)img * (kConvPrefix - kOffW2) - kOffW2;  // indexed with flat offsets >= kOffW2
    float* partc1 = part1 + (size_t)img * kOffW2;

    // ================= conv2 weight gradient over the pooled grid (K16 wgrad, K15) =================
    // zero the dense padded gradient plane for the data gradient in the same pass
    for (int i = tid; i < 16 * 18 * 18; i += kFBThreads) s.dc2p[i] = 0.f;
    if (tid < 480) {
      const int o = tid / 30, rem = tid - o * 30, c = rem / 5, kh = rem - c * 5;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll 5
      for (int pp = 0; pp < 25; ++pp) {
        const float g = s.g2[o * 25 + pp];
        const int id = s.idx2[o * 25 + pp];
        const int ph = pp / 5, pw = pp - ph * 5;
        const float* row = s.p1 + c * 196 + (2 * ph + (id >> 1) + kh) * 14 + 2 * pw + (id & 1);
        a0 += g * row[0]; a1 += g * row[1]; a2 += g * row[2]; a3 += g * row[3]; a4 += g * row[4];
      }
      float* dst = part + kOffW2 + o * 150 + c * 25 + kh * 5;
      dst[0] = a0; dst[1] = a1; dst[2] = a2; dst[3] = a3; dst[4] = a4;
    } else if (tid < 496) {
      const int o = tid - 480;
      float a = 0.f;
      for (int pp = 0; pp < 25; ++pp) a += s.g2[o * 25 + pp];
      part[kOffB2 + o] = a;
    }
    __syncthreads();
    if (tid < 400) {  // scatter pooled gradient to its argmax position (K15)
      const int o = tid / 25, pp = tid - o * 25, ph = pp / 5, pw = pp - ph * 5;
      const int id = s.idx2[tid];
      s.dc2p[o * 324 + (2 * ph + (id >> 1) + 4) * 18 + (2 * pw + (id & 1) + 4)] = s.g2[tid];
    }
    __syncthreads();

    // ================= conv2 data gradient (K16 dgrad), then ReLU'/pool' of conv1 =================
    {
      const int half = tid / 224, pix = tid - half * 224;
      if (tid < 448 && pix < 196) {
        const int ih = pix / 14, iw = pix - ih * 14;
        float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
        for (int oo = 0; oo < 8; ++oo) {
          const int o = half * 8 + oo;
#pragma unroll
          for (int kh = 0; kh < 5; ++kh) {
            const int r = ih + 4 - kh;
            if (r < 4 || r > 13) continue;  // rows of the zero border
            const float* row = s.dc2p + o * 324 + r * 18 + iw + 4;
            const float* wb = s.w2d + (o * 25 + kh * 5) * 8;
#pragma unroll
            for (int kw = 0; kw < 5; ++kw) {
              const float v = row[-kw];
              const float4 wa = *reinterpret_cast<const float4*>(wb + kw * 8);
              const float2 wc = *reinterpret_cast<const float2*>(wb + kw * 8 + 4);
              acc[0] += v * wa.x; acc[1] += v * wa.y; acc[2] += v * wa.z; acc[3] += v * wa.w;
              acc[4] += v * wc.x; acc[5] += v * wc.y;
            }
          }
        }
#pragma unroll
        for (int c = 0; c < 6; ++c) s.red[half * 1176 + c * 196 + pix] = acc[c];
      }
    }
    __syncthreads();
    for (int i = tid; i < 1176; i += kFBThreads) {
      const float v = s.red[i] + s.red[1176 + i];
      s.g1[i] = s.p1[i] > 0.f ? v : 0.f;
    }
    __syncthreads();

    // ================= conv1 weight gradient over the pooled grid (K17, K15) =================
    if (tid < 450) {
      const int pg = tid / 90, rem = tid - pg * 90;
      const int o = rem / 15, r2 = rem - o * 15, c = r2 / 5, kh = r2 - c * 5;
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll 4
      for (int pp = pg; pp < 196; pp += 5) {
        const float g = s.g1[o * 196 + pp];
        const int id = s.idx1[o * 196 + pp];
        const int ph = pp / 14, pw = pp - ph * 14;
        const float* row = s.x + c * 1024 + (2 * ph + (id >> 1) + kh) * 32 + 2 * pw + (id & 1);
        a0 += g * row[0]; a1 += g * row[1]; a2 += g * row[2]; a3 += g * row[3]; a4 += g * row[4];
      }
      float* dst = s.red + pg * 450 + o * 75 + c * 25 + kh * 5;  // red was consumed above
      dst[0] = a0; dst[1] = a1; dst[2] = a2; dst[3] = a3; dst[4] = a4;
    } else if (warp >= 15) {
      // warp 15 (threads 480..511): conv1 bias gradient
      for (int o = 0; o < 6; ++o) {
        float a = 0.f;
        for (int pp = lane; pp < 196; pp += 32) a += s.g1[o * 196 + pp];
        a = warp_sum(a);
        if (lane == 0) partc1[kOffB1 + o] = a;
      }
    }
    __syncthreads();
    if (tid < 450) {
      float a = 0.f;
#pragma unroll
      for (int pg = 0; pg < 5; ++pg) a += s.red[pg * 450 + tid];
      partc1[kOffW1 + tid] = a;
    }
    __syncthreads();  // s.x may be overwritten now
    if (tid == 0 && img + gridDim.x < B) {
      mbar_expect_tx(&s.mbar, kImgNumel * 4);
      bulk_g2s(s.x, x + (size_t)(img + gridDim.x) * kImgNumel, kImgNumel * 4, &s.mbar);
    }
  }
}

// ============================================================================
// update_kernel
// ============================================================================
#ifndef DNN_U_VEC
#define DNN_U_VEC 2
#endif
constexpr int kUVec = DNN_U_VEC;              // consecutive floats per thread (2 or 4)
constexpr int kUpdThreads = kSlice / kUVec;   // 256 or 128 threads per 512-float slice
static_assert(kUVec == 2 || kUVec == 4, "update kernel vector width");

#ifndef DNN_GRAD_UNROLL
#define DNN_GRAD_UNROLL 16
#endif
constexpr int kGradUnroll = DNN_GRAD_UNROLL;   // batch loop: a whole batch of 16 has all its loads in flight at once

// d(loss)/d(param[i]) and param[i+1] summed over the local batch. i is even and both
// elements lie in the same tensor row (all row lengths are even).
__device__ __forceinline__ float2 local_grad2(const RankCtx& cx, int i, int B, int n_part1) {
  float2 g = make_float2(0.f, 0.f);
  if (i < kOffW2) {
    const float* p = cx.part1 + i;
#pragma unroll kGradUnroll
    for (int b = 0; b < n_part1; ++b) {
      const float2 t = __ldg(reinterpret_cast<const float2*>(p + (size_t)b * kOffW2));
      g.x += t.x; g.y += t.y;
    }
    return g;
  }
  if (i < kConvPrefix) {
    const float* p = cx.part2 + (i - kOffW2);
#pragma unroll kGradUnroll
    for (int b = 0; b < B; ++b) {
      const float2 t = __ldg(reinterpret_cast<const float2*>(p + (size_t)b * (kConvPrefix - kOffW2)));
      g.x += t.x; g.y += t.y;
    }
    return g;
  }
  int off_d, off_a, j, k;
  bool is_bias = false;
  if (i < kOffF1B) {
    const int e = i - kOffF1W; j = e / kFlat; k = e - j * kFlat; off_d = kSavD1; off_a = kSavP2;
  } else if (i < kOffF2W) {
    j = i - kOffF1B; k = 0; off_d = kSavD1; off_a = 0; is_bias = true;
    if (j >= kF1) return g;
  } else if (i < kOffF2B) {
    const int e = i - kOffF2W; j = e / kF1; k = e - j * kF1; off_d = kSavD2; off_a = kSavH1;
  } else if (i < kOffF3W) {
    j = i - kOffF2B; k = 0; off_d = kSavD2; off_a = 0; is_bias = true;
    if (j >= kF2) return g;
  } else if (i < kOffF3B) {
    const int e = i - kOffF3W; j = e / kF2; k = e - j * kF2; off_d = kSavD3; off_a = kSavH2;
    if (e >= kF3 * kF2) return g;
  } else {
    j = i - kOffF3B; k = 0; off_d = kSavD3; off_a = 0; is_bias = true;
    if (j >= kF3) return g;
  }
  const float* sv = cx.saved;
  if (is_bias) {
    // j even, j+1 in range for all three biases (120, 84, 10 are even)
#pragma unroll kGradUnroll
    for (int b = 0; b < B; ++b) {
      const float2 d = __ldg(reinterpret_cast<const float2*>(sv + (size_t)b * kSavStride + off_d + j));
      g.x += d.x; g.y += d.y;
    }
  } else {
#pragma unroll kGradUnroll
    for (int b = 0; b < B; ++b) {
      const float d = __ldg(sv + (size_t)b * kSavStride + off_d + j);
      const float2 a = __ldg(reinterpret_cast<const float2*>(sv + (size_t)b * kSavStride + off_a + k));
      g.x += d * a.x; g.y += d * a.y;
    }
  }
  return g;
}

// kEmulated: several ranks emulated on one GPU wait on each other inside ONE cooperative launch.
// kMinBlocks (CTAs per SM the register budget must allow):
//   7  ranks emulated on one GPU: all 8 x 122 CTAs of the cooperative launch must be co-resident
//   1  one GPU, no exchange: no constraint (fastest gradient reduction)
//   3  real peers: all 122 CTAs must fit on the 52 SMs the forward/backward kernel (96 CTAs, one per SM, nearly
//      the whole register file each) leaves free. Otherwise the last CTAs land on SMs that kernel has just left,
//      and the NEXT step's clusters cannot be placed there until those CTAs have exited -- which, after NVLink
//      stores, takes ~3 us longer than their last instruction (the stores drain first). Measured at 2 GPUs:
//      25.3 -> 22.4 us per step; on one GPU the same budget costs 0.7 us, hence two instantiations.
// flat index i of a conv weight -> its places in the packed B operands (layout.h); biases and gaps have none
__device__ __forceinline__ void pack_conv_weight(float* __restrict__ packed, int i, float w) {
  if (i < kOffW1 + 450) {
    const int n = i / 75, rem = i - n * 75, c = rem / 25, tap = rem - c * 25;
    packed[kPackB1 + ((n >> 3) * kB1Chunks + tap) * 32 + (n & 7) * 4 + c] = w;
  } else if (i >= kOffW2 && i < kOffW2 + 2400) {
    const int e = i - kOffW2, o = e / 150, rem = e - o * 150, ch = rem / 25, tap = rem - ch * 25;
    packed[kPackB2 + ((o >> 3) * kB2Chunks + 2 * tap + (ch >> 2)) * 32 + (o & 7) * 4 + (ch & 3)] = w;
    packed[kPackBd + (4 * (24 - tap) + (o >> 2)) * 32 + ch * 4 + (o & 3)] = w;
  }
}

// kExchange = false: the instantiation for a rank without peers carries no exchange code at all (and no register
// pressure from it).
template <bool kEmulated, int kMinBlocks, bool kExchange>
__global__ void __launch_bounds__(kUpdThreads, kMinBlocks)
update_kernel(const RankCtx* __restrict__ ctxs, int B, int n_part1, int loss_slot, int flags,
              float* __restrict__ host_loss) {
  __shared__ typename std::conditional<kExchange, ExchangeShared, int>::type xsh;
  __shared__ RankCtx cx;        // this rank's context and ...
  __shared__ StepCtl ctl_snap;  // ... the host-set fields of its control block, staged BEFORE the dependency wait
  const int slice = blockIdx.x, tid = threadIdx.x;
  const int i = slice * kSlice + kUVec * tid;
  const bool in_range = i < kFlatNumel;
  long long t_enter = 0, t_dep = 0;
  if (tid == 0) t_enter = (long long)globaltimer_ns();
  // Programmatic dependent launch: the next step's forward/backward kernel may be scheduled from now on (it
  // loads its input batch and then waits for us). Everything the host configured before this step's
  // forward/backward kernel -- the context (pointers, peer tables) and lr / momentum / membership / timeout /
  // exchange variant -- is visible already (that kernel only lets us start after its own dependency wait), so it
  // is copied to shared memory now: after the wait, nothing but the gradient records themselves is loaded from
  // global memory before the first remote store goes out.
  if (!(flags & 2)) pdl_trigger();
  {
    static_assert(sizeof(RankCtx) % 4 == 0 && sizeof(StepCtl) % 4 == 0, "word copies");
    const unsigned int* src = reinterpret_cast<const unsigned int*>(ctxs + blockIdx.y);
    unsigned int* dst = reinterpret_cast<unsigned int*>(&cx);
    for (int w = tid; w < (int)(sizeof(RankCtx) / 4); w += kUpdThreads) dst[w] = __ldg(src + w);
  }
  __syncthreads();
  StepCtl* ctl = cx.ctl;
  if (tid < (int)(sizeof(StepCtl) / 4))
    reinterpret_cast<unsigned int*>(&ctl_snap)[tid] = reinterpret_cast<const volatile unsigned int*>(ctl)[tid];
  __syncthreads();
  // ranks that have announced their departure (leave markers in my flag rows): looked up now, off the critical path
  unsigned int left_now = 0u;
  if constexpr (kExchange) {
    if (cx.world > 1) left_now = leave_mask(cx, ctl_snap.active_mask, (unsigned int)(ctl_snap.step + 1ull));
  }
  // parameters and momentum of my elements were written by the PREVIOUS update kernel, which had completed (and
  // fenced) before this step's forward/backward kernel let us start: fetch them now (from L2, the coherence
  // point), so that the SGD update after the exchange has no load left in front of its stores
  const bool grad_only = (flags & 1) != 0;  // tests: leave the local batch gradient in cx.grads, touch nothing else
  float2 m[kUVec / 2], p[kUVec / 2];
#pragma unroll
  for (int k = 0; k < kUVec; k += 2) {
    m[k / 2] = make_float2(0.f, 0.f);
    p[k / 2] = make_float2(0.f, 0.f);
    if (in_range && !grad_only) {
      m[k / 2] = __ldcg(reinterpret_cast<const float2*>(cx.momentum + i + k));
      p[k / 2] = __ldcg(reinterpret_cast<const float2*>(cx.params + i + k));
    }
  }
  pdl_wait();
  long long* tr = cx.trace;
  if (tr && tid == 0) t_dep = (long long)globaltimer_ns();
  // the previous update kernel stored `step` before it released the counter our predecessor waited on, and the
  // host only changes it with stream-ordered kernels before that predecessor: the staged copy is current
  const unsigned long long step = ctl_snap.step;
  const unsigned int seq = (unsigned int)(step + 1ull);
  if (tr) {
    // flags bit 2 (value 4): keep the stamps of the last 1024 steps ([step % 1024][slice][8]) instead of one step
    if (flags & 4) tr += (size_t)(step & 1023ull) * gridDim.x * 8;
    if (tid == 0) { tr[slice * 8 + 0] = t_enter; tr[slice * 8 + 1] = t_dep; }
  }

  // Per-step mean loss (fixed summation order), by the last warp of the last slice: its elements lie beyond the
  // end of the parameter vector, so it has nothing else to do, and the bookkeeping (a read-modify-write of the
  // control block, a 4-byte store to the caller's pinned host buffer) runs beside the gradient reduction instead
  // of behind it. (It used to sit on thread 0 of slice 0, 0.8 us on the step's critical path.)
  static_assert(kFlatNumel - (kNumSlices - 1) * kSlice <= (kUpdThreads - 32) * kUVec, "last warp of last slice idle");
  if (slice == (int)gridDim.x - 1 && tid >= kUpdThreads - 32) {
    const int lane = tid & 31;
    float a = 0.f;
    for (int b = lane; b < B; b += 32) a += cx.img_loss[b];
    a = warp_sum(a);
    if (lane == 0) {
      const float mean = a / (float)B;
      cx.loss_ring[loss_slot & cx.loss_ring_mask] = mean;
      // the step's result goes straight to the caller's pinned host buffer: a 4-byte store over PCIe, no
      // copy-engine operation and no extra stream dependency on the step path (visible to the host once the
      // stream has been synchronised)
      if (host_loss != nullptr) *reinterpret_cast<volatile float*>(host_loss) = mean;
      if (!grad_only) {
        ctl->loss_sum += mean;
        ctl->loss_steps += 1u;
      }
    }
  }

  float v[kUVec];
#pragma unroll
  for (int k = 0; k < kUVec; ++k) v[k] = 0.f;
  if (in_range) {
#pragma unroll
    for (int k = 0; k < kUVec; k += 2) {
      const float2 g = local_grad2(cx, i + k, B, n_part1);
      v[k] = g.x; v[k + 1] = g.y;
    }
  }
  int n = 1;
  if (tr && tid == 0) tr[slice * 8 + 2] = (long long)globaltimer_ns();
  if constexpr (kExchange) if (cx.world > 1 && !grad_only) {
    const unsigned int variant = ctl_snap.two_shot;   // 0 one-shot, 1 two-shot, 2 NVLS, 3 one-shot with 128-byte lines
    ExchangeCfg cfg;
    cfg.configured = ctl_snap.active_mask;
    cfg.timeout_ns = ctl_snap.timeout_ns;
    cfg.left = left_now;
    if (variant == 2u && cx.nvls_mc != nullptr) exchange_slice_nvls<kUVec>(cx, cfg, &xsh, slice, seq, i, in_range, v, &n);
    else if (variant == 1u) exchange_slice_two_shot<kUVec>(cx, cfg, &xsh, slice, seq, i, in_range, v, &n);
    else if (variant == 3u) {
      const int left = kFlatNumel - slice * kSlice;
      exchange_slice_ll128<kUVec>(cx, cfg, &xsh, slice, seq, left < kSlice ? left : kSlice, v, &n);
    }
    else exchange_slice<kUVec>(cx, cfg, &xsh, slice, seq, i, in_range, v, &n);
  }
  if (tr && tid == 0) {
    tr[slice * 8 + 3] = (long long)globaltimer_ns();
    tr[slice * 8 + 7] = n;   // how many ranks entered this slice's sum (the divisor unless the host overrides it)
  }

  if (in_range && grad_only) {
    if (cx.grads) {
#pragma unroll
      for (int k = 0; k < kUVec; k += 2) *reinterpret_cast<float2*>(cx.grads + i + k) = make_float2(v[k], v[k + 1]);
    }
  } else if (in_range) {
    const unsigned int dov = ctl_snap.divisor_override;
    const float inv_n = 1.f / (float)(dov ? (int)dov : n);
    const float lr = ctl_snap.lr, mu = ctl_snap.mu;
#pragma unroll
    for (int k = 0; k < kUVec; k += 2) {
      const float g0 = v[k] * inv_n, g1 = v[k + 1] * inv_n;
      float2 mm = m[k / 2], pp = p[k / 2];
      mm.x = mu * mm.x + g0;  // zeroed momentum makes the first step buf = g, as torch.optim.SGD
      mm.y = mu * mm.y + g1;
      pp.x -= lr * mm.x;
      pp.y -= lr * mm.y;
      *reinterpret_cast<float2*>(cx.momentum + i + k) = mm;
      *reinterpret_cast<float2*>(cx.params + i + k) = pp;
      if (cx.packed != nullptr && i + k < kConvPrefix) {   // (slices 0..5 only) the tensor-core view of the conv weights
        pack_conv_weight(cx.packed, i + k, pp.x);
        pack_conv_weight(cx.packed, i + k + 1, pp.y);
      }
      if (cx.grads) *reinterpret_cast<float2*>(cx.grads + i + k) = make_float2(g0, g1);
    }
  }

  if (tr && tid == 0) tr[slice * 8 + 4] = (long long)globaltimer_ns();
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    if (tr) tr[slice * 8 + 5] = (long long)globaltimer_ns();
    const unsigned int done = atomicAdd(&ctl->ctas_done, 1u);
    if (tr) tr[slice * 8 + 6] = (long long)globaltimer_ns();
    if (done == gridDim.x - 1) {  // last CTA of this rank: everyone has read `step`
      ctl->ctas_done = 0u;
      if (!grad_only) *reinterpret_cast<volatile unsigned long long*>(&ctl->step) = step + 1ull;
      // "parameters of this step are in place": what the next forward/backward kernel spins on (fused_cluster.cu).
      // Every CTA fenced its stores before it bumped ctas_done, and this CTA has observed all the bumps; the
      // release also orders the two stores above (one fence on this path, not three).
      st_release_gpu(&ctl->updates_done, ctl_snap.updates_done + 1u);
      // the same number for the HOST (a 4-byte store over PCIe, nobody waits for it here): everything this step read
      // -- its input slot included -- has been consumed
      if (cx.host_progress != nullptr) *reinterpret_cast<volatile unsigned int*>(cx.host_progress) = ctl_snap.updates_done + 1u;
    }
  }
}

// ============================================================================
// eval reduction: mean over batches of the batch-mean loss, and the hit count
// (reference: data_parallelism_train.py:157-183 -- np.mean(losses), 100*corrects/total)
// ============================================================================
__global__ void eval_reduce_kernel(const float* __restrict__ img_loss, const int* __restrict__ img_correct,
                                   int N, int bs, double* __restrict__ out /* [2] */) {
  __shared__ double s_loss[32];
  __shared__ long long s_hit[32];
  const int nb = (N + bs - 1) / bs;
  double acc = 0.0;
  long long hit = 0;
  for (int b = threadIdx.x; b < nb; b += blockDim.x) {
    const int lo = b * bs, hi = min(N, lo + bs);
    float t = 0.f;
    for (int k = lo; k < hi; ++k) { t += img_loss[k]; hit += img_correct[k]; }
    acc += (double)(t / (float)(hi - lo));
  }
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    hit += __shfl_xor_sync(0xffffffffu, hit, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_loss[warp] = acc; s_hit[warp] = hit; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0; long long h = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a += s_loss[w]; h += s_hit[w]; }
    out[0] = a / (double)nb;
    out[1] = (double)h;
  }
}

// ============================================================================
// host launchers (used by runtime.cu and noalloc.cu)
// ============================================================================
int launch_fwd_bwd(bool train, const float* params, const float* x, const long long* labels,
                   float* saved, float* part1, float* part2, float* img_loss, int* img_correct,
                   float* logits_out, int B, int max_ctas, cudaStream_t stream) {
  static PerDeviceOnce attr_set[2];
  const size_t smem = sizeof(FBSmem);
  auto kern = train ? fwd_bwd_kernel<true> : fwd_bwd_kernel<false>;
  if (int rc = ensure_max_dynamic_smem(attr_set[train ? 1 : 0], kern, smem)) return rc;
  const int grid = B < max_ctas ? B : max_ctas;
  kern<<<grid, kFBThreads, smem, stream>>>(params, x, labels, saved, part1, part2, img_loss, img_correct,
                                           logits_out, B, 1.f / (float)B);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_update(const RankCtx* d_ctxs, int n_virtual_ranks, int B, int n_part1, int loss_slot, int flags,
                  float* host_loss, cudaStream_t stream) {
  dim3 grid(kNumSlices, n_virtual_ranks);
  if (n_virtual_ranks > 1) {
    // ranks emulated on ONE device wait on each other: every CTA must be resident
    void* args[] = {(void*)&d_ctxs, (void*)&B, (void*)&n_part1, (void*)&loss_slot, (void*)&flags, (void*)&host_loss};
    DNN_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)update_kernel<true, 7, true>, grid, dim3(kUpdThreads), args, 0, stream));
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kUpdThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (flags & kUpdFlagPeers)
      DNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, update_kernel<false, 3, true>, d_ctxs, B, n_part1, loss_slot, flags, host_loss));
    else if (flags & kUpdFlagHasGroup)   // peers, but the A/B switch asked for the unconstrained register budget
      DNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, update_kernel<false, 1, true>, d_ctxs, B, n_part1, loss_slot, flags, host_loss));
    else
      DNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, update_kernel<false, 1, false>, d_ctxs, B, n_part1, loss_slot, flags, host_loss));
  }
  return kOk;
}

int launch_eval_reduce(const float* img_loss, const int* img_correct, int N, int bs, double* out,
                       cudaStream_t stream) {
  eval_reduce_kernel<<<1, 256, 0, stream>>>(img_loss, img_correct, N, bs, out);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

size_t fwd_bwd_smem_bytes() { return sizeof(FBSmem); }

}  // namespace dnnb200
