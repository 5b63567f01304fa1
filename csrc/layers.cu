// Layer-at-a-time fp32 kernels behind the autograd Functions of ops/functional.py.
//
// These serve the general nn.Module path (any batch size, user-driven
// forward/backward/optimizer, reference: models/model.py:20-27) and double as the
// in-tree GPU oracle for the tensor-core kernels in tc_*.cu. The flagship training
// step does not go through them -- it uses the two fused kernels of fused_step.cu.
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"

namespace dnnb200 {

// ---------------------------------------------------------------------------
// conv 5x5 valid + bias + ReLU + maxpool 2x2/2 (K1-K3, K4-K6). NCHW fp32.
// One thread per pooled output; weights of one output channel staged in smem.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
conv_relu_pool_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                          const float* __restrict__ b, float* __restrict__ y,
                          unsigned char* __restrict__ idx, int N, int Ci, int H, int W, int Co) {
  extern __shared__ float sw[];  // [Ci*25] weights of channel o
  const int o = blockIdx.y;
  const int Hp = (H - 4) / 2, Wp = (W - 4) / 2;
  for (int i = threadIdx.x; i < Ci * 25; i += blockDim.x) sw[i] = w[o * Ci * 25 + i];
  __syncthreads();
  const int per_img = Hp * Wp;
  const long long total = (long long)N * per_img;
  const float bias = b[o];
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(t / per_img), pp = (int)(t - (long long)n * per_img);
    const int ph = pp / Wp, pw = pp - ph * Wp;
    float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f;
    for (int c = 0; c < Ci; ++c) {
      const float* xb = x + ((size_t)(n * Ci + c) * H + 2 * ph) * W + 2 * pw;
      float patch[6][6];
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int q = 0; q < 6; ++q) patch[r][q] = __ldg(xb + r * W + q);
      const float* wc = sw + c * 25;
#pragma unroll
      for (int kh = 0; kh < 5; ++kh)
#pragma unroll
        for (int kw = 0; kw < 5; ++kw) {
          const float wv = wc[kh * 5 + kw];
          a00 += wv * patch[kh][kw];
          a01 += wv * patch[kh][kw + 1];
          a10 += wv * patch[kh + 1][kw];
          a11 += wv * patch[kh + 1][kw + 1];
        }
    }
    float best = fmaxf(a00 + bias, 0.f);
    int bi = 0;
    float v = fmaxf(a01 + bias, 0.f);
    if (v > best) { best = v; bi = 1; }
    v = fmaxf(a10 + bias, 0.f);
    if (v > best) { best = v; bi = 2; }
    v = fmaxf(a11 + bias, 0.f);
    if (v > best) { best = v; bi = 3; }
    const size_t oi = ((size_t)(n * Co + o)) * per_img + pp;
    y[oi] = best;
    idx[oi] = (unsigned char)bi;
  }
}

// wgrad + bgrad over the pooled grid (K16/K17 wgrad with K14/K15 folded in).
// One CTA per (o, c) and batch chunk (blockIdx.y); each thread keeps the 25 taps and walks a strided subset of
// (n, pooled pixel) of its chunk. With one chunk the sums are stored (fixed order); with several they are added
// to a zeroed dw/db with fp32 atomics -- large batches would otherwise run on Co*Ci (18 for conv1) CTAs only.
__global__ void __launch_bounds__(256)
conv_relu_pool_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ y,
                            const unsigned char* __restrict__ idx, const float* __restrict__ dy,
                            float* __restrict__ dw, float* __restrict__ db, int N, int Ci, int H, int W,
                            int Co, int imgs_per_chunk) {
  const int o = blockIdx.x / Ci, c = blockIdx.x - o * Ci;
  const int Hp = (H - 4) / 2, Wp = (W - 4) / 2, per_img = Hp * Wp;
  float acc[25];
#pragma unroll
  for (int t = 0; t < 25; ++t) acc[t] = 0.f;
  float bsum = 0.f;
  const int n_lo = blockIdx.y * imgs_per_chunk, n_hi = min(N, n_lo + imgs_per_chunk);
  const long long total = (long long)(n_hi - n_lo) * per_img;
  for (long long t = threadIdx.x; t < total; t += blockDim.x) {
    const int n = n_lo + (int)(t / per_img), pp = (int)(t % per_img);
    const size_t oi = ((size_t)(n * Co + o)) * per_img + pp;
    const float g = y[oi] > 0.f ? dy[oi] : 0.f;
    if (g == 0.f) continue;
    bsum += g;
    const int id = idx[oi];
    const int ph = pp / Wp, pw = pp - ph * Wp;
    const float* xb = x + ((size_t)(n * Ci + c) * H + 2 * ph + (id >> 1)) * W + 2 * pw + (id & 1);
#pragma unroll
    for (int kh = 0; kh < 5; ++kh)
#pragma unroll
      for (int kw = 0; kw < 5; ++kw) acc[kh * 5 + kw] += g * __ldg(xb + kh * W + kw);
  }
  __shared__ float red[8][26];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int t = 0; t < 25; ++t) {
    const float s = warp_sum(acc[t]);
    if (lane == 0) red[warp][t] = s;
  }
  {
    const float s = warp_sum(bsum);
    if (lane == 0) red[warp][25] = s;
  }
  __syncthreads();
  if (threadIdx.x < 26) {
    float s = 0.f;
    for (int wv = 0; wv < (int)(blockDim.x >> 5); ++wv) s += red[wv][threadIdx.x];
    if (gridDim.y == 1) {
      if (threadIdx.x < 25) dw[(o * Ci + c) * 25 + threadIdx.x] = s;
      else if (c == 0) db[o] = s;
    } else {
      if (threadIdx.x < 25) atomicAdd(dw + (o * Ci + c) * 25 + threadIdx.x, s);
      else if (c == 0) atomicAdd(db + o, s);
    }
  }
}

// dgrad (K16): one thread per input element; visits the <= 25 conv outputs that saw it and
// takes the gradient only where that output was its window's argmax and passed the ReLU.
__global__ void __launch_bounds__(256)
conv_relu_pool_dgrad_kernel(const float* __restrict__ w, const float* __restrict__ y,
                            const unsigned char* __restrict__ idx, const float* __restrict__ dy,
                            float* __restrict__ dx, int N, int Ci, int H, int W, int Co) {
  const int Hc = H - 4, Wc = W - 4, Hp = Hc / 2, Wp = Wc / 2, per_img = Hp * Wp;
  const long long total = (long long)N * Ci * H * W;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int iw = (int)(t % W);
    const int ih = (int)((t / W) % H);
    const int c = (int)((t / ((long long)W * H)) % Ci);
    const int n = (int)(t / ((long long)W * H * Ci));
    float acc = 0.f;
    for (int o = 0; o < Co; ++o) {
      const size_t ob = ((size_t)(n * Co + o)) * per_img;
      const float* wo = w + ((size_t)(o * Ci + c)) * 25;
      for (int kh = 0; kh < 5; ++kh) {
        const int oh = ih - kh;
        if (oh < 0 || oh >= 2 * Hp) continue;
        for (int kw = 0; kw < 5; ++kw) {
          const int ow = iw - kw;
          if (ow < 0 || ow >= 2 * Wp) continue;
          const size_t oi = ob + (oh >> 1) * Wp + (ow >> 1);
          if (idx[oi] == (unsigned char)(((oh & 1) << 1) | (ow & 1)) && y[oi] > 0.f)
            acc += dy[oi] * __ldg(wo + kh * 5 + kw);
        }
      }
    }
    dx[t] = acc;
  }
}

// ---------------------------------------------------------------------------
// linear (+bias, +ReLU): y[n,m] = act(b[m] + x[n,:] . w[m,:])   (K8-K10)
// one warp per output element
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                  float* __restrict__ y, int N, int M, int K, int relu) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long t = warp0; t < (long long)N * M; t += nwarps) {
    const int n = (int)(t / M), m = (int)(t - (long long)n * M);
    const float* xr = x + (size_t)n * K;
    const float* wr = w + (size_t)m * K;
    float acc = 0.f;
    for (int k = lane; k < K; k += 32) acc += __ldg(xr + k) * __ldg(wr + k);
    acc = warp_sum(acc);
    if (lane == 0) {
      acc += b[m];
      y[t] = relu ? fmaxf(acc, 0.f) : acc;
    }
  }
}

// dx[n,k] = sum_m dy'[n,m] w[m,k]   with dy' = dy * [y > 0] when the layer had a ReLU (K13, K14)
__global__ void __launch_bounds__(256)
linear_dx_kernel(const float* __restrict__ w, const float* __restrict__ y, const float* __restrict__ dy,
                 float* __restrict__ dx, int N, int M, int K, int relu) {
  const long long total = (long long)N * K;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(t / K), k = (int)(t - (long long)n * K);
    float acc = 0.f;
    for (int m = 0; m < M; ++m) {
      float d = dy[(size_t)n * M + m];
      if (relu && !(y[(size_t)n * M + m] > 0.f)) d = 0.f;
      acc += d * __ldg(w + (size_t)m * K + k);
    }
    dx[t] = acc;
  }
}

// dw[m,k] = sum_n dy'[n,m] x[n,k];  db[m] = sum_n dy'[n,m]
// blockIdx.y = batch chunk: one chunk stores (fixed order), several add to a zeroed dw/db with fp32 atomics
__global__ void __launch_bounds__(256)
linear_dw_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ dy,
                 float* __restrict__ dw, float* __restrict__ db, int N, int M, int K, int relu, int rows_per_chunk) {
  const long long total = (long long)M * K;
  const int n_lo = blockIdx.y * rows_per_chunk, n_hi = min(N, n_lo + rows_per_chunk);
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int m = (int)(t / K), k = (int)(t - (long long)m * K);
    float acc = 0.f, bacc = 0.f;
    for (int n = n_lo; n < n_hi; ++n) {
      float d = dy[(size_t)n * M + m];
      if (relu && !(y[(size_t)n * M + m] > 0.f)) d = 0.f;
      acc += d * __ldg(x + (size_t)n * K + k);
      bacc += d;
    }
    if (gridDim.y == 1) {
      dw[t] = acc;
      if (k == 0) db[m] = bacc;
    } else {
      atomicAdd(dw + t, acc);
      if (k == 0) atomicAdd(db + m, bacc);
    }
  }
}

// ---------------------------------------------------------------------------
// softmax cross-entropy (mean) forward + backward in one pass (K11, K12)
// single CTA (32 warps), one warp per row, fixed summation order => deterministic loss
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
softmax_xent_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                    float* __restrict__ loss, float* __restrict__ dlogits, int N, int C) {
  __shared__ float s_part[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const float invN = 1.f / (float)N;
  float part = 0.f;
  for (int n = warp; n < N; n += nw) {
    const float* z = logits + (size_t)n * C;
    float m = -INFINITY;
    for (int c = lane; c < C; c += 32) m = fmaxf(m, z[c]);
    m = warp_max(m);
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) sum += __expf(z[c] - m);
    sum = warp_sum(sum);
    const int label = (int)target[n];
    for (int c = lane; c < C; c += 32)
      dlogits[(size_t)n * C + c] = (__expf(z[c] - m) / sum - (c == label ? 1.f : 0.f)) * invN;
    if (lane == 0) part += (m + __logf(sum)) - z[label];
  }
  if (lane == 0) s_part[warp] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int wv = 0; wv < nw; ++wv) a += s_part[wv];
    *loss = a * invN;
  }
}

// ---------------------------------------------------------------------------
// SGD with momentum on a flat buffer (K19): buf = mu*buf + g*scale ; p -= lr*buf
// first_step: buf = g*scale (torch.optim.SGD's first-step rule)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sgd_momentum_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf, int n,
                    float lr, float mu, int first_step, float scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float gv = g[i] * scale;
  const float b = first_step ? gv : mu * buf[i] + gv;
  buf[i] = b;
  p[i] -= lr * b;
}

// ---------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------
static inline int grid_for(long long work, int block, int cap = 148 * 8) {
  long long g = (work + block - 1) / block;
  if (g < 1) g = 1;
  if (g > cap) g = cap;
  return (int)g;
}

int launch_conv_relu_pool_fwd(const float* x, const float* w, const float* b, float* y, unsigned char* idx,
                              int N, int Ci, int H, int W, int Co, cudaStream_t s) {
  DNN_ARG_CHECK(H >= 6 && W >= 6 && ((H - 4) % 2 == 0) && ((W - 4) % 2 == 0), "conv_relu_pool: H,W must be even and >= 6");
  const int per_img = ((H - 4) / 2) * ((W - 4) / 2);
  dim3 grid(grid_for((long long)N * per_img, 256, 148 * 4), Co);
  conv_relu_pool_fwd_kernel<<<grid, 256, Ci * 25 * sizeof(float), s>>>(x, w, b, y, idx, N, Ci, H, W, Co);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_tc_conv2_dgrad(const float* w, const float* y, const unsigned char* idx, const float* dy, float* dx, int N,
                          cudaStream_t s);

int launch_tc_conv_wgrad(const float* x, const float* y, const unsigned char* idx, const float* dy, float* dw,
                         float* db, int N, int Ci, int H, int W, int Co, cudaStream_t s);

int launch_conv_relu_pool_bwd(const float* x, const float* w, const float* y, const unsigned char* idx,
                              const float* dy, float* dx, float* dw, float* db, int N, int Ci, int H, int W,
                              int Co, int tensor_core_dgrad, cudaStream_t s) {
  // TF32 mode, conv2, from a thousand images on: the tensor-core weight gradient is the faster one (189 vs 585 us at
  // 10 000 images, 33 vs 54 us at 1 024, 14 vs 11 us at 16: profiles/layer_bench_r2.json); for conv1 the two tie at
  // 10 000 images (999 vs 932 us) and the pooled-grid FFMA kernel wins below
  if (tensor_core_dgrad && Ci == 6 && H == 14 && W == 14 && Co == 16 && N >= 1024) {
    const int rc = launch_tc_conv_wgrad(x, y, idx, dy, dw, db, N, Ci, H, W, Co, s);
    if (rc) return rc;
  } else {
    // enough (o, c, batch chunk) CTAs for two waves; small batches keep one chunk and the fixed summation order
    int chunks = N >= 64 ? (2 * 148 + Co * Ci - 1) / (Co * Ci) : 1;
    if (chunks > N / 8) chunks = N / 8 > 0 ? N / 8 : 1;
    const int per = (N + chunks - 1) / chunks;
    chunks = (N + per - 1) / per;
    if (chunks > 1) {
      DNN_CUDA_CHECK(cudaMemsetAsync(dw, 0, (size_t)Co * Ci * 25 * sizeof(float), s));
      DNN_CUDA_CHECK(cudaMemsetAsync(db, 0, (size_t)Co * sizeof(float), s));
    }
    conv_relu_pool_wgrad_kernel<<<dim3(Co * Ci, chunks), 256, 0, s>>>(x, y, idx, dy, dw, db, N, Ci, H, W, Co, per);
    DNN_CUDA_CHECK(cudaGetLastError());
  }
  if (dx && tensor_core_dgrad && Ci == 6 && H == 14 && W == 14 && Co == 16)
    return launch_tc_conv2_dgrad(w, y, idx, dy, dx, N, s);   // tcgen05 implicit GEMM (TF32)
  if (dx) {
    conv_relu_pool_dgrad_kernel<<<grid_for((long long)N * Ci * H * W, 256), 256, 0, s>>>(w, y, idx, dy, dx, N, Ci,
                                                                                         H, W, Co);
    DNN_CUDA_CHECK(cudaGetLastError());
  }
  return kOk;
}

int launch_linear_fwd(const float* x, const float* w, const float* b, float* y, int N, int M, int K, int relu,
                      cudaStream_t s) {
  linear_fwd_kernel<<<grid_for((long long)N * M * 32, 256), 256, 0, s>>>(x, w, b, y, N, M, K, relu);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dx, float* dw,
                      float* db, int N, int M, int K, int relu, cudaStream_t s) {
  if (dx) {
    linear_dx_kernel<<<grid_for((long long)N * K, 256), 256, 0, s>>>(w, y, dy, dx, N, M, K, relu);
    DNN_CUDA_CHECK(cudaGetLastError());
  }
  {
    const int gx = grid_for((long long)M * K, 256);
    int chunks = N >= 256 ? (2 * 148 + gx - 1) / gx : 1;
    if (chunks > N / 32) chunks = N / 32 > 0 ? N / 32 : 1;
    const int per = (N + chunks - 1) / chunks;
    chunks = (N + per - 1) / per;
    if (chunks > 1) {
      DNN_CUDA_CHECK(cudaMemsetAsync(dw, 0, (size_t)M * K * sizeof(float), s));
      DNN_CUDA_CHECK(cudaMemsetAsync(db, 0, (size_t)M * sizeof(float), s));
    }
    linear_dw_kernel<<<dim3(gx, chunks), 256, 0, s>>>(x, y, dy, dw, db, N, M, K, relu, per);
    DNN_CUDA_CHECK(cudaGetLastError());
  }
  return kOk;
}

int launch_softmax_xent(const float* logits, const long long* target, float* loss, float* dlogits, int N, int C,
                        cudaStream_t s) {
  softmax_xent_kernel<<<1, N >= 256 ? 1024 : 256, 0, s>>>(logits, target, loss, dlogits, N, C);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_sgd_momentum(float* p, const float* g, float* buf, int n, float lr, float mu, int first_step,
                        float scale, cudaStream_t s) {
  sgd_momentum_kernel<<<(n + 255) / 256, 256, 0, s>>>(p, g, buf, n, lr, mu, first_step, scale);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

}  // namespace dnnb200
