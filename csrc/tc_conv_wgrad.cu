// conv5x5 weight (and bias) gradient on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators in
// TMEM): K16 / K17 of SURVEY section 2.7, the `convolution_backward` weight half of
// data_parallelism_train.py:198 for both conv layers of models/model.py:13-14.
//
//   dW[o][c][kh][kw] = sum over images and output pixels (oh, ow) of  dC[o][oh][ow] * X[c][oh+kh][ow+kw]
//
// is a GEMM whose REDUCTION index is the pixel: D[m][n] with m = (c, kh, kw) (75 rows per group of three input
// channels, padded to one 128-row tile), n = o (padded to 16), K = every output pixel of every image. The tensor core
// wants both operands K-major in 16-byte chunks, i.e. four consecutive PIXELS of one row m next to each other -- which
// is the transposed im2col matrix, a layout that exists nowhere (for fixed (c, kh, kw) consecutive pixels are 4 bytes
// apart in the image, but the eight rows of a core matrix must then be 16 bytes apart; see DESIGN.md section 4 for the
// MN-major alternatives that were ruled out). So it is materialised: per block of 112 output pixels the threads gather
//   sA[chunk][16 row groups][8 rows][4 pixels]   A[m][k] = X[c][oh+kh][ow+kw]   (+ a row of ones: its product is db)
//   sB[chunk][ 2 row groups][8 rows][4 pixels]   B[n][k] = dC[o][oh][ow] = dy at the argmax of its 2x2 window if the
//                                                 pooled output is positive (max-pool and ReLU backward), else 0
// into the canonical no-swizzle K-major form, one thread issues 14 MMAs (M=128, N=16, K=8) on the block, and the
// accumulators stay in tensor memory across ALL images a CTA processes; they are added to dw / db with fp32 atomics
// once, at the end. Persistent CTAs over the batch, the next image arrives by TMA bulk copy while the last block of
// the current one is multiplied.
//
// What it costs is in profiles/layer_bench_r2.json (10 000 images): conv2 189 us against 585 us for the pooled-grid FFMA
// kernel of layers.cu and 540 us for cuDNN given the dense gradient; conv1 999 us against 932 and 1 890. The autograd
// path uses it for conv2 from 1 024 images on (layers.cu: launch_conv_relu_pool_bwd); `native.tc_conv_wgrad` calls it
// directly. For conv1 the gather writes as many elements per image as the FFMA kernel executes multiply-adds and every
// K=8 slab of pixels costs a 47..64-cycle MMA whatever N is (bench/tc_mma_rate.py), so the two tie.
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace dnnb200 {

using namespace tc;

template <int CI, int HIN, int CO>
struct WgradCfg {
  static constexpr int WIN = HIN, HOUT = HIN - 4, WOUT = HIN - 4, HP = HOUT / 2;
  static constexpr int P = HOUT * WOUT;                    // output pixels per image = reduction length per image
  static constexpr int TILES = (CI + 2) / 3;               // 128-row tiles: three input channels (75 rows) each
  static constexpr int MT = 75;                            // real rows per tile; row 75 of tile 0 is all ones (-> db)
  static constexpr int GROUPS = 10;                        // 8-row groups that hold anything (rows 0..79)
  static constexpr int KB_PIX = 112;                       // pixels per block (28 chunks, 14 MMAs)
  static constexpr int KB_CH = KB_PIX / 4;
  static constexpr int BLOCKS = (P + KB_PIX - 1) / KB_PIX; // blocks per image
  static constexpr int NG = (CO + 7) / 8;                  // 8-row groups of B that hold anything
  static constexpr int STAGE_FLOATS = CI * HIN * WIN;
  static constexpr size_t kABytes = (size_t)KB_CH * 16 * 128;   // one tile's A block
  static constexpr size_t kBBytes = (size_t)KB_CH * 2 * 128;
  static constexpr size_t kSmemBytes = (size_t)STAGE_FLOATS * 4 + kABytes + kBBytes + 64;
  static_assert((STAGE_FLOATS * 4) % 16 == 0, "bulk copy granularity");
  static_assert(TILES * 16 <= 32, "TMEM columns");
};

template <int CI, int HIN, int CO>
__global__ void __launch_bounds__(256)
tc_conv_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ y, const unsigned char* __restrict__ idx,
                     const float* __restrict__ dy, float* __restrict__ dw, float* __restrict__ db, int N) {
  using C = WgradCfg<CI, HIN, CO>;
  extern __shared__ __align__(128) unsigned char smem[];
  float* stage = reinterpret_cast<float*>(smem);                       // CHW image as it is in global memory
  float* sA = stage + C::STAGE_FLOATS;
  float* sB = sA + C::kABytes / 4;
  __shared__ uint64_t bar_img, bar_mma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) { mbar_init(&bar_img, 1); mbar_init(&bar_mma, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc<32>(&tmem_slot);
  for (int i = tid; i < (int)((C::kABytes + C::kBBytes) / 16); i += 256)   // rows that never hold anything stay zero
    reinterpret_cast<float4*>(sA)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  constexpr uint32_t kIdesc = idesc_tf32(128, 16);
  int img = blockIdx.x;
  if (tid == 0 && img < N) {
    mbar_expect_tx(&bar_img, C::STAGE_FLOATS * 4);
    bulk_g2s(stage, x + (size_t)img * C::STAGE_FLOATS, C::STAGE_FLOATS * 4, &bar_img);
  }
  uint32_t img_phase = 0, mma_phase = 0;
  uint32_t started = 0u;   // bit t: tile t's accumulator has been written (its first MMA overwrites, the rest add)

  for (; img < N; img += gridDim.x) {
    mbar_wait(&bar_img, img_phase);
    img_phase ^= 1u;
    const float* yb = y + (size_t)img * CO * C::HP * C::HP;
    const float* dyb = dy + (size_t)img * CO * C::HP * C::HP;
    const unsigned char* ib = idx + (size_t)img * CO * C::HP * C::HP;
#pragma unroll 1
    for (int blk = 0; blk < C::BLOCKS; ++blk) {
      const int k0 = blk * C::KB_PIX;
#pragma unroll 1
      for (int t = 0; t < C::TILES; ++t) {
        // ---- A block of tile t. A thread keeps ONE row m = (c, kh, kw) for the whole kernel (its base address in the
        // image is computed once) and walks that row's K-chunks, three threads per row: four pixels per item, one
        // 16-byte store. (The first version did the index arithmetic per element: ncu showed the kernel bound by the
        // issue slots of this gather -- 57-69 % busy, tensor pipe 1.5 % -- not by its MMAs.)
        if (tid < 3 * C::GROUPS * 8) {
          const int m = tid % (C::GROUPS * 8), q0 = tid / (C::GROUPS * 8);
          const int cl = m / 25, tap = m - cl * 25, c = t * 3 + cl;
          const bool real = m < C::MT && c < CI, ones = m == C::MT && t == 0;
          const float* src = stage + (c * HIN + tap / 5) * C::WIN + tap % 5;   // (only dereferenced for real rows)
          for (int q = q0; q < C::KB_CH; q += 3) {
            float v[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int k = k0 + q * 4 + e, oh = k / C::WOUT, ow = k - oh * C::WOUT;
              v[e] = k < C::P ? (real ? src[oh * C::WIN + ow] : (ones ? 1.f : 0.f)) : 0.f;
            }
            *reinterpret_cast<float4*>(sA + (q * 16 + (m >> 3)) * 32 + (m & 7) * 4) = make_float4(v[0], v[1], v[2], v[3]);
          }
        }
        if (t == 0) {
          // ---- B block (shared by the tiles): dC of these pixels, from the pooled gradient and the argmax ----
          for (int it = tid; it < C::KB_CH * C::NG * 8; it += 256) {
            const int q = it / (C::NG * 8), o = it - q * (C::NG * 8);
            float v[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int k = k0 + q * 4 + e, oh = k / C::WOUT, ow = k - oh * C::WOUT;
              v[e] = 0.f;
              if (k < C::P && o < CO) {
                const int pi = (o * C::HP + (oh >> 1)) * C::HP + (ow >> 1);
                if ((int)__ldg(ib + pi) == (oh & 1) * 2 + (ow & 1) && __ldg(yb + pi) > 0.f) v[e] = __ldg(dyb + pi);
              }
            }
            *reinterpret_cast<float4*>(sB + (q * 2 + (o >> 3)) * 32 + (o & 7) * 4) = make_float4(v[0], v[1], v[2], v[3]);
          }
        }
        fence_proxy_async();   // operands were written with st.shared; the tensor core reads through the async proxy
        __syncthreads();
        if (tid == 0) {
          tc_fence_after();
          if (blk == C::BLOCKS - 1 && t == C::TILES - 1 && img + (int)gridDim.x < N) {
            // the image is not read any more: fetch the next one under these MMAs
            mbar_expect_tx(&bar_img, C::STAGE_FLOATS * 4);
            bulk_g2s(stage, x + (size_t)(img + gridDim.x) * C::STAGE_FLOATS, C::STAGE_FLOATS * 4, &bar_img);
          }
          const uint64_t a0 = smem_desc_nosw(smem_u32(sA), 16 * 128, 128);
          const uint64_t b0 = smem_desc_nosw(smem_u32(sB), 2 * 128, 128);
#pragma unroll
          for (int j = 0; j < C::KB_CH / 2; ++j)   // K-chunks 2j and 2j+1: 2 x 16 groups x 128 bytes further on in A
            mma_tf32(tmem + (uint32_t)t * 16, a0 + (uint64_t)(j * 2 * 16 * 128 / 16), b0 + (uint64_t)(j * 2 * 2 * 128 / 16),
                     kIdesc, ((started >> t) & 1u) != 0u || j > 0);
          mma_commit(&bar_mma);
        }
        started |= 1u << t;
        mbar_wait(&bar_mma, mma_phase);   // the block is consumed: the operand buffers may be rewritten
        mma_phase ^= 1u;
        tc_fence_after();
      }
    }
  }
  // ---- epilogue: this CTA's share of dW (and db, the row of ones) -> global, fp32 atomics ----
  if (blockIdx.x < N && warp < 4) {
#pragma unroll
    for (int t = 0; t < C::TILES; ++t) {
      float v[16];
      tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)t * 16, v);
      const int m = warp * 32 + lane;
      if (m < C::MT) {
        const int cl = m / 25, tap = m - cl * 25, c = t * 3 + cl;
        if (c < CI) {
#pragma unroll
          for (int o = 0; o < CO; ++o) atomicAdd(dw + ((size_t)o * CI + c) * 25 + tap, v[o]);
        }
      } else if (m == C::MT && t == 0) {
#pragma unroll
        for (int o = 0; o < CO; ++o) atomicAdd(db + o, v[o]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_free<32>(tmem);
}

template <int CI, int HIN, int CO>
static int launch_tc_conv_wgrad_t(const float* x, const float* y, const unsigned char* idx, const float* dy, float* dw,
                                  float* db, int N, cudaStream_t s) {
  using C = WgradCfg<CI, HIN, CO>;
  static PerDeviceOnce attr_set;
  auto kern = tc_conv_wgrad_kernel<CI, HIN, CO>;
  if (int rc = ensure_max_dynamic_smem(attr_set, kern, C::kSmemBytes)) return rc;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int per_sm = (int)(220 * 1024 / C::kSmemBytes) < 1 ? 1 : (int)(220 * 1024 / C::kSmemBytes);
  const int grid = N < sms * per_sm ? N : sms * per_sm;
  DNN_CUDA_CHECK(cudaMemsetAsync(dw, 0, (size_t)CO * CI * 25 * sizeof(float), s));
  DNN_CUDA_CHECK(cudaMemsetAsync(db, 0, (size_t)CO * sizeof(float), s));
  kern<<<grid, 256, C::kSmemBytes, s>>>(x, y, idx, dy, dw, db, N);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_tc_conv_wgrad(const float* x, const float* y, const unsigned char* idx, const float* dy, float* dw,
                         float* db, int N, int Ci, int H, int W, int Co, cudaStream_t s) {
  if (Ci == 3 && H == 32 && W == 32 && Co == 6) return launch_tc_conv_wgrad_t<3, 32, 6>(x, y, idx, dy, dw, db, N, s);
  if (Ci == 6 && H == 14 && W == 14 && Co == 16) return launch_tc_conv_wgrad_t<6, 14, 16>(x, y, idx, dy, dw, db, N, s);
  set_last_error("tc_conv_wgrad: only the CNN's two conv shapes are instantiated (got Ci=%d H=%d W=%d Co=%d)", Ci, H, W, Co);
  return kErrArg;
}

}  // namespace dnnb200
