// conv5x5 + bias + ReLU + maxpool2x2 forward as an implicit GEMM on the 5th-generation tensor
// cores (tcgen05.mma kind::tf32, accumulators in TMEM), for both conv layers of the CNN
// (reference: models/model.py:21-22; kernels K1-K6 of SURVEY §2.7).
//
// The im2col matrix is never built. The image is kept in shared memory as NHWC with the channels
// padded to groups of 4 floats, so one pixel of one channel group is exactly one 16-byte K-chunk
// of a K-major UMMA operand. With the GEMM row index m taken as the LINEAR pixel index of the
// input plane (m = oh*W_in + ow), the A operand of filter tap (kh,kw) is the same plane shifted by
// (kh*W_in + kw) pixels: 8 consecutive pixels are 8 rows 16 bytes apart (a core matrix), groups of
// 8 pixels are 128 bytes apart (SBO), and the two K-chunks of one MMA are two taps (conv1) or the two
// channel groups of one tap (conv2), LBO apart. Every MMA therefore just gets a different START
// ADDRESS into the same plane -- the shared-memory descriptor does the im2col. Rows with
// ow >= W_out are computed on wrapped pixels and discarded in the epilogue (12.5% / 49% of the tile
// for conv1 / conv2; the tensor pipe is nowhere near the bottleneck at these sizes).
//
//   GEMM per image   conv1: M = 896 (7 tiles of 128), N = 16 (6 used), K = 26 chunks x 4 = 104
//                    conv2: M = 256 (2 tiles),        N = 16,          K = 50 chunks x 4 = 200
//
// Pipeline per CTA (persistent over images): TMA bulk copy of the CHW image -> re-layout to NHWC4
// planes -> one thread issues all tcgen05.mma of the image -> tcgen05.commit -> mbarrier ->
// 4 warps tcgen05.ld their 32 TMEM lanes -> bias+ReLU -> shared tile -> 2x2 max/argmax -> global.
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace dnnb200 {

using namespace tc;

// ------------------------------------------------------------------------------------------
// probe: D[128][16] = A[128][K] * B[16][K]^T with every descriptor field exercised the way the
// conv kernel uses them (development aid + unit test of the encodings)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
tc_gemm_probe_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int K) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int kc = K / 4;                               // K-chunks of 16 bytes
  float* sA = reinterpret_cast<float*>(smem);         // [kc][16 groups][8 rows][4]
  float* sB = sA + (size_t)kc * 128 * 4;              // [kc][2 groups][8 rows][4]
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 128 * K; i += 128) {
    const int m = i / K, k = i - m * K;
    sA[((size_t)(k >> 2) * 16 + (m >> 3)) * 32 + (m & 7) * 4 + (k & 3)] = A[i];
  }
  for (int i = tid; i < 16 * K; i += 128) {
    const int n = i / K, k = i - n * K;
    sB[((size_t)(k >> 2) * 2 + (n >> 3)) * 32 + (n & 7) * 4 + (k & 3)] = B[i];
  }
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc<32>(&tmem_slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    const uint32_t idesc = idesc_tf32(128, 16);
    for (int ks = 0; ks < K / 8; ++ks) {
      const uint64_t ad = smem_desc_nosw(smem_u32(sA) + (uint32_t)(2 * ks) * 16 * 128, 16 * 128, 128);
      const uint64_t bd = smem_desc_nosw(smem_u32(sB) + (uint32_t)(2 * ks) * 2 * 128, 2 * 128, 128);
      mma_tf32(tmem, ad, bd, idesc, ks > 0);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  float v[16];
  tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16), v);
#pragma unroll
  for (int n = 0; n < 16; ++n) D[(warp * 32 + lane) * 16 + n] = v[n];
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_free<32>(tmem);
}

// ------------------------------------------------------------------------------------------
// the conv kernel
// ------------------------------------------------------------------------------------------
template <int CI, int HIN, int CO>
struct ConvCfg {
  static constexpr int G = (CI + 3) / 4;                       // channel groups of 4 floats (one K-chunk each)
  static constexpr int WIN = HIN;
  static constexpr int HOUT = HIN - 4, WOUT = HIN - 4, HP = HOUT / 2;
  static constexpr int P = HIN * WIN;                          // pixels of one input plane
  static constexpr int MEXT = (HOUT - 1) * WIN + WOUT;         // GEMM rows that contain a valid output
  static constexpr int TILES = (MEXT + 127) / 128;
  static constexpr int PLANE_PIX = ((TILES * 128 + 4 * WIN + 4 + 8 + 7) / 8) * 8;  // + reach of the last tap
  static constexpr int CHUNKS = 25 * G;                        // real K-chunks
  static constexpr int NMMA = (CHUNKS + 1) / 2;                // MMAs per tile (K = 2 chunks each)
  static constexpr int TMEM_COLS = TILES * 16 <= 32 ? 32 : (TILES * 16 <= 64 ? 64 : 128);
  static constexpr int STAGE_FLOATS = CI * P;
  static constexpr size_t kPlaneBytes = (size_t)PLANE_PIX * 16;
  static constexpr size_t kBBytes = (size_t)2 * (2 * NMMA) * 128;   // [2 n-groups][2*NMMA chunks][8][4] floats
  static constexpr size_t kSmemBytes = (size_t)STAGE_FLOATS * 4 + G * kPlaneBytes + kBBytes + (size_t)CO * HOUT * WOUT * 4 + 64;
  static_assert(TILES * 16 <= 128, "TMEM budget");
  static_assert((STAGE_FLOATS * 4) % 16 == 0, "bulk copy granularity");
};

template <int CI, int HIN, int CO>
__global__ void __launch_bounds__(128)
tc_conv_relu_pool_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                             float* __restrict__ y, unsigned char* __restrict__ idx, int N) {
  using C = ConvCfg<CI, HIN, CO>;
  extern __shared__ __align__(128) unsigned char smem[];
  float* stage = reinterpret_cast<float*>(smem);                                  // CHW image as it is in global
  float* planes = stage + C::STAGE_FLOATS;                                        // G planes of NHWC4 pixels
  float* sB = planes + (size_t)C::G * C::PLANE_PIX * 4;                           // weights, canonical K-major
  float* cbuf = sB + C::kBBytes / 4;                                              // relu(conv) of the image [CO][HOUT][WOUT]
  float* sbias = cbuf + CO * C::HOUT * C::WOUT;
  __shared__ uint64_t bar_img, bar_mma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) { mbar_init(&bar_img, 1); mbar_init(&bar_mma, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc<C::TMEM_COLS>(&tmem_slot);
  __syncthreads();
  int img = blockIdx.x;
  if (tid == 0 && img < N) {
    mbar_expect_tx(&bar_img, C::STAGE_FLOATS * 4);
    bulk_g2s(stage, x + (size_t)img * C::STAGE_FLOATS, C::STAGE_FLOATS * 4, &bar_img);
  }
  // weights -> canonical no-swizzle K-major B operand; zero the planes once (their padding stays zero)
  for (int i = tid; i < (int)(C::kBBytes / 4); i += 128) {
    const int c = i & 3, nr = (i >> 2) & 7, q = (i >> 5) % (2 * C::NMMA), ng = (i >> 5) / (2 * C::NMMA);
    const int n = ng * 8 + nr;
    float v = 0.f;
    if (n < CO && q < C::CHUNKS) {
      const int tap = q / C::G, ch = (q % C::G) * 4 + c;
      if (ch < CI) v = __ldg(w + ((size_t)n * CI + ch) * 25 + tap);
    }
    sB[i] = v;
  }
  for (int i = tid; i < C::G * C::PLANE_PIX * 4; i += 128) planes[i] = 0.f;
  if (tid < 16) sbias[tid] = tid < CO ? __ldg(b + tid) : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  constexpr uint32_t kIdesc = idesc_tf32(128, 16);
  uint32_t phase = 0;

  for (; img < N; img += gridDim.x) {
    mbar_wait(&bar_img, phase);
    // ---- CHW -> NHWC4 planes (one 16-byte pixel per channel group) ----
    for (int p = tid; p < C::P; p += 128) {
#pragma unroll
      for (int g = 0; g < C::G; ++g) {
        float4 v;
        v.x = stage[(g * 4 + 0) * C::P + p];
        v.y = (g * 4 + 1 < CI) ? stage[(g * 4 + 1) * C::P + p] : 0.f;
        v.z = (g * 4 + 2 < CI) ? stage[(g * 4 + 2) * C::P + p] : 0.f;
        v.w = (g * 4 + 3 < CI) ? stage[(g * 4 + 3) * C::P + p] : 0.f;
        reinterpret_cast<float4*>(planes)[(size_t)g * C::PLANE_PIX + p] = v;
      }
    }
    fence_proxy_async();   // the planes were written with st.shared; the tensor core reads them through the async proxy
    __syncthreads();
    if (tid == 0) {
      if (img + (int)gridDim.x < N) {  // staging is free again: prefetch the next image under the MMAs
        mbar_expect_tx(&bar_img, C::STAGE_FLOATS * 4);
        bulk_g2s(stage, x + (size_t)(img + gridDim.x) * C::STAGE_FLOATS, C::STAGE_FLOATS * 4, &bar_img);
      }
      // Issue loop kept to the bone (bench/tc_mma_rate.py: an MMA of this shape costs its issuing thread 64 cycles when
      // the thread does nothing else, 90+ with descriptor arithmetic in between): the two descriptor bases are built
      // once, every tap offset is a compile-time constant, per MMA two 64-bit adds remain.
      const uint64_t b0 = smem_desc_nosw(smem_u32(sB), 128, (uint32_t)(2 * C::NMMA) * 128);
      const uint64_t a_in = smem_desc_nosw(smem_u32(planes), C::G == 1 ? 16u : (uint32_t)C::kPlaneBytes, 128);
      const uint64_t a_x = smem_desc_nosw(smem_u32(planes), (uint32_t)(C::WIN - 4) * 16, 128);   // (G == 1) taps 2j, 2j+1 in different kernel rows
#pragma unroll 1
      for (int t = 0; t < C::TILES; ++t) {
        const uint64_t at = (uint64_t)(t * 128);
#pragma unroll
        for (int j = 0; j < C::NMMA; ++j) {
          // G == 1: K-chunks = taps 2j and 2j+1 of the single plane (chunk 25 does not exist: B is zero there);
          // G == 2: K-chunks = the two channel groups of tap j, a plane apart
          const int tap0 = C::G == 1 ? 2 * j : j;
          const int off0 = (tap0 / 5) * C::WIN + (tap0 % 5);
          const bool cross = C::G == 1 && (tap0 % 5) == 4 && tap0 + 1 < 25;
          mma_tf32(tmem + (uint32_t)t * 16, (cross ? a_x : a_in) + at + (uint64_t)off0, b0 + (uint64_t)(16 * j), kIdesc, j > 0);
        }
      }
      mma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, phase);
    tc_fence_after();
    // ---- epilogue 1: TMEM -> registers -> bias + ReLU -> shared [CO][HOUT][WOUT] ----
#pragma unroll 1
    for (int t = 0; t < C::TILES; ++t) {
      const int m = t * 128 + warp * 32 + lane;
      const int oh = m / C::WIN, ow = m - oh * C::WIN;
      const bool valid = oh < C::HOUT && ow < C::WOUT;
      const uint32_t ta = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)t * 16;
      if (CO <= 8) {
        float v[8];
        tmem_ld8(ta, v);
        if (valid) {
#pragma unroll
          for (int o = 0; o < CO; ++o) cbuf[(o * C::HOUT + oh) * C::WOUT + ow] = fmaxf(v[o] + sbias[o], 0.f);
        }
      } else {
        float v[16];
        tmem_ld16(ta, v);
        if (valid) {
#pragma unroll
          for (int o = 0; o < CO; ++o) cbuf[(o * C::HOUT + oh) * C::WOUT + ow] = fmaxf(v[o] + sbias[o], 0.f);
        }
      }
    }
    tc_fence_before();   // TMEM reads are done before the next image's MMAs may overwrite the accumulators
    __syncthreads();
    // ---- epilogue 2: 2x2 max + argmax (first maximum wins), to global NCHW ----
    for (int i = tid; i < CO * C::HP * C::HP; i += 128) {
      const int o = i / (C::HP * C::HP), pp = i - o * (C::HP * C::HP), ph = pp / C::HP, pw = pp - ph * C::HP;
      const float* c0 = cbuf + (o * C::HOUT + 2 * ph) * C::WOUT + 2 * pw;
      float best = c0[0];
      int bi = 0;
      if (c0[1] > best) { best = c0[1]; bi = 1; }
      if (c0[C::WOUT] > best) { best = c0[C::WOUT]; bi = 2; }
      if (c0[C::WOUT + 1] > best) { best = c0[C::WOUT + 1]; bi = 3; }
      const size_t oi = (size_t)img * CO * C::HP * C::HP + i;
      y[oi] = best;
      idx[oi] = (unsigned char)bi;
    }
    __syncthreads();     // cbuf and the planes are rewritten by the next image
    phase ^= 1u;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_free<C::TMEM_COLS>(tmem);
}

// ------------------------------------------------------------------------------------------
// conv2 data gradient (K16 dgrad) with ReLU'/maxpool' folded into the prologue (K14/K15):
//   dX[c][ih][iw] = sum_{o,kh,kw} dC[o][ih-kh][iw-kw] W[o][c][kh][kw]
// is a correlation of the zero-padded dense gradient plane dCp (18x18, border 4) with the FLIPPED
// filter: dX[ih][iw] = sum_{a,b} dCp[ih+a][iw+b] W[4-a][4-b]. So it is the same implicit GEMM as the
// forward: the 16 conv2 channels are four NHWC4 planes, GEMM row m = ih*18 + iw, the A operand of
// tap (a,b) is the plane shifted by a*18+b pixels, K = 25 taps x 4 planes = 100 chunks (50 MMAs per
// tile of 128 rows), N = 16 (6 input channels used). The prologue scatters the pooled gradient to the
// argmax position of each 2x2 window, masked by ReLU, directly into the planes.
// ------------------------------------------------------------------------------------------
struct DgradCfg {
  static constexpr int WP = 18;                           // padded plane width/height
  static constexpr int TILES = 2;                         // rows 0 .. 13*18+13 = 247
  static constexpr int PLANE_PIX = 2 * 128 + 4 * 18 + 4 + 12;   // reach of the last tap, multiple of 8 -> 344
  static constexpr int NMMA = 50;
  static constexpr size_t kPlaneBytes = (size_t)PLANE_PIX * 16;
  static constexpr size_t kBBytes = (size_t)2 * 100 * 128;
  static constexpr size_t kSmemBytes = 4 * kPlaneBytes + kBBytes;
};

__global__ void __launch_bounds__(128)
tc_conv2_dgrad_kernel(const float* __restrict__ w, const float* __restrict__ y, const unsigned char* __restrict__ idx,
                      const float* __restrict__ dy, float* __restrict__ dx, int N) {
  using C = DgradCfg;
  extern __shared__ __align__(128) unsigned char smem[];
  float* planes = reinterpret_cast<float*>(smem);                 // 4 planes of NHWC4 pixels
  float* sB = planes + 4 * C::PLANE_PIX * 4;                      // flipped weights, canonical K-major
  __shared__ uint64_t bar_mma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) { mbar_init(&bar_mma, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc<32>(&tmem_slot);
  // B[n = c][chunk q = tap*4 + g][k = channel o%4]: W[o = g*4+k][c][4-a][4-b] with tap = a*5+b
  for (int i = tid; i < (int)(C::kBBytes / 4); i += 128) {
    const int k = i & 3, nr = (i >> 2) & 7, q = (i >> 5) % 100, ng = (i >> 5) / 100;
    const int c = ng * 8 + nr, tap = q >> 2, g = q & 3, o = g * 4 + k;
    const int a = tap / 5, b = tap - a * 5;
    sB[i] = c < 6 ? __ldg(w + ((size_t)o * 6 + c) * 25 + (4 - a) * 5 + (4 - b)) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  constexpr uint32_t kIdesc = idesc_tf32(128, 16);
  uint32_t phase = 0;
  for (int img = blockIdx.x; img < N; img += gridDim.x) {
    for (int i = tid; i < 4 * C::PLANE_PIX; i += 128) reinterpret_cast<float4*>(planes)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    for (int i = tid; i < 400; i += 128) {   // maxpool' + ReLU': gradient goes to the argmax of its window
      const int o = i / 25, pp = i - o * 25, ph = pp / 5, pw = pp - ph * 5;
      const size_t gi = (size_t)img * 400 + i;
      const float g = y[gi] > 0.f ? dy[gi] : 0.f;
      const int id = idx[gi];
      const int pix = (2 * ph + (id >> 1) + 4) * C::WP + 2 * pw + (id & 1) + 4;
      planes[((size_t)(o >> 2) * C::PLANE_PIX + pix) * 4 + (o & 3)] = g;
    }
    fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
      // (lean issue loop: descriptor bases once, compile-time offsets, two 64-bit adds per MMA -- see the forward kernel)
      const uint64_t a0 = smem_desc_nosw(smem_u32(planes), (uint32_t)C::kPlaneBytes, 128);
      const uint64_t b0 = smem_desc_nosw(smem_u32(sB), 128, 100 * 128);
#pragma unroll 1
      for (int t = 0; t < C::TILES; ++t) {
        const uint64_t at = (uint64_t)(t * 128);
#pragma unroll
        for (int j = 0; j < C::NMMA; ++j) {
          const int tap = j >> 1, gp = (j & 1) * 2;   // this MMA: planes gp and gp+1 of tap
          const int off = (tap / 5) * C::WP + (tap % 5) + gp * C::PLANE_PIX;
          mma_tf32(tmem + (uint32_t)t * 16, a0 + at + (uint64_t)off, b0 + (uint64_t)(16 * j), kIdesc, j > 0);
        }
      }
      mma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, phase);
    tc_fence_after();
#pragma unroll 1
    for (int t = 0; t < C::TILES; ++t) {
      const int m = t * 128 + warp * 32 + lane, ih = m / C::WP, iw = m - ih * C::WP;
      float v[8];
      tmem_ld8(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)t * 16, v);
      if (ih < 14 && iw < 14) {
#pragma unroll
        for (int c = 0; c < 6; ++c) dx[((size_t)img * 6 + c) * 196 + ih * 14 + iw] = v[c];
      }
    }
    tc_fence_before();
    __syncthreads();
    phase ^= 1u;
  }
  if (warp == 0) tmem_free<32>(tmem);
}

int launch_tc_conv2_dgrad(const float* w, const float* y, const unsigned char* idx, const float* dy, float* dx, int N,
                          cudaStream_t s) {
  static PerDeviceOnce attr_set;
  if (int rc = ensure_max_dynamic_smem(attr_set, tc_conv2_dgrad_kernel, DgradCfg::kSmemBytes)) return rc;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = N < sms * 4 ? N : sms * 4;
  tc_conv2_dgrad_kernel<<<grid, 128, DgradCfg::kSmemBytes, s>>>(w, y, idx, dy, dx, N);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

template <int CI, int HIN, int CO>
static int launch_tc_conv_t(const float* x, const float* w, const float* b, float* y, unsigned char* idx, int N,
                            cudaStream_t s) {
  using C = ConvCfg<CI, HIN, CO>;
  static PerDeviceOnce attr_set;
  auto kern = tc_conv_relu_pool_fwd_kernel<CI, HIN, CO>;
  if (int rc = ensure_max_dynamic_smem(attr_set, kern, C::kSmemBytes)) return rc;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int per_sm = (int)(200 * 1024 / C::kSmemBytes) < 512 / C::TMEM_COLS ? (int)(200 * 1024 / C::kSmemBytes)
                                                                             : 512 / C::TMEM_COLS;
  const int grid = N < sms * per_sm ? N : sms * per_sm;
  kern<<<grid, 128, C::kSmemBytes, s>>>(x, w, b, y, idx, N);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_tc_conv_relu_pool_fwd(const float* x, const float* w, const float* b, float* y, unsigned char* idx, int N,
                                 int Ci, int H, int W, int Co, cudaStream_t s) {
  if (Ci == 3 && H == 32 && W == 32 && Co == 6) return launch_tc_conv_t<3, 32, 6>(x, w, b, y, idx, N, s);
  if (Ci == 6 && H == 14 && W == 14 && Co == 16) return launch_tc_conv_t<6, 14, 16>(x, w, b, y, idx, N, s);
  set_last_error("tc_conv_relu_pool_fwd: only the two conv layers of the CNN are instantiated (3x32x32->6, 6x14x14->16)");
  return kErrArg;
}

int launch_tc_gemm_probe(const float* A, const float* B, float* D, int K, cudaStream_t s) {
  DNN_ARG_CHECK(K >= 8 && K % 8 == 0 && K <= 256, "tc_gemm_probe: K must be a multiple of 8 in [8, 256]");
  const size_t smem = (size_t)(K / 4) * (128 + 16) * 16;
  static PerDeviceOnce attr_set;
  if (int rc = ensure_max_dynamic_smem(attr_set, tc_gemm_probe_kernel, 160 * 1024)) return rc;
  tc_gemm_probe_kernel<<<1, 128, smem, s>>>(A, B, D, K);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

}  // namespace dnnb200
