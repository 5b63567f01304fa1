// fwd_bwd_cluster_kernel: the forward/backward half of the training step with each
// IMAGE spread over a thread-block cluster of CS CTAs (CS = 2..8; one CTA per image is fused_step.cu).
//
// Why: at the reference's batch sizes (1..64 per rank, data_parallelism_train.py:263,
// run_training.sh) one CTA per image leaves most of the 148 SMs idle and makes the step a
// long serial chain on a few SMs. A cluster splits every layer of one image CS ways and
// re-assembles the small activations with distributed-shared-memory stores:
//
//   conv1+ReLU+pool   split by pooled rows      -> p1 pushed to every CTA of the cluster
//   conv2+ReLU+pool   split by output channel   -> p2 pushed
//   fc1 / fc2         split by output row       -> h1 / h2 pushed   (also splits the 192 KB weight stream)
//   fc3, softmax-CE, fc3 dX                      redundantly in every CTA (tiny)
//   fc2 dX / fc1 dX   split by input column     -> d1 / g2 pushed
//   conv2 wgrad       split by output channel   (needs only p1, g2: local)
//   conv2 dgrad       split by input pixel      -> g1 stays local because ...
//   conv1 wgrad       ... is split by the same pixel ranges; each CTA emits a partial
//
// The six all-gathers use st.async: every 4-byte DSMEM store also credits the destination
// CTA's mbarrier (complete_tx), so a consumer simply waits until the expected byte count has
// landed. No fence and no cluster-wide barrier sits between layers -- a first version with
// cluster.sync() spent 45% of its time in the barrier's release fence (profiles/). All math is fp32 FFMA, bit-compatible in structure with the
// one-CTA kernel of fused_step.cu (which stays as the CS-independent oracle and eval path).
//
// reference: models/model.py:20-27 forward; autograd of data_parallelism_train.py:195-198.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "layout.h"
#include "tc_common.cuh"

namespace cg = cooperative_groups;

namespace dnnb200 {

constexpr int kCLThreads = 512;
constexpr int kXRow = 36;                  // padded image row (floats): 16-byte aligned rows, and the five
constexpr int kXCh = 32 * kXRow + 8;       //   kernel rows of a wgrad gather fall into different banks


// stage barriers: one mbarrier per all-gather, completed by the st.async traffic itself
enum { kBarP1 = 0, kBarP2, kBarH1, kBarH2, kBarD1, kBarG2, kBarW1, kNumStageBars };

// kMode: 0 = fp32 FFMA throughout; 1 = conv1, conv2 and the conv2 data gradient on tcgen05 (TF32 operands);
//        2 = fp32 FFMA forward and weight gradients, ONLY the conv2 data gradient on tcgen05
template <int kMode>
struct CLSmemT;

template <>
struct __align__(16) CLSmemT<0> {
  alignas(16) float xd[kImgNumel];    // image as it is in global memory (ONE 12 KB TMA bulk copy)
  alignas(16) float x[3 * kXCh];      // image, CHW with padded rows (re-laid out by the threads)
  alignas(16) float cw[kConvPrefix];  // conv1.weight | conv1.bias | conv2.weight | conv2.bias exactly as in the flat
                                      // parameter vector (offsets kOffW1, kOffB1, kOffW2, kOffB2): one TMA bulk copy
  alignas(16) float w2d[16 * 25 * 8]; // conv2.weight as [(o,kh,kw)][c8] for the data gradient
  alignas(16) float p1[1176 + 8];
  alignas(16) float p2[400];
  alignas(16) float h1[120];
  alignas(16) float h2[88];
  alignas(16) float logits[16];
  alignas(16) float d3[16];
  alignas(16) float d2[88];
  alignas(16) float d1[120];
  alignas(16) float g2[400];
  alignas(16) float g1[1176 + 8];
  alignas(16) float dc2p[16 * 18 * 18];
  alignas(16) int idx1[1176 + 8];     // 2-bit argmax of the pooling windows (one word each: st.async granularity)
  alignas(16) int idx2[400];
  alignas(16) int poff[196];          // (2*ph)*kXRow + 2*pw of pooled pixel q
  alignas(8) unsigned long long mbar;                  // image bulk copies
  alignas(8) unsigned long long cbar;                  // conv weights bulk copy (once per launch)
  alignas(8) unsigned long long sbar[kNumStageBars];   // all-gather completion
};

// ---- tensor-core instantiation (kTC): geometry of the three implicit GEMMs ------------------------------------
// conv1 forward, conv2 forward and the conv2 data gradient run on tcgen05 (kind::tf32, fp32 accumulators in
// TMEM); everything else (FC chain, both weight gradients, softmax) stays fp32 FFMA. All three GEMMs use the
// trick of tc_conv.cu: the activation plane is NHWC with 4 floats per pixel, so one pixel of one channel group is
// one 16-byte K-chunk of a no-swizzle K-major operand and a filter tap is just a different START ADDRESS into the
// same plane. New here: the 8-row groups of a tile are laid along image ROWS (SBO = one plane row), so a 128-row
// tile is 16 conv rows x 8 columns and the four pixels of every 2x2 pooling window sit in ONE warp's TMEM lanes
// (lanes l, l^1, l^8, l^9): bias + ReLU + max-pool + argmax are register shuffles, no shared-memory round trip.
constexpr int kXPitch = 33;                            // conv1 plane: pixels per row (33, not 32: the five kernel
                                                       //   rows of a wgrad gather then fall into different banks)
constexpr int kXPlanePix = 36 * kXPitch + 12;          // rows 0..35 are reachable (tile rows 16..31 + 4) + tap reach
constexpr int kP1Pitch = 14;
constexpr int kP1PlanePix = 20 * kP1Pitch + 24;        // conv2 tiles reach rows 0..19, columns up to 19
constexpr int kDPitch = 18;                            // conv2 output gradient, zero border of 4
constexpr int kDPlanePix = 2 * 128 + 4 * kDPitch + 4 + 12;   // two tiles + reach of the last tap (344)
constexpr int kTmemCols = 128;                         // [0,64) conv1 tiles, [64,96) conv2 tiles, [96,128) dgrad tiles
constexpr uint32_t kTmemC2 = 64, kTmemDg = 96;
enum { kTBarC1 = 0, kTBarC2, kTBarDg, kNumTcBars };

template <>
struct __align__(128) CLSmemT<1> {
  alignas(128) float xpl[kXPlanePix * 4];        // image, NHWC4 (x, y, z, 0) per pixel, row pitch kXPitch: conv1 A operand
                                                 //   and the conv1 weight gradient's gather
  alignas(128) float dpl[4 * kDPlanePix * 4];    // dL/d(conv2 output), 4 planes of 4 channels, zero-padded 18x18: dgrad A
  alignas(128) float p1n[2 * kP1PlanePix * 4];   // pooled conv1 output, 2 planes of 4 channels (6, 7 stay zero): conv2 A.
                                                 //   The two low mantissa bits of every value carry the pooling argmax
  alignas(128) float sb1[2 * kB1Chunks * 32];    // conv1.weight as canonical K-major B operand [n-group][chunk][8][4]
  alignas(128) float sb2[2 * kB2Chunks * 32];    // conv2.weight, K = (tap, channel group)
  alignas(128) float sbd[2 * kBdChunks * 32];    // conv2.weight flipped and transposed for the data gradient, K = (tap, o)
  alignas(16) float xd[kImgNumel];               // image as it is in global memory (ONE 12 KB TMA bulk copy)
  alignas(16) float cw[kConvPrefix];
  alignas(16) float p2[400];
  alignas(16) float h1[120];
  alignas(16) float h2[88];
  alignas(16) float logits[16];
  alignas(16) float d3[16];
  alignas(16) float d2[88];
  alignas(16) float d1[120];
  alignas(16) float g2[400];
  alignas(16) float g1[1176 + 8];
  alignas(16) int idx2[400];
  alignas(16) int poff[196];          // (2*ph)*kXPitch + 2*pw of pooled pixel q
  alignas(8) unsigned long long mbar;
  alignas(8) unsigned long long cbar;
  alignas(8) unsigned long long sbar[kNumStageBars];
  alignas(8) unsigned long long tbar[kNumTcBars];      // tcgen05.commit targets
  alignas(8) unsigned long long pbar;                  // bulk copy of sb2 | sbd (packed operands from global memory)
  uint32_t tmem_slot;
};

// fp32 layout with the data gradient's tensor-core operands in place of its FFMA buffers (dc2p, w2d)
template <>
struct __align__(128) CLSmemT<2> {
  alignas(128) float dpl[4 * kDPlanePix * 4];    // dL/d(conv2 output), 4 planes of 4 channels, zero-padded 18x18: dgrad A
  alignas(128) float sbd[2 * kBdChunks * 32];    // conv2.weight flipped and transposed for the data gradient (B operand)
  alignas(16) float xd[kImgNumel];    // image as it is in global memory (ONE 12 KB TMA bulk copy)
  alignas(16) float x[3 * kXCh];      // image, CHW with padded rows (re-laid out by the threads)
  alignas(16) float cw[kConvPrefix];  // conv1.weight | conv1.bias | conv2.weight | conv2.bias exactly as in the flat
                                      // parameter vector (offsets kOffW1, kOffB1, kOffW2, kOffB2): one TMA bulk copy
  alignas(16) float p1[1176 + 8];
  alignas(16) float p2[400];
  alignas(16) float h1[120];
  alignas(16) float h2[88];
  alignas(16) float logits[16];
  alignas(16) float d3[16];
  alignas(16) float d2[88];
  alignas(16) float d1[120];
  alignas(16) float g2[400];
  alignas(16) float g1[1176 + 8];
  alignas(16) int idx1[1176 + 8];     // 2-bit argmax of the pooling windows (one word each: st.async granularity)
  alignas(16) int idx2[400];
  alignas(16) int poff[196];          // (2*ph)*kXRow + 2*pw of pooled pixel q
  alignas(8) unsigned long long mbar;                  // image bulk copies
  alignas(8) unsigned long long cbar;                  // conv weights bulk copy (once per launch)
  alignas(8) unsigned long long sbar[kNumStageBars];   // all-gather completion
  alignas(8) unsigned long long tbar[kNumTcBars];      // tcgen05.commit target (only kTBarDg is used)
  alignas(8) unsigned long long pbar;                  // bulk copy of sbd (packed operand from global memory)
  uint32_t tmem_slot;
};

__device__ __forceinline__ uint32_t cl_smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void cl_mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(cl_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void cl_mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(cl_smem_u32(bar)), "r"(bytes)
               : "memory");
}
// wait for a phase; acquire at cluster scope so data stored by peer CTAs (st.async) is visible
__device__ __forceinline__ void cl_mbar_wait(unsigned long long* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(cl_smem_u32(bar)),
      "r"(phase)
      : "memory");
}
__device__ __forceinline__ void cl_bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          cl_smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(cl_smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ uint32_t cl_mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// 4-byte store into a peer CTA's shared memory that also credits 4 bytes to that CTA's mbarrier:
// the consumer's wait completes when every expected byte has landed -- no fence, no cluster barrier.
__device__ __forceinline__ void cl_st_async(uint32_t remote_addr, uint32_t v, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(remote_addr),
               "r"(v), "r"(remote_bar)
               : "memory");
}
// 16 / 8 bytes at once (one pixel of an NHWC4 plane): same crediting of the destination's mbarrier
__device__ __forceinline__ void cl_st_async_v4(uint32_t remote_addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                               uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(
                   remote_addr),
               "r"(a), "r"(b), "r"(c), "r"(d), "r"(remote_bar)
               : "memory");
}
__device__ __forceinline__ void cl_st_async_v2(uint32_t remote_addr, uint32_t a, uint32_t b, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(
                   remote_addr),
               "r"(a), "r"(b), "r"(remote_bar)
               : "memory");
}
__device__ __forceinline__ void cl_cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
// no memory ordering needed (the mbarrier inits are published by fence.mbarrier_init): much cheaper
__device__ __forceinline__ void cl_cluster_arrive_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}
__device__ __forceinline__ void cl_cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// shared::cluster base address of CLSmem in every CTA of the cluster
template <int CS>
struct Remote {
  uint32_t base[CS];
};
// (used inside the kernel, where SM names the instantiation's CLSmemT)
#define CL_OFF(field) ((uint32_t)offsetof(SM, field))
#define CL_BAR(i) ((uint32_t)(offsetof(SM, sbar) + 8 * (i)))

template <int CS>
__device__ __forceinline__ void push_all(const Remote<CS>& R, uint32_t field_off, uint32_t bar_off, uint32_t v) {
#pragma unroll
  for (int d = 0; d < CS; ++d) cl_st_async(R.base[d] + field_off, v, R.base[d] + bar_off);
}

// rows [m0, m1) of y = act(b + W . in); each result is pushed to every CTA of the cluster
template <int CS, int K, bool kRelu>
__device__ __forceinline__ void fc_rows_push(const float* __restrict__ W, const float* __restrict__ b,
                                             const float* in, int m0, int m1, const Remote<CS>& R,
                                             uint32_t field_off, uint32_t bar_off) {
  constexpr int K4 = K / 4;
  constexpr int kWarps = kCLThreads / 32;
  constexpr int kIter = (K4 + 31) / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t my_dst = 0;  // lane d pushes to CTA d: select its base once, without divergent branches in the loop
#pragma unroll
  for (int d = 0; d < CS; ++d) my_dst = lane == d ? R.base[d] : my_dst;
  const float4* in4 = reinterpret_cast<const float4*>(in);
  for (int r0 = m0 + warp; r0 < m1; r0 += 2 * kWarps) {
    float4 w[2][kIter];
    float bias[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = r0 + q * kWarps;
      bias[q] = r < m1 ? __ldg(b + r) : 0.f;  // issued with the weights, not after the reduction
#pragma unroll
      for (int it = 0; it < kIter; ++it) {
        const int k4 = lane + it * 32;
        w[q][it] = (r < m1 && k4 < K4) ? __ldg(reinterpret_cast<const float4*>(W + (size_t)r * K) + k4)
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
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
      if (r < m1) {
        acc += bias[q];
        const float y = kRelu ? fmaxf(acc, 0.f) : acc;
        if (lane < CS) cl_st_async(my_dst + field_off + 4u * r, __float_as_uint(y), my_dst + bar_off);  // lane d -> CTA d
      }
    }
  }
}

// same rows, result kept in this CTA only (fc3 is computed redundantly by every CTA)
template <int M, int K>
__device__ __forceinline__ void fc_rows_local(const float* __restrict__ W, const float* __restrict__ b,
                                              const float* in, float* out) {
  constexpr int K4 = K / 4;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < M) {
    const float4 w = lane < K4 ? __ldg(reinterpret_cast<const float4*>(W + (size_t)warp * K) + lane)
                               : make_float4(0.f, 0.f, 0.f, 0.f);
    const float bias = __ldg(b + warp);
    float acc = 0.f;
    if (lane < K4) {
      const float4 v = reinterpret_cast<const float4*>(in)[lane];
      acc = w.x * v.x + w.y * v.y + w.z * v.z + w.w * v.w;
    }
    acc = warp_sum(acc);
    if (lane == 0) out[warp] = acc + bias;
  }
  static_assert(K4 <= 32 && M <= kCLThreads / 32, "fc_rows_local: one warp per row, one float4 per lane");
}

// dx[k] = [act[k] > 0] * sum_m W[m,k] dy[m] for the float4 columns [k4lo, k4hi); pushed to every CTA
template <int CS, int M, int K>
__device__ __forceinline__ void fc_dx_split_push(const float* __restrict__ W, const float* dy, const float* act,
                                                 int k4lo, int k4hi, float* red, const Remote<CS>& R,
                                                 uint32_t field_off, uint32_t bar_off) {
  constexpr int G = 32;  // row groups
  const int tid = threadIdx.x;
  const int nk4 = k4hi - k4lo;
  for (int t = tid; t < nk4 * G; t += kCLThreads) {
    const int g = t / nk4, j = t - g * nk4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int m = g; m < M; m += G) {
      const float4 w = __ldg(reinterpret_cast<const float4*>(W + (size_t)m * K) + k4lo + j);
      const float d = dy[m];
      acc.x += w.x * d; acc.y += w.y * d; acc.z += w.z * d; acc.w += w.w * d;
    }
    reinterpret_cast<float4*>(red)[g * nk4 + j] = acc;
  }
  __syncthreads();
  for (int t = tid; t < nk4 * 4; t += kCLThreads) {
    float s = 0.f;
#pragma unroll 8
    for (int g = 0; g < G; ++g) s += red[g * nk4 * 4 + t];
    const int k = k4lo * 4 + t;
    const float v = act[k] > 0.f ? s : 0.f;
    push_all<CS>(R, field_off + 4u * k, bar_off, __float_as_uint(v));
  }
}

template <int M, int K>
__device__ __forceinline__ void fc_dx_local(const float* __restrict__ W, const float* dy, const float* act,
                                            float* dx) {
  const int tid = threadIdx.x;
  if (tid < K) {
    float s = 0.f;
#pragma unroll
    for (int m = 0; m < M; ++m) s += __ldg(W + m * K + tid) * dy[m];
    dx[tid] = act[tid] > 0.f ? s : 0.f;
  }
}


// ---- FC weights resident in shared memory (clusters of 4 and 8) ---------------------------------
// The FC layers are a chain of dependent matrix-vector products; reading their weights from L2 at
// each link made the chain 41% of the kernel (profiles/). Each CTA therefore prefetches, with TMA
// bulk copies issued at kernel start and hidden under the conv layers, exactly the ROW slices of
// fc1/fc2 it owns (and all of the tiny fc3). The SAME slices serve the backward pass: a CTA
// multiplies its rows by its part of the incoming delta, which yields a partial sum over its rows
// for every input column; the CS partials are exchanged with st.async and added in CTA order.
template <int CS>
struct FCSmem {
  static constexpr int kR1 = (kF1 + CS - 1) / CS;
  static constexpr int kR2 = (kF2 + CS - 1) / CS;
  alignas(16) float wf1[kR1 * kFlat];
  alignas(16) float wf2[kR2 * kF1];
  alignas(16) float wf3[kF3 * kF2];
  alignas(16) float bf1[kF1];
  alignas(16) float bf2[kF2 + 4];
  alignas(16) float bf3[16];
  alignas(16) float rx_d1[CS * kF1];    // partial dL/dh1 from every CTA of the cluster
  alignas(16) float rx_g2[CS * kFlat];  // partial dL/dp2
  alignas(8) unsigned long long wbar;   // weight prefetch
};
template <int kMode>
__host__ __device__ constexpr size_t fc_off() {
  return (sizeof(CLSmemT<kMode>) + 127) & ~(size_t)127;
}
// conv1 weight-gradient partials of the CTAs of a cluster, collected by CTA 0 (row d = CTA d: 450 weights + 6 biases)
constexpr int kW1Row = 456;
// scratch for split reductions, sized for the cluster size: conv2 dgrad partials 16*6*nq + conv2 wgrad partials
// 5*no*150 live together; every other use (conv2 forward partials, conv1 wgrad partials, CS<4 FC splits) is smaller
// (tensor-core instantiation: the data gradient and the conv2 forward need no partials, only the two weight
// gradients do -- 5*no*150 and 5*450+30 floats; clusters below 4 still split the FC data gradients through it)
template <int CS, int kMode = 0>
struct W1Rx {
  static constexpr int kNq = (196 + CS - 1) / CS, kNo = (16 + CS - 1) / CS;
  static constexpr int kWg = 5 * kNo * 150 > 2304 ? 5 * kNo * 150 : 2304;        // the two weight gradients' partials
  static constexpr int kC2 = kNo * 600;                                          // FFMA conv2 forward partials
  static constexpr int kRed0 = kMode == 1 ? kWg : (kMode == 2 ? (kWg > kC2 ? kWg : kC2) : 96 * kNq + 5 * kNo * 150);
  static constexpr int kRedMin = (kMode != 0 && CS >= 4) ? 0 : 6400;
  static constexpr int kRed = kRed0 > kRedMin ? kRed0 : kRedMin;
  alignas(16) float red[kRed];
  alignas(16) float rx[CS * kW1Row];
};
template <int CS, int kMode = 0>
__host__ __device__ constexpr size_t w1rx_offset() {
  return CS >= 4 ? ((fc_off<kMode>() + sizeof(FCSmem<CS>) + 127) & ~(size_t)127) : fc_off<kMode>();
}
template <int CS, int kMode = 0>
__host__ __device__ constexpr size_t cluster_smem_bytes() {
  return w1rx_offset<CS, kMode>() + sizeof(W1Rx<CS, kMode>);
}
static_assert(cluster_smem_bytes<4>() <= 227 * 1024 && cluster_smem_bytes<2>() <= 227 * 1024 &&
              cluster_smem_bytes<8>() <= 227 * 1024, "shared memory budget");
static_assert(cluster_smem_bytes<2, 1>() <= 227 * 1024 && cluster_smem_bytes<3, 1>() <= 227 * 1024 &&
              cluster_smem_bytes<4, 1>() <= 227 * 1024 && cluster_smem_bytes<5, 1>() <= 227 * 1024 &&
              cluster_smem_bytes<6, 1>() <= 227 * 1024 && cluster_smem_bytes<7, 1>() <= 227 * 1024 &&
              cluster_smem_bytes<8, 1>() <= 227 * 1024, "shared memory budget (tensor-core instantiation)");
static_assert(cluster_smem_bytes<2, 2>() <= 227 * 1024 && cluster_smem_bytes<3, 2>() <= 227 * 1024 &&
              cluster_smem_bytes<4, 2>() <= 227 * 1024 && cluster_smem_bytes<5, 2>() <= 227 * 1024 &&
              cluster_smem_bytes<6, 2>() <= 227 * 1024 && cluster_smem_bytes<7, 2>() <= 227 * 1024 &&
              cluster_smem_bytes<8, 2>() <= 227 * 1024, "shared memory budget (tensor-core data gradient only)");

// rows [m0, m1) of y = act(b + W . in) with the rows in shared memory (wsm[0] is row m0)
template <int CS, int K, bool kRelu>
__device__ __forceinline__ void fc_rows_smem_push(const float* wsm, const float* bsm, const float* in, int m0, int m1,
                                                  const Remote<CS>& R, uint32_t field_off, uint32_t bar_off,
                                                  long long* dbg = nullptr) {
  constexpr int K4 = K / 4;
  constexpr int kIter = (K4 + 31) / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t my_dst = 0;
#pragma unroll
  for (int d = 0; d < CS; ++d) my_dst = lane == d ? R.base[d] : my_dst;
  const float4* in4 = reinterpret_cast<const float4*>(in);
  int dbg_n = 0;
  for (int r = m0 + warp; r < m1; r += kCLThreads / 32) {
    if (dbg && threadIdx.x == 0 && dbg_n < 2) dbg[dbg_n * 4 + 0] = clock64();
    const float4* w4 = reinterpret_cast<const float4*>(wsm + (size_t)(r - m0) * K);
    float acc = 0.f;
#pragma unroll
    for (int it = 0; it < kIter; ++it) {
      const int k4 = lane + it * 32;
      if (k4 < K4) {
        const float4 w = w4[k4], v = in4[k4];
        acc += w.x * v.x + w.y * v.y + w.z * v.z + w.w * v.w;
      }
    }
    if (dbg && threadIdx.x == 0 && dbg_n < 2) dbg[dbg_n * 4 + 1] = clock64();
    acc = warp_sum(acc) + bsm[r];
    const float y = kRelu ? fmaxf(acc, 0.f) : acc;
    if (dbg && threadIdx.x == 0 && dbg_n < 2) dbg[dbg_n * 4 + 2] = clock64();
    if (lane < CS) cl_st_async(my_dst + field_off + 4u * r, __float_as_uint(y), my_dst + bar_off);  // lane d -> CTA d
    if (dbg && threadIdx.x == 0 && dbg_n < 2) dbg[dbg_n * 4 + 3] = clock64();
    ++dbg_n;
  }
}

// partial[k] = sum over my rows m of W[m,k] * dy[m], pushed into slot `rank` of every CTA's rx buffer
template <int CS, int K>
__device__ __forceinline__ void fc_dx_partial_push(const float* wsm, const float* dy, int m0, int m1, int rank,
                                                   const Remote<CS>& R, uint32_t rx_off, uint32_t bar_off) {
  for (int k = threadIdx.x; k < K; k += kCLThreads) {
    float a0 = 0.f, a1 = 0.f;
    int m = m0;
    for (; m + 1 < m1; m += 2) {
      a0 += wsm[(size_t)(m - m0) * K + k] * dy[m];
      a1 += wsm[(size_t)(m + 1 - m0) * K + k] * dy[m + 1];
    }
    if (m < m1) a0 += wsm[(size_t)(m - m0) * K + k] * dy[m];
    push_all<CS>(R, rx_off + 4u * (uint32_t)(rank * K + k), bar_off, __float_as_uint(a0 + a1));
  }
}

// MMAs J0..J1-1 of an implicit-GEMM tile, everything but the two descriptor bases a compile-time constant: per MMA
// the issuing thread executes two 64-bit adds and the instruction. A operand: tap (kh, kw) = the plane shifted by
// kh*PITCH + kw pixels (16 bytes each). PLANE2 == 0: MMA j is tap j (its two K-chunks are the descriptor's LBO apart);
// PLANE2 != 0: MMA j is tap j/2 and odd j reads the second pair of planes, PLANE2 pixels further on.
template <int J0, int J1, int PITCH, int PLANE2>
__device__ __forceinline__ void tc_issue_taps(uint32_t d_tmem, uint64_t a0, uint64_t b0, uint32_t idesc) {
#pragma unroll
  for (int j = J0; j < J1; ++j) {
    const int tap = PLANE2 ? (j >> 1) : j;
    const int off = (tap / 5) * PITCH + (tap % 5) + (PLANE2 ? (j & 1) * PLANE2 : 0);
    tc::mma_tf32(d_tmem, a0 + (uint64_t)off, b0 + (uint64_t)(16 * j), idesc, j > J0);
  }
}

// development aid: thread 0 of every CTA stamps the SM clock at stage boundaries (trace == nullptr in production)
#define CL_TRACE(k) do { if (trace != nullptr && tid == 0) trace[blockIdx.x * 32 + (k)] = clock64(); } while (0)

// Register budget (experiment switch): 512 threads x 96 registers would leave a quarter of the register file
// free, exactly one CTA of the update kernel at 256 threads x 64, so that the kernels of consecutive steps can
// share an SM. (__maxnreg__ and __launch_bounds__ are mutually exclusive.)
#ifndef DNN_F_MAXNREG
#define DNN_F_MAXNREG 0
#endif
template <int CS, int kMode>
#if DNN_F_MAXNREG > 0
__global__ void __maxnreg__(DNN_F_MAXNREG)
#else
__global__ void __launch_bounds__(kCLThreads, 1)
#endif
fwd_bwd_cluster_kernel(const float* __restrict__ params, const float* __restrict__ x,
                       const long long* __restrict__ labels, float* __restrict__ saved,
                       float* __restrict__ part1, float* __restrict__ part2, float* __restrict__ img_loss,
                       int* __restrict__ img_correct, int B, float inv_batch, long long* __restrict__ trace,
                       const unsigned int* __restrict__ ready_ctr, unsigned int ready_expect,
                       const float* __restrict__ packed, const unsigned int* __restrict__ in_flag, unsigned int in_expect) {
  static_assert(CS >= 2 && CS <= 8, "the one-CTA case is fused_step.cu's fwd_bwd_kernel");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr bool kTC = kMode == 1;    // conv1, conv2 forward on tcgen05 (and the layouts that go with it)
  constexpr bool kTCD = kMode != 0;   // conv2 data gradient on tcgen05
  using SM = CLSmemT<kMode>;
  SM& s = *reinterpret_cast<SM*>(smem_raw);
  constexpr size_t kFcOff = fc_off<kMode>();
  cg::cluster_group cluster = cg::this_cluster();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (trace != nullptr && tid == 0) trace[blockIdx.x * 32 + 14] = (long long)globaltimer_ns();
  CL_TRACE(16);
  const int r = (int)cluster.block_rank();
  const int n_clusters = gridDim.x / CS;
  int img = blockIdx.x / CS;

  constexpr bool kPre = CS >= 4;  // FC weight slices resident in shared memory
  FCSmem<kPre ? CS : 4>& f = *reinterpret_cast<FCSmem<kPre ? CS : 4>*>(smem_raw + kFcOff);
  constexpr uint32_t kRxD1 = (uint32_t)(kFcOff + offsetof(FCSmem<kPre ? CS : 4>, rx_d1));
  constexpr uint32_t kRxG2 = (uint32_t)(kFcOff + offsetof(FCSmem<kPre ? CS : 4>, rx_g2));
  using Tail = W1Rx<CS, kMode>;
  Tail& tail = *reinterpret_cast<Tail*>(smem_raw + w1rx_offset<CS, kMode>());
  constexpr uint32_t kRxW1 = (uint32_t)(w1rx_offset<CS, kMode>() + offsetof(Tail, rx));
  const float* rx_w1 = tail.rx;
  float* const sred = tail.red;

  Remote<CS> R;
#pragma unroll
  for (int d = 0; d < CS; ++d) R.base[d] = cl_mapa(cl_smem_u32(&s), (uint32_t)d);

  if (tid == 0) {
    cl_mbar_init(&s.mbar, 1);
    cl_mbar_init(&s.cbar, 1);
#pragma unroll
    for (int i = 0; i < kNumStageBars; ++i) cl_mbar_init(&s.sbar[i], 1);
    if (kPre) cl_mbar_init(&f.wbar, 1);
    if constexpr (kTCD) {   // one arrival per issuing thread: conv1 one per tile of mine, conv2 and the data gradient two
      cl_mbar_init(&s.tbar[kTBarC1], kTC ? (8 - r + CS - 1) / CS : 1);
      cl_mbar_init(&s.tbar[kTBarC2], 2);
      cl_mbar_init(&s.tbar[kTBarDg], 2);
      cl_mbar_init(&s.pbar, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  uint32_t tmem = 0;
  if constexpr (kTC) {
    // tensor memory for the accumulators of the three GEMMs (one warp allocates, the same warp frees at the end);
    // and everything in the operand buffers that is padding is zeroed ONCE, long before any weight or image arrives:
    // B-operand rows/chunks without a weight, the pad column/rows of the image plane, channels 6..7 of p1
    // (shared-memory stores run at 128 bytes per cycle: zeroing all 96 KB of operand buffers cost 800 cycles of
    // every launch. Zeroed now is only what a VALID output row can read and nothing else writes.)
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    if (packed == nullptr) {   // B operands are packed in this kernel: their padding must be zero beforehand
      for (int i = tid; i < (int)(sizeof(s.sb1) + sizeof(s.sb2) + sizeof(s.sbd)) / 16; i += kCLThreads)
        reinterpret_cast<float4*>(s.sb1)[i] = z;   // sb1 | sb2 | sbd are contiguous (static_assert below)
    }
    // image plane: the pad column (the 26th, weightless K-chunk of conv1 reads one pixel past tap 24) and the rows
    // below the image (only rows of the tile that are thrown away read them, but NaN patterns are best kept out)
    for (int i = tid; i < 32; i += kCLThreads) reinterpret_cast<float4*>(s.xpl)[i * kXPitch + 32] = z;
    for (int i = 32 * kXPitch + tid; i < kXPlanePix; i += kCLThreads) reinterpret_cast<float4*>(s.xpl)[i] = z;
    for (int i = tid; i < 2 * kP1PlanePix; i += kCLThreads) reinterpret_cast<float4*>(s.p1n)[i] = z;
    static_assert(offsetof(SM, sb2) == offsetof(SM, sb1) + sizeof(s.sb1) &&
                  offsetof(SM, sbd) == offsetof(SM, sb2) + sizeof(s.sb2), "B operands are zeroed as one block");
  } else if constexpr (kTCD) {
    // rows 8..15 of the data gradient's B operand never hold a weight (always zeroed here); rows 0..7 arrive ready-made
    // or, when this kernel packs them itself, need their padding (channels 6, 7) zeroed too
    for (int i = tid + (packed != nullptr ? (int)sizeof(s.sbd) / 32 : 0); i < (int)sizeof(s.sbd) / 16; i += kCLThreads)
      reinterpret_cast<float4*>(s.sbd)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncthreads();
  if constexpr (kTCD) {
    // (allocation takes a few hundred cycles: warp 1 does it while the others go on to the image and the dependency
    // wait; the address is picked up after the block barrier at the top of the image loop)
    if (warp == 1) {
      tc::tmem_alloc<kTmemCols>(&s.tmem_slot);
      tc::tc_fence_before();
    }
  }
  // tell the cluster "my barriers exist"; the matching wait sits just before the first remote store
  CL_TRACE(17);
  cl_cluster_arrive_relaxed();
  CL_TRACE(18);

  uint32_t phase = 0;
  bool first_img = true;
  // One bulk copy per image. (Copying the 96 rows separately into the padded layout was tried first:
  // 96 tiny cp.async.bulk serialise in the TMA unit and cost 4.3 us per launch -- see profiles/.)
  auto load_image = [&](int im) {
    if (tid == 0) {
      cl_mbar_expect_tx(&s.mbar, kImgNumel * 4);
      cl_bulk_g2s(s.xd, x + (size_t)im * kImgNumel, kImgNumel * 4, &s.mbar);
    }
  };
  // dense CHW image -> padded rows (36 floats, 16-byte aligned) for the conv1 patch loads
  auto relayout_image = [&]() {
    if constexpr (kTC) {   // dense CHW -> NHWC4 pixels (x, y, z, 0) at row pitch kXPitch: the conv1 A operand
      for (int px = tid; px < 1024; px += kCLThreads) {
        const float4 v = make_float4(s.xd[px], s.xd[1024 + px], s.xd[2048 + px], 0.f);
        *reinterpret_cast<float4*>(s.xpl + ((px >> 5) * kXPitch + (px & 31)) * 4) = v;
      }
    } else {
      for (int i = tid; i < kImgNumel / 4; i += kCLThreads) {
        const float4 v = reinterpret_cast<const float4*>(s.xd)[i];
        const int e = i * 4, c = e >> 10, h = (e >> 5) & 31, w = e & 31;
        *reinterpret_cast<float4*>(s.x + c * kXCh + h * kXRow + w) = v;
      }
    }
  };
  // dense dL/d(conv2 output), zero-padded: rewritten (scattered into) for every image, so zeroed for every image
  auto zero_grad_plane = [&]() {
    if constexpr (kTCD) {
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int i = tid; i < 4 * kDPlanePix; i += kCLThreads) reinterpret_cast<float4*>(s.dpl)[i] = z;
    } else {
      for (int i = tid; i < 16 * 18 * 18; i += kCLThreads) s.dc2p[i] = 0.f;
    }
  };
  // Host-fed steps: the batch is copied in by the copy engine on ANOTHER stream, and instead of a stream event in front
  // of this kernel (which would end the programmatic overlap with its predecessor and the chaining of steps) the copy
  // stream writes a sequence number behind the data; the one thread that fetches images waits for it here.
  if (in_flag != nullptr && tid == 0) {
    while ((int)(ld_acquire_sys(in_flag) - in_expect) < 0) {}
    asm volatile("fence.proxy.async;" ::: "memory");   // the bulk copies below read through the async proxy
  }
  load_image(img);   // the input batch does not depend on the previous kernel of the stream,
  if (img < B) {     // ... so the first image is landed AND re-laid out before the dependency is even looked at
    cl_mbar_wait(&s.mbar, phase);
    relayout_image();
  }
  // but the weights do: they are written by the previous step's update kernel. Two ways to wait for it:
  //  * griddepcontrol.wait -- until that GRID has completed and flushed. With peers, its completion includes the
  //    acknowledgement of its NVLink stores: measured 3.3 us after its last instruction on the rank that pushed last.
  //  * (steps chained inside one engine call, all weights read by TMA) spin on the counter the update kernel's
  //    last CTA releases once every CTA has stored its parameters; the grid's drain then overlaps this kernel.
  // Only thread 0 reads weights (bulk copies; everybody else sees them through the mbarriers they complete), so
  // only thread 0 waits, and only it lets the successor be scheduled: when the update kernel of THIS step starts,
  // everything launched before this kernel is complete and visible (it stages its control block before its wait).
  if (ready_ctr != nullptr) {
    if (tid == 0) {
      while ((int)(ld_acquire_gpu(ready_ctr) - ready_expect) < 0) {}
      asm volatile("fence.proxy.async;" ::: "memory");  // the bulk copies below read through the async proxy
      pdl_trigger();
    }
  } else {
    pdl_wait();
    if (tid == 0) pdl_trigger();
  }
  if (tid == 0) {
    // every weight this CTA needs arrives by TMA bulk copy, issued before anything else: the conv block now,
    // (clusters >= 4) my FC row slices + the FC biases, which land while the conv layers run
    if constexpr (kTC) {
      if (packed != nullptr) {   // conv weights as ready-made B operands: conv1's with the first barrier, the rest later
        cl_mbar_expect_tx(&s.cbar, kConvPrefix * 4 + (uint32_t)sizeof(s.sb1));
        cl_bulk_g2s(s.sb1, packed + kPackB1, (uint32_t)sizeof(s.sb1), &s.cbar);
        cl_bulk_g2s(s.cw, params, kConvPrefix * 4, &s.cbar);   // (biases)
        cl_mbar_expect_tx(&s.pbar, (uint32_t)(sizeof(s.sb2) + sizeof(s.sbd)));
        cl_bulk_g2s(s.sb2, packed + kPackB2, (uint32_t)(sizeof(s.sb2) + sizeof(s.sbd)), &s.pbar);
      } else {
        cl_mbar_expect_tx(&s.cbar, kConvPrefix * 4);
        cl_bulk_g2s(s.cw, params, kConvPrefix * 4, &s.cbar);
      }
    } else {
      cl_mbar_expect_tx(&s.cbar, kConvPrefix * 4);
      cl_bulk_g2s(s.cw, params, kConvPrefix * 4, &s.cbar);
    }
    if (kPre) {
      const int a0 = r * kF1 / CS, a1 = (r + 1) * kF1 / CS, c0 = r * kF2 / CS, c1 = (r + 1) * kF2 / CS;
      const uint32_t n1 = (uint32_t)(a1 - a0) * kFlat * 4, n2 = (uint32_t)(c1 - c0) * kF1 * 4, n3 = kF3 * kF2 * 4;
      cl_mbar_expect_tx(&f.wbar, n1 + n2 + n3 + kF1 * 4 + kF2 * 4 + 48);
      cl_bulk_g2s(f.wf1, params + kOffF1W + (size_t)a0 * kFlat, n1, &f.wbar);
      cl_bulk_g2s(f.wf2, params + kOffF2W + (size_t)c0 * kF1, n2, &f.wbar);
      cl_bulk_g2s(f.wf3, params + kOffF3W, n3, &f.wbar);
      cl_bulk_g2s(f.bf1, params + kOffF1B, kF1 * 4, &f.wbar);
      cl_bulk_g2s(f.bf2, params + kOffF2B, kF2 * 4, &f.wbar);
      cl_bulk_g2s(f.bf3, params + kOffF3B, 48, &f.wbar);  // 10 biases + 2 floats of the (zero) alignment gap
    }
    if constexpr (kTCD && !kTC) {
      if (packed != nullptr) {
        // the data gradient's B operand, ready-made: needed two thirds into the kernel, so it is fetched behind
        // everything the forward pass waits for; only rows 0..7 hold weights (six input channels), rows 8..15 were zeroed
        cl_mbar_expect_tx(&s.pbar, (uint32_t)sizeof(s.sbd) / 2);
        cl_bulk_g2s(s.sbd, packed + kPackBd, (uint32_t)sizeof(s.sbd) / 2, &s.pbar);
      }
    }
  }

  if (tid < 196) s.poff[tid] = (2 * (tid / 14)) * (kTC ? kXPitch : kXRow) + 2 * (tid % 14);
  CL_TRACE(19);
  CL_TRACE(20);

  // static work split of this CTA
  const int o0 = r * 16 / CS, o1 = (r + 1) * 16 / CS;            // conv2 / conv2 wgrad: output channels
  const int f1m0 = r * kF1 / CS, f1m1 = (r + 1) * kF1 / CS;      // fc1 rows
  const int f2m0 = r * kF2 / CS, f2m1 = (r + 1) * kF2 / CS;      // fc2 rows
  const int d1k0 = r * (kF1 / 4) / CS, d1k1 = (r + 1) * (kF1 / 4) / CS;      // fc2 dX float4 columns (of 30)
  const int g2k0 = r * (kFlat / 4) / CS, g2k1 = (r + 1) * (kFlat / 4) / CS;  // fc1 dX float4 columns (of 100)
  const int q0 = r * 196 / CS, q1 = (r + 1) * 196 / CS;          // conv2 dgrad / conv1 wgrad: pooled pixels
  const int nq = q1 - q0;

  CL_TRACE(0);
  cl_cluster_wait();  // every CTA of the cluster is running and has initialised its barriers
  CL_TRACE(1);

  for (; img < B; img += n_clusters) {
    // the label is only needed by the softmax, six stages from here: fetch it now, off that stage's critical path
    const int label = (int)__ldg(labels + img);
    if (tid == 0) {  // arm this image's all-gathers: bytes each one will deliver into THIS CTA
      cl_mbar_expect_tx(&s.sbar[kBarP1], kTC ? 1176 * 4 : 1176 * 8);   // (kTC: the argmax rides in the value's low bits)
      if (!kTC) cl_mbar_expect_tx(&s.sbar[kBarP2], 400 * 8);
      else if ((r ^ 1) < CS)   // kTC: paired CTAs compute one conv2 tile each; I receive the OTHER tile's pooled outputs:
        cl_mbar_expect_tx(&s.sbar[kBarP2], ((r & 1) ? 16 * 5 * 4 : 16 * 5 * 1) * 8);   // columns 0..3 (tile 0) or 4 (tile 1)
      cl_mbar_expect_tx(&s.sbar[kBarH1], kF1 * 4);
      cl_mbar_expect_tx(&s.sbar[kBarH2], kF2 * 4);
      cl_mbar_expect_tx(&s.sbar[kBarD1], (kPre ? CS : 1) * kF1 * 4);
      cl_mbar_expect_tx(&s.sbar[kBarG2], (kPre ? CS : 1) * kFlat * 4);
      if (r == 0) cl_mbar_expect_tx(&s.sbar[kBarW1], CS * kW1Row * 4);
    }
    if (!first_img) {
      cl_mbar_wait(&s.mbar, phase);  // image landed
      relayout_image();
    }
    cl_mbar_wait(&s.cbar, 0);      // conv weights landed (first image; complete ever after)
    __syncthreads();
    if constexpr (kTCD) {
      tc::tc_fence_after();
      tmem = s.tmem_slot;
    }
    CL_TRACE(2);

    if constexpr (kTC) {
      // ================= conv1 and conv2 on the tensor cores =================
      // What a tiny MMA costs (bench/tc_mma_rate.py): M=128, N=16, K=8 takes ~64 cycles from one issuing thread and
      // ~47 from two or more (the tensor pipe re-reads 4 KB of A for 16k MACs), whatever N or the layout. So the
      // stages are sized in MMAs: each issuing thread does as little else as possible (a 64-bit add per descriptor,
      // the offsets are compile-time constants), and two threads issue wherever there are two independent tiles.
      constexpr uint32_t kIdesc = tc::idesc_tf32(128, 16);
      const int n_tile1 = (8 - r + CS - 1) / CS;   // conv1: 8 tiles per image (2 row blocks x 4 column blocks), dealt round-robin
      // conv2: two tiles (columns 0..7 / 8..15). Paired CTAs (r, r^1) compute one each and hand the result to the
      // partner; a CTA without partner (last one of an odd cluster) computes both.
      const bool paired = (r ^ 1) < CS;
      const bool pack_here = first_img && packed == nullptr;   // no ready-made operands: pack the weights in here
      if (pack_here) {
        // conv1.weight -> B operand (450 real entries, the rest stays zero): element (n, chunk = tap, c)
        for (int i = tid; i < 450; i += kCLThreads) {
          const int n = i / 75, rem = i - n * 75, c = rem / 25, tap = rem - c * 25;
          s.sb1[((n >> 3) * kB1Chunks + tap) * 32 + (n & 7) * 4 + c] = s.cw[kOffW1 + i];
        }
      }
      tc::fence_proxy_async();   // plane and weights were written with st.shared; the tensor core reads through the async proxy
      __syncthreads();
      if (lane == 0 && warp < n_tile1) {           // one issuing thread per tile of mine
        tc::tc_fence_after();
        const int t = r + warp * CS;
        const uint32_t a_addr = tc::smem_u32(s.xpl) + (uint32_t)((16 * (t >> 2)) * kXPitch + 8 * (t & 3)) * 16;
        // K-chunks of MMA j = taps 2j and 2j+1: 16 bytes apart inside a kernel row, (pitch - 4) pixels across rows
        const uint64_t a_in = tc::smem_desc_nosw(a_addr, 16, kXPitch * 16);
        const uint64_t a_x = tc::smem_desc_nosw(a_addr, (kXPitch - 4) * 16, kXPitch * 16);
        const uint64_t b0 = tc::smem_desc_nosw(tc::smem_u32(s.sb1), 128, kB1Chunks * 128);
        const uint32_t d = tmem + (uint32_t)warp * 16;
#pragma unroll
        for (int j = 0; j < kB1Chunks / 2; ++j) {
          const int t0 = 2 * j;
          const int off0 = (t0 / 5) * kXPitch + (t0 % 5);                // compile-time after unrolling
          const bool cross = (t0 % 5) == 4;                              // tap 2j+1 starts the next kernel row
          tc::mma_tf32(d, (cross ? a_x : a_in) + (uint64_t)off0, b0 + (uint64_t)(16 * j), kIdesc, j > 0);
        }                                                                // (tap 25 does not exist: its B chunk is zero)
        tc::mma_commit(reinterpret_cast<uint64_t*>(&s.tbar[kTBarC1]));
        if (warp == 0) CL_TRACE(30);
      }
      if (pack_here) {
        // While the tensor core works: the other two B operands (needed from conv2 on). A warp writes 32 consecutive
        // floats of an operand per step -- (row n%8, component) of one K-chunk -- so the stores are conflict-free.
        //   sb2: conv2.weight, chunk = 2*tap + ch/4, element (n = o, c = ch%4)
        //   sbd: the same weights for the data gradient dX[ch] = sum_{tap,o} dCp[o](shifted by tap) W[o][ch][24 - tap]:
        //        n = ch, chunk = 4*tap' + o/4 with tap' = 24 - tap, component o%4
        // (warps 4..15: the first four may be issuing conv1; four items at a time, loads ahead of the stores)
        if (warp >= 4) {
          for (int it0 = warp - 4; it0 < 200; it0 += 4 * 12) {
            float w[4];
            int dst[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int it = it0 + 12 * u;
              dst[u] = -1;
              w[u] = 0.f;
              if (it < 100) {
                const int og = it / 50, rem = it - og * 50, tap = rem >> 1, cg = rem & 1;
                const int o = og * 8 + (lane >> 2), ch = cg * 4 + (lane & 3);
                if (ch < 6) { dst[u] = (og * kB2Chunks + rem) * 32 + lane; w[u] = s.cw[kOffW2 + (o * 6 + ch) * 25 + tap]; }
              } else if (it < 200) {
                const int v = it - 100, tp = v >> 2, og = v & 3;     // tp = flipped tap, og = group of 4 output channels
                const int ch = lane >> 2, o = og * 4 + (lane & 3);
                if (ch < 6) {
                  dst[u] = (int)(sizeof(s.sb2) / 4) + (4 * tp + og) * 32 + lane;   // sbd follows sb2
                  w[u] = s.cw[kOffW2 + (o * 6 + ch) * 25 + (24 - tp)];
                }
              }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
              if (dst[u] >= 0) s.sb2[dst[u]] = w[u];
          }
        }
        tc::fence_proxy_async();
      }
      CL_TRACE(3);
      cl_mbar_wait(&s.tbar[kTBarC1], phase);
      tc::tc_fence_after();
      {
        // epilogue: lane = (conv row 4*(warp%4) + lane/8 of the tile, column lane%8); 2x2 windows by shuffles. The
        // top-left lane of a window sends the pooled pixel of one NHWC4 plane with ONE vector st.async per CTA.
        // With at most two tiles the 16 warps split the planes (channels 0-3 / 4-5); with more, a warp does both.
        int slot, c0, nc;
        if (n_tile1 <= 2) { slot = (warp >> 2) & 1; c0 = (warp >> 3) * 4; nc = (warp >> 3) ? 2 : 4; }
        else { slot = warp >> 2; c0 = 0; nc = 6; }
        if (slot < n_tile1) {
          const int t = r + slot * CS;
          float v[8];
          tc::tmem_ld8(tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)slot * 16, v);
          const int row = 16 * (t >> 2) + 4 * (warp & 3) + (lane >> 3), col = 8 * (t & 3) + (lane & 7);
          const int ph = row >> 1, pw = col >> 1;
          const bool mine = !(lane & 1) && !((lane >> 3) & 1) && ph < 14 && pw < 14;   // top-left lane of a valid window
          uint32_t word[6];
#pragma unroll
          for (int c = 0; c < 6; ++c) {
            word[c] = 0u;
            if (c < c0 || c >= c0 + nc) continue;   // warp-uniform
            const float a00 = fmaxf(v[c] + s.cw[kOffB1 + c], 0.f);
            const float a01 = __shfl_xor_sync(0xffffffffu, a00, 1);
            const float a10 = __shfl_xor_sync(0xffffffffu, a00, 8);
            const float a11 = __shfl_xor_sync(0xffffffffu, a00, 9);
            float best = a00;
            uint32_t bi = 0;
            if (a01 > best) { best = a01; bi = 1; }
            if (a10 > best) { best = a10; bi = 2; }
            if (a11 > best) { best = a11; bi = 3; }
            // the argmax rides in the two low mantissa bits (far below the TF32 rounding of the value's inputs)
            word[c] = (__float_as_uint(best) & ~3u) | bi;
          }
          if (mine) {
            const uint32_t px = (uint32_t)(ph * 14 + pw) * 16;
            if (c0 == 0) {
#pragma unroll
              for (int d = 0; d < CS; ++d)
                cl_st_async_v4(R.base[d] + CL_OFF(p1n) + px, word[0], word[1], word[2], word[3], R.base[d] + CL_BAR(kBarP1));
            }
            if (c0 + nc == 6) {
#pragma unroll
              for (int d = 0; d < CS; ++d)
                cl_st_async_v2(R.base[d] + CL_OFF(p1n) + (uint32_t)kP1PlanePix * 16 + px, word[4], word[5],
                               R.base[d] + CL_BAR(kBarP1));
            }
          }
        }
      }
      tc::tc_fence_before();
      if (pack_here) __syncthreads();   // sb2 / sbd are complete (and fenced) before anybody issues conv2
      CL_TRACE(4);
      if (lane == 0 && warp < 2) {
        if (packed != nullptr) cl_mbar_wait(&s.pbar, 0);   // sb2 | sbd have landed (long ago; stays complete)
        cl_mbar_wait(&s.sbar[kBarP1], phase);      // every CTA's part of p1 has landed (st.async from the whole cluster)
        tc::fence_proxy_async();
        tc::tc_fence_after();
        // paired: my tile (columns 8*(r&1)..), its 25 taps split over the two issuing threads and two accumulators
        // (added in the epilogue); alone: issuing thread w computes tile w
        const int cb = paired ? (r & 1) : warp;
        const uint64_t a0 = tc::smem_desc_nosw(tc::smem_u32(s.p1n) + (uint32_t)(8 * cb) * 16, kP1PlanePix * 16, kP1Pitch * 16);
        const uint64_t b0 = tc::smem_desc_nosw(tc::smem_u32(s.sb2), 128, kB2Chunks * 128);
        const uint32_t d = tmem + kTmemC2 + (uint32_t)warp * 16;
        // K-chunks of MMA j = the two channel groups of tap j
        if (!paired) tc_issue_taps<0, 25, kP1Pitch, 0>(d, a0, b0, kIdesc);
        else if (warp == 0) tc_issue_taps<0, 13, kP1Pitch, 0>(d, a0, b0, kIdesc);
        else tc_issue_taps<13, 25, kP1Pitch, 0>(d, a0, b0, kIdesc);
        tc::mma_commit(reinterpret_cast<uint64_t*>(&s.tbar[kTBarC2]));
        if (warp == 0) CL_TRACE(31);
      }
      cl_mbar_wait(&s.tbar[kTBarC2], phase);
      tc::tc_fence_after();
      cl_mbar_wait(&s.sbar[kBarP1], phase);        // (complete by now) acquire: later stages read p1 with ordinary loads
      CL_TRACE(5);
      {
        // 16 warps = 4 lane quarters x 2 halves of the channels x (paired: nothing more / alone: the 2 tiles)
        const int half = (warp >> 2) & 1, cb = paired ? (r & 1) : (warp >> 3);
        if (paired ? warp < 8 : true) {
          float v[8];
          const uint32_t lanes = (uint32_t)((warp & 3) * 32) << 16;
          if (paired) {
            float u[8];
            tc::tmem_ld8(tmem + lanes + kTmemC2 + (uint32_t)(half * 8), v);
            tc::tmem_ld8(tmem + lanes + kTmemC2 + 16 + (uint32_t)(half * 8), u);
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] += u[c];
          } else {
            tc::tmem_ld8(tmem + lanes + kTmemC2 + (uint32_t)(cb * 16 + half * 8), v);
          }
          const int row = 4 * (warp & 3) + (lane >> 3), col = 8 * cb + (lane & 7);
          const int ph = row >> 1, pw = col >> 1;
          const bool mine = !(lane & 1) && !((lane >> 3) & 1) && ph < 5 && pw < 5;
          const uint32_t partner = R.base[paired ? (r ^ 1) : r];
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int o = half * 8 + c;
            const float a00 = fmaxf(v[c] + s.cw[kOffB2 + o], 0.f);
            const float a01 = __shfl_xor_sync(0xffffffffu, a00, 1);
            const float a10 = __shfl_xor_sync(0xffffffffu, a00, 8);
            const float a11 = __shfl_xor_sync(0xffffffffu, a00, 9);
            float best = a00;
            int bi = 0;
            if (a01 > best) { best = a01; bi = 1; }
            if (a10 > best) { best = a10; bi = 2; }
            if (a11 > best) { best = a11; bi = 3; }
            if (mine) {
              const int oi = o * 25 + ph * 5 + pw;
              s.p2[oi] = best;
              s.idx2[oi] = bi;
              if (paired) {   // my tile's share of p2 for the partner, which computed the other tile
                cl_st_async(partner + CL_OFF(p2) + 4u * (uint32_t)oi, __float_as_uint(best), partner + CL_BAR(kBarP2));
                cl_st_async(partner + CL_OFF(idx2) + 4u * (uint32_t)oi, (uint32_t)bi, partner + CL_BAR(kBarP2));
              }
            }
          }
        }
      }
      tc::tc_fence_before();
      if (paired) cl_mbar_wait(&s.sbar[kBarP2], phase);   // the partner's tile has landed
      __syncthreads();   // p2 / idx2 complete in this CTA
    } else {
      // ================= conv1 + bias + ReLU + maxpool, my pooled pixels =================
      // (split by pooled PIXEL range, not by row: 196 pixels divide evenly over any cluster size)
      {
        const int npx = nq;
        for (int item = tid; item < npx * 6; item += kCLThreads) {
          const int ch = item / npx, px = item - ch * npx;
          const int ph = (q0 + px) / 14, pw = (q0 + px) % 14;
          float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f;
  #pragma unroll 1
          for (int c = 0; c < 3; ++c) {
            const float* xb = s.x + c * kXCh + (2 * ph) * kXRow + 2 * pw;
            float patch[6][6];
  #pragma unroll
            for (int rr = 0; rr < 6; ++rr)
  #pragma unroll
              for (int q = 0; q < 3; ++q) {
                const float2 v = *reinterpret_cast<const float2*>(xb + rr * kXRow + 2 * q);
                patch[rr][2 * q] = v.x;
                patch[rr][2 * q + 1] = v.y;
              }
            const float* wb = s.cw + kOffW1 + (ch * 3 + c) * 25;
  #pragma unroll
            for (int kh = 0; kh < 5; ++kh)
  #pragma unroll
              for (int kw = 0; kw < 5; ++kw) {
                const float w = wb[kh * 5 + kw];
                a00 += w * patch[kh][kw];
                a01 += w * patch[kh][kw + 1];
                a10 += w * patch[kh + 1][kw];
                a11 += w * patch[kh + 1][kw + 1];
              }
          }
          const float bias = s.cw[kOffB1 + ch];
          float best = fmaxf(a00 + bias, 0.f);
          int bi = 0;
          float v = fmaxf(a01 + bias, 0.f);
          if (v > best) { best = v; bi = 1; }
          v = fmaxf(a10 + bias, 0.f);
          if (v > best) { best = v; bi = 2; }
          v = fmaxf(a11 + bias, 0.f);
          if (v > best) { best = v; bi = 3; }
          const uint32_t oi = (uint32_t)(ch * 196 + ph * 14 + pw);
          push_all<CS>(R, CL_OFF(p1) + 4u * oi, CL_BAR(kBarP1), __float_as_uint(best));
          push_all<CS>(R, CL_OFF(idx1) + 4u * oi, CL_BAR(kBarP1), (uint32_t)bi);
        }
      }
      if constexpr (!kTCD) {
        if (first_img) {  // conv2 weights in data-gradient layout [(o,tap)][c8]: built once, off the critical path
          for (int ot = tid; ot < 400; ot += kCLThreads) {
            const int o = ot / 25, t = ot - o * 25;
            const float* src = s.cw + kOffW2 + o * 150 + t;
            reinterpret_cast<float4*>(s.w2d)[ot * 2] = make_float4(src[0], src[25], src[50], src[75]);
            reinterpret_cast<float4*>(s.w2d)[ot * 2 + 1] = make_float4(src[100], src[125], 0.f, 0.f);
          }
        }
      } else {
        if (first_img && packed == nullptr) {
          // no ready-made operand (first step of an engine call): pack the data gradient's B operand here, off the
          // critical path -- n = input channel, K-chunk = 4*(flipped tap) + o/4, component o%4; a warp writes the 32
          // consecutive floats of one chunk
          for (int u = warp; u < 100; u += kCLThreads / 32) {
            const int tp = u >> 2, og = u & 3, ch = lane >> 2, o = og * 4 + (lane & 3);
            if (ch < 6) s.sbd[(4 * tp + og) * 32 + lane] = s.cw[kOffW2 + (o * 6 + ch) * 25 + (24 - tp)];
          }
          tc::fence_proxy_async();   // (a block barrier follows long before the data gradient reads it)
        }
      }
      CL_TRACE(3);
      cl_mbar_wait(&s.sbar[kBarP1], phase);
      CL_TRACE(4);

      // ================= conv2 + bias + ReLU + maxpool, my output channels =================
      {
        const int no = o1 - o0;
        for (int item = tid; item < no * 150; item += kCLThreads) {
          const int pp = item % 25, oc = item / 25, c = oc % 6, ol = oc / 6;
          const int ph = pp / 5, pw = pp - ph * 5;
          const float* xb = s.p1 + c * 196 + (2 * ph) * 14 + 2 * pw;
          float patch[6][6];
  #pragma unroll
          for (int rr = 0; rr < 6; ++rr)
  #pragma unroll
            for (int q = 0; q < 3; ++q) {
              const float2 v = *reinterpret_cast<const float2*>(xb + rr * 14 + 2 * q);
              patch[rr][2 * q] = v.x;
              patch[rr][2 * q + 1] = v.y;
            }
          const float* wb = s.cw + kOffW2 + ((o0 + ol) * 6 + c) * 25;
          float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f;
  #pragma unroll
          for (int kh = 0; kh < 5; ++kh)
  #pragma unroll
            for (int kw = 0; kw < 5; ++kw) {
              const float w = wb[kh * 5 + kw];
              a00 += w * patch[kh][kw];
              a01 += w * patch[kh][kw + 1];
              a10 += w * patch[kh + 1][kw];
              a11 += w * patch[kh + 1][kw + 1];
            }
          *reinterpret_cast<float4*>(sred + item * 4) = make_float4(a00, a01, a10, a11);
        }
        __syncthreads();
        for (int t = tid; t < no * 25; t += kCLThreads) {
          const int ol = t / 25, pp = t - ol * 25, o = o0 + ol;
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  #pragma unroll
          for (int c = 0; c < 6; ++c) {
            const float4 q = *reinterpret_cast<const float4*>(sred + ((ol * 6 + c) * 25 + pp) * 4);
            v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
          }
          const float bias = s.cw[kOffB2 + o];
          float best = fmaxf(v.x + bias, 0.f);
          int bi = 0;
          float u = fmaxf(v.y + bias, 0.f);
          if (u > best) { best = u; bi = 1; }
          u = fmaxf(v.z + bias, 0.f);
          if (u > best) { best = u; bi = 2; }
          u = fmaxf(v.w + bias, 0.f);
          if (u > best) { best = u; bi = 3; }
          const uint32_t oi = (uint32_t)(o * 25 + pp);
          push_all<CS>(R, CL_OFF(p2) + 4u * oi, CL_BAR(kBarP2), __float_as_uint(best));
          push_all<CS>(R, CL_OFF(idx2) + 4u * oi, CL_BAR(kBarP2), (uint32_t)bi);
        }
      }
    CL_TRACE(5);
      cl_mbar_wait(&s.sbar[kBarP2], phase);
    }
    CL_TRACE(6);

    // ================= fc1, fc2 (rows split), fc3 + softmax-CE (redundant) =================
    if (kPre) {
      cl_mbar_wait(&f.wbar, 0);  // weight slices have landed (completed long ago; stays complete)
      fc_rows_smem_push<CS, kFlat, true>(f.wf1, f.bf1, s.p2, f1m0, f1m1, R, CL_OFF(h1), CL_BAR(kBarH1),
                                         trace ? trace + blockIdx.x * 32 + 22 : nullptr);
      cl_mbar_wait(&s.sbar[kBarH1], phase);
      CL_TRACE(7);
      fc_rows_smem_push<CS, kF1, true>(f.wf2, f.bf2, s.h1, f2m0, f2m1, R, CL_OFF(h2), CL_BAR(kBarH2));
      cl_mbar_wait(&s.sbar[kBarH2], phase);
      CL_TRACE(8);
      if (warp < kF3) {  // fc3 redundantly in every CTA: one warp per logit
        const float4 w = lane < kF2 / 4 ? reinterpret_cast<const float4*>(f.wf3 + warp * kF2)[lane]
                                        : make_float4(0.f, 0.f, 0.f, 0.f);
        float acc = 0.f;
        if (lane < kF2 / 4) {
          const float4 v = reinterpret_cast<const float4*>(s.h2)[lane];
          acc = w.x * v.x + w.y * v.y + w.z * v.z + w.w * v.w;
        }
        acc = warp_sum(acc);
        if (lane == 0) s.logits[warp] = acc + f.bf3[warp];
      }
    } else {
      fc_rows_push<CS, kFlat, true>(params + kOffF1W, params + kOffF1B, s.p2, f1m0, f1m1, R, CL_OFF(h1),
                                    CL_BAR(kBarH1));
      cl_mbar_wait(&s.sbar[kBarH1], phase);
      CL_TRACE(7);
      fc_rows_push<CS, kF1, true>(params + kOffF2W, params + kOffF2B, s.h1, f2m0, f2m1, R, CL_OFF(h2), CL_BAR(kBarH2));
      cl_mbar_wait(&s.sbar[kBarH2], phase);
      CL_TRACE(8);
      fc_rows_local<kF3, kF2>(params + kOffF3W, params + kOffF3B, s.h2, s.logits);
    }
    __syncthreads();
    if (warp == 0) {
      const float z = lane < 10 ? s.logits[lane] : -INFINITY;
      const float m = warp_max(z);
      const float e = lane < 10 ? __expf(z - m) : 0.f;
      const float sum = warp_sum(e);
      const float lse = m + __logf(sum);
      const float zl = __shfl_sync(0xffffffffu, z, label & 31);
      const unsigned int hit = __ballot_sync(0xffffffffu, lane < 10 && z == m);
      const int pred = __ffs(hit) - 1;
      if (lane < 10) s.d3[lane] = (e / sum - (lane == label ? 1.f : 0.f)) * inv_batch;
      if (lane == 0 && r == 0) {
        img_loss[img] = lse - zl;
        if (img_correct) img_correct[img] = (pred == label) ? 1 : 0;
      }
    }
    __syncthreads();

    // ================= FC data gradients =================
    if (kPre) {
      if (tid < kF2) {  // dL/dh2, redundantly: 10 x 84 from shared memory
        float a = 0.f;
#pragma unroll
        for (int m = 0; m < kF3; ++m) a += f.wf3[m * kF2 + tid] * s.d3[m];
        s.d2[tid] = s.h2[tid] > 0.f ? a : 0.f;
      }
      __syncthreads();
      // dL/dh1: partial sums over MY fc2 rows for all 120 columns, exchanged and added in CTA order
      fc_dx_partial_push<CS, kF1>(f.wf2, s.d2, f2m0, f2m1, r, R, kRxD1, CL_BAR(kBarD1));
      cl_mbar_wait(&s.sbar[kBarD1], phase);
      if (tid < kF1) {
        float a = 0.f;
#pragma unroll
        for (int d = 0; d < CS; ++d) a += f.rx_d1[d * kF1 + tid];
        s.d1[tid] = s.h1[tid] > 0.f ? a : 0.f;
      }
      __syncthreads();
      CL_TRACE(9);
      fc_dx_partial_push<CS, kFlat>(f.wf1, s.d1, f1m0, f1m1, r, R, kRxG2, CL_BAR(kBarG2));
      zero_grad_plane();  // while the partials are in flight
      cl_mbar_wait(&s.sbar[kBarG2], phase);
      if (tid < kFlat) {
        float a = 0.f;
#pragma unroll
        for (int d = 0; d < CS; ++d) a += f.rx_g2[d * kFlat + tid];
        s.g2[tid] = s.p2[tid] > 0.f ? a : 0.f;
      }
      __syncthreads();
      CL_TRACE(10);
    } else {
      fc_dx_local<kF3, kF2>(params + kOffF3W, s.d3, s.h2, s.d2);
      __syncthreads();
      fc_dx_split_push<CS, kF2, kF1>(params + kOffF2W, s.d2, s.h1, d1k0, d1k1, sred, R, CL_OFF(d1), CL_BAR(kBarD1));
      cl_mbar_wait(&s.sbar[kBarD1], phase);
      CL_TRACE(9);
      __syncthreads();  // `red` is reused by the next split
      fc_dx_split_push<CS, kF1, kFlat>(params + kOffF1W, s.d1, s.p2, g2k0, g2k1, sred, R, CL_OFF(g2), CL_BAR(kBarG2));
      zero_grad_plane();  // while the g2 traffic is in flight
      cl_mbar_wait(&s.sbar[kBarG2], phase);
      __syncthreads();
      CL_TRACE(10);
    }
    // from here on nothing crosses CTAs any more

    if (r == 0) {  // FC record for the update kernel
      float* sv = saved + (size_t)img * kSavStride;
      if (tid < 400) sv[kSavP2 + tid] = s.p2[tid];
      if (tid < 120) { sv[kSavH1 + tid] = s.h1[tid]; sv[kSavD1 + tid] = s.d1[tid]; }
      if (tid < 84) { sv[kSavH2 + tid] = s.h2[tid]; sv[kSavD2 + tid] = s.d2[tid]; }
      if (tid < 10) sv[kSavD3 + tid] = s.d3[tid];
    }
    if (tid < 400) {  // scatter the pooled gradient to its argmax position (maxpool backward)
      const int o = tid / 25, pp = tid - o * 25, ph = pp / 5, pw = pp - ph * 5;
      const int id = s.idx2[tid];
      if constexpr (kTCD) {
        s.dpl[(((o >> 2) * kDPlanePix) + (2 * ph + (id >> 1) + 4) * kDPitch + (2 * pw + (id & 1) + 4)) * 4 + (o & 3)] = s.g2[tid];
        tc::fence_proxy_async();
      } else {
        s.dc2p[o * 324 + (2 * ph + (id >> 1) + 4) * 18 + (2 * pw + (id & 1) + 4)] = s.g2[tid];
      }
    }
    __syncthreads();  // the data-gradient items below read the plane just scattered

    // ================= conv2 data gradient (my pooled pixels) + conv2 weight gradient (my channels) ==========
    // The two are independent, so they share ONE pair of phases: the longer data-gradient items are handed out
    // first, the weight-gradient items fill the remaining threads, and only two block barriers are needed.
    {
      const int no = o1 - o0;
      const int n_dg = kTCD ? 0 : nq * 16;  // data gradient: (o, pixel) -> 6 channel sums (kTCD: on the tensor core instead)
      const int n_wg = no * 30 * 5;         // weight gradient: (pixel row, o, c, kh) -> 5 taps
      float* red_dg = sred;                            // [16*6][nq]
      float* red_wg = kTCD ? sred : sred + 96 * nq;    // [5][no*150]
      // kTCD: dX[c][ih][iw] = sum_{o,a,b} dCp[o][ih+a][iw+b] W[o][c][4-a][4-b] as ONE 128-row tile of the implicit GEMM
      // over the zero-padded gradient planes: GEMM row m = ih*18 + iw, my pixel rows start at row dg_m0 of the plane
      // (clamped so that the tile ends inside the plane; any cluster size's pixel range spans fewer than 128 rows).
      const int dg_m0 = min((q0 / 14) * kDPitch, 128);
      if constexpr (kTCD) {
        if (lane == 0 && warp < 2) {   // two issuing threads, 25 MMAs each, into two accumulators (added in the epilogue)
          if (packed != nullptr) cl_mbar_wait(&s.pbar, 0);   // the B operand has landed (long ago; stays complete)
          tc::tc_fence_after();
          constexpr uint32_t kIdesc = tc::idesc_tf32(128, 16);
          // K-chunks of MMA j = two of the four channel planes of tap j/2: planes 0,1 for even j, 2,3 for odd j
          const uint64_t a0 = tc::smem_desc_nosw(tc::smem_u32(s.dpl) + (uint32_t)dg_m0 * 16, kDPlanePix * 16, 128);
          const uint64_t b0 = tc::smem_desc_nosw(tc::smem_u32(s.sbd), 128, kBdChunks * 128);
          const uint32_t d = tmem + kTmemDg + (uint32_t)warp * 16;
          if (warp == 0) tc_issue_taps<0, 25, kDPitch, 2 * kDPlanePix>(d, a0, b0, kIdesc);
          else tc_issue_taps<25, 50, kDPitch, 2 * kDPlanePix>(d, a0, b0, kIdesc);
          tc::mma_commit(reinterpret_cast<uint64_t*>(&s.tbar[kTBarDg]));
          if (warp == 0) CL_TRACE(21);
        }
      }
      // (kTCD: warps 0 and 1 hold the two issuing threads; their other lanes would only start on weight-gradient
      // items once the issue loop is through, so the items go to warps 2..15)
      for (int item = kTCD ? tid - 64 : tid; item < n_dg + n_wg; item += kTCD ? kCLThreads - 64 : kCLThreads) {
        if (item < 0) break;
        if (item < n_dg) {
         if constexpr (!kTCD) {
          const int o = item / nq, ql = item - o * nq, q = q0 + ql;
          const int ih = q / 14, iw = q - ih * 14;
          float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int kh = 0; kh < 5; ++kh) {
            const int rr = ih + 4 - kh;
            if (rr < 4 || rr > 13) continue;
            const float* row = s.dc2p + o * 324 + rr * 18 + iw + 4;
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
#pragma unroll
          for (int c = 0; c < 6; ++c) red_dg[(o * 6 + c) * nq + ql] = acc[c];
         }
        } else {
          const int t = item - n_dg, nitem = no * 30;
          const int pg = t / nitem, rem = t - pg * nitem;
          const int ol = rem / 30, r2 = rem - ol * 30, c = r2 / 5, kh = r2 - c * 5, o = o0 + ol;
          float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            const int pp = pg * 5 + k;  // pooled row pg, column k
            const float g = s.g2[o * 25 + pp];
            const int id = s.idx2[o * 25 + pp];
            if constexpr (kTC) {   // p1 is NHWC4 here: consecutive pixels are 4 floats apart
              const float* row = s.p1n + (((c >> 2) * kP1PlanePix) + (2 * pg + (id >> 1) + kh) * 14 + 2 * k + (id & 1)) * 4 + (c & 3);
              a0 += g * row[0]; a1 += g * row[4]; a2 += g * row[8]; a3 += g * row[12]; a4 += g * row[16];
            } else {
              const float* row = s.p1 + c * 196 + (2 * pg + (id >> 1) + kh) * 14 + 2 * k + (id & 1);
              a0 += g * row[0]; a1 += g * row[1]; a2 += g * row[2]; a3 += g * row[3]; a4 += g * row[4];
            }
          }
          float* dst = red_wg + pg * (no * 150) + ol * 150 + c * 25 + kh * 5;
          dst[0] = a0; dst[1] = a1; dst[2] = a2; dst[3] = a3; dst[4] = a4;
        }
      }
      __syncthreads();
      CL_TRACE(11);
      if constexpr (kTCD) {
        cl_mbar_wait(&s.tbar[kTBarDg], phase);
        tc::tc_fence_after();
        if (warp < 4) {   // TMEM lane = GEMM row: dL/dp1 for my pixels, masked by ReLU (stays in this CTA)
          float v[8], u[8];
          tc::tmem_ld8(tmem + ((uint32_t)(warp * 32) << 16) + kTmemDg, v);
          tc::tmem_ld8(tmem + ((uint32_t)(warp * 32) << 16) + kTmemDg + 16, u);
#pragma unroll
          for (int c = 0; c < 6; ++c) v[c] += u[c];
          const int m = dg_m0 + warp * 32 + lane, ih = m / kDPitch, iw = m - ih * kDPitch, q = ih * 14 + iw;
          if (iw < 14 && q >= q0 && q < q1) {
#pragma unroll
            for (int c = 0; c < 6; ++c) {
              if constexpr (kTC) {
                const uint32_t pbits = __float_as_uint(s.p1n[(((c >> 2) * kP1PlanePix) + q) * 4 + (c & 3)]);
                s.g1[c * 196 + q] = (pbits & ~3u) != 0u ? v[c] : 0.f;   // p1 > 0, ignoring the argmax bits
              } else {
                s.g1[c * 196 + q] = s.p1[c * 196 + q] > 0.f ? v[c] : 0.f;
              }
            }
          }
        }
        tc::tc_fence_before();
      }
      float* dst = part2 + (size_t)img * (kConvPrefix - kOffW2);
      const int n_g1 = kTCD ? 0 : nq * 6, n_w2 = no * 150;
      for (int t = tid; t < n_g1 + n_w2 + no; t += kCLThreads) {
        if (t < n_g1) {                    // dL/dp1 for my pixels, masked by ReLU (stays in this CTA)
         if constexpr (!kTCD) {
          const int c = t / nq, ql = t - c * nq;
          float v = 0.f;
#pragma unroll
          for (int o = 0; o < 16; ++o) v += red_dg[(o * 6 + c) * nq + ql];
          const int gi = c * 196 + q0 + ql;
          s.g1[gi] = s.p1[gi] > 0.f ? v : 0.f;
         }
        } else if (t < n_g1 + n_w2) {      // conv2.weight gradient of this image, my channels
          const int u = t - n_g1;
          float a = 0.f;
#pragma unroll
          for (int pg = 0; pg < 5; ++pg) a += red_wg[pg * n_w2 + u];
          dst[o0 * 150 + u] = a;
        } else {                           // conv2.bias gradient
          const int o = o0 + (t - n_g1 - n_w2);
          float a = 0.f;
          for (int pp = 0; pp < 25; ++pp) a += s.g2[o * 25 + pp];
          dst[(kOffB2 - kOffW2) + o] = a;
        }
      }
    }
    __syncthreads();
    if constexpr (kTCD && !kTC) {
      // tensor memory is not touched again after the last image's data gradient: give it back here instead of behind
      // the kernel's last stage (the tcgen05.ld of every warp precede the block barrier above)
      if (img + n_clusters >= B && warp == 1) tc::tmem_free<kTmemCols>(tmem);
    }

    CL_TRACE(12);
    // ================= conv1 weight gradient over my pooled pixels (partial) =================
    // thread = (pixel group of 5, o, c, kh) with the five kw taps in registers
    {
      if (tid < 450) {
        const int pg = tid / 90, rem = tid - pg * 90;
        const int o = rem / 15, r2 = rem - o * 15, c = r2 / 5, kh = r2 - c * 5;
        const float* gp = s.g1 + o * 196;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
        if constexpr (kTC) {   // argmax from the low bits of p1, image from the NHWC4 plane (pixels 4 floats apart)
          const float* pv = s.p1n + ((o >> 2) * kP1PlanePix) * 4 + (o & 3);
          const float* xb = s.xpl + (kh * kXPitch) * 4 + c;
#pragma unroll 4
          for (int q = q0 + pg; q < q1; q += 5) {
            const float g = gp[q];
            const int id = (int)(__float_as_uint(pv[q * 4]) & 3u);
            const float* row = xb + (s.poff[q] + (id >> 1) * kXPitch + (id & 1)) * 4;
            a0 += g * row[0]; a1 += g * row[4]; a2 += g * row[8]; a3 += g * row[12]; a4 += g * row[16];
          }
        } else {
          const int* ip = s.idx1 + o * 196;
          const float* xb = s.x + c * kXCh + kh * kXRow;
#pragma unroll 4
          for (int q = q0 + pg; q < q1; q += 5) {
            const float g = gp[q];
            const int id = ip[q];
            const float* row = xb + s.poff[q] + (id >> 1) * kXRow + (id & 1);
            a0 += g * row[0]; a1 += g * row[1]; a2 += g * row[2]; a3 += g * row[3]; a4 += g * row[4];
          }
        }
        float* dst = sred + pg * 450 + o * 75 + c * 25 + kh * 5;
        dst[0] = a0; dst[1] = a1; dst[2] = a2; dst[3] = a3; dst[4] = a4;
      } else if (warp == 15 && lane < 30) {
        // bias partial: (channel, pixel group of 5) per lane, same two-step shape as the weights. (Six warp-wide
        // reductions one after the other in this warp used to be what the barrier below waited for.)
        const int o = lane / 5, pg = lane - o * 5;
        float a = 0.f;
        for (int q = q0 + pg; q < q1; q += 5) a += s.g1[o * 196 + q];
        sred[5 * 450 + pg * 6 + o] = a;
      }
      __syncthreads();
      if (tid < kW1Row) {   // 450 weight taps + 6 biases: add the five pixel groups, send the partial to CTA 0
        float a = 0.f;
        if (tid < 450) {
#pragma unroll
          for (int pg = 0; pg < 5; ++pg) a += sred[pg * 450 + tid];
        } else {
#pragma unroll
          for (int pg = 0; pg < 5; ++pg) a += sred[5 * 450 + pg * 6 + (tid - 450)];
        }
        cl_st_async(R.base[0] + kRxW1 + 4u * (uint32_t)(r * kW1Row + tid), __float_as_uint(a),
                    R.base[0] + CL_BAR(kBarW1));
      }
      if (r == 0) {
        // CTA 0 adds the CS partials in CTA order and writes ONE conv1 gradient row per image, so the update
        // kernel's first slice sums B rows instead of B*CS (it was that kernel's critical path)
        cl_mbar_wait(&s.sbar[kBarW1], phase);
        if (tid < kW1Row) {
          float a = 0.f;
#pragma unroll
          for (int d = 0; d < CS; ++d) a += rx_w1[d * kW1Row + tid];
          part1[(size_t)img * kOffW2 + (tid < 450 ? kOffW1 + tid : kOffB1 + (tid - 450))] = a;
        }
      }
    }
    CL_TRACE(13);
    first_img = false;
    if (trace != nullptr && tid == 0) trace[blockIdx.x * 32 + 15] = (long long)globaltimer_ns();
    phase ^= 1u;
    if (img + n_clusters < B) {
      // several images per cluster: nobody may push the next image's activations (or reload the
      // image buffer) while a slower CTA of the cluster still reads this image's
      __syncthreads();
      cl_cluster_arrive();
      cl_cluster_wait();
      load_image(img + n_clusters);
    }
  }
  if constexpr (kTC) {
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_free<kTmemCols>(tmem);
  }
}

template <int CS, int kMode>
static int launch_cluster_t(const float* params, const float* x, const long long* labels, float* saved, float* part1,
                            float* part2, float* img_loss, int* img_correct, int B, int max_clusters,
                            long long* trace, const unsigned int* ready_ctr, unsigned int ready_expect,
                            const float* packed, const unsigned int* in_flag, unsigned int in_expect, cudaStream_t stream) {
  static PerDeviceOnce attr_set;
  auto kern = fwd_bwd_cluster_kernel<CS, kMode>;
  const size_t smem = cluster_smem_bytes<CS, kMode>();
  if (int rc = ensure_max_dynamic_smem(attr_set, kern, smem)) return rc;
  const int n_clusters = B < max_clusters ? B : max_clusters;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(n_clusters * CS);
  cfg.blockDim = dim3(kCLThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  // programmatic dependent launch: this kernel may be scheduled while the previous step's update kernel is
  // still running; it loads its inputs, then waits (griddepcontrol.wait) before it reads the weights
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  const float inv_batch = 1.f / (float)B;
  DNN_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, params, x, labels, saved, part1, part2, img_loss, img_correct, B,
                                    inv_batch, trace, CS >= 4 ? ready_ctr : nullptr, ready_expect,
                                    kMode != 0 ? packed : nullptr, in_flag, in_expect));
  return kOk;
}

// Cluster size from the batch: use as many SMs as the batch allows in ONE wave.
// (clusters of 8 pack 2 per GPC -> 16 resident; of 4 -> ~33; of 2 -> 74.)
int pick_cluster_size(int B, int num_sms) {
  if (B * 8 <= num_sms - 20) return 8;
  if (B * 4 <= num_sms - 16) return 4;
  if (B * 2 <= num_sms) return 2;
  return 1;
}

// how many clusters of CS CTAs of this kernel the device can hold at once
template <int CS, int kMode>
static int max_active_clusters_t() {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(CS * 64);
  cfg.blockDim = dim3(kCLThreads);
  cfg.dynamicSmemBytes = cluster_smem_bytes<CS, kMode>();
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CS;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  cudaFuncSetAttribute(fwd_bwd_cluster_kernel<CS, kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       (int)cluster_smem_bytes<CS, kMode>());
  if (cudaOccupancyMaxActiveClusters(&n, fwd_bwd_cluster_kernel<CS, kMode>, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
template <int kMode>
static int max_active_clusters_p(int cs) {
  switch (cs) {
    case 2: return max_active_clusters_t<2, kMode>();
    case 3: return max_active_clusters_t<3, kMode>();
    case 4: return max_active_clusters_t<4, kMode>();
    case 5: return max_active_clusters_t<5, kMode>();
    case 6: return max_active_clusters_t<6, kMode>();
    case 7: return max_active_clusters_t<7, kMode>();
    case 8: return max_active_clusters_t<8, kMode>();
    default: return 0;
  }
}
int max_active_clusters(int cs, int mode) {
  return mode == 1 ? max_active_clusters_p<1>(cs) : (mode == 2 ? max_active_clusters_p<2>(cs) : max_active_clusters_p<0>(cs));
}

template <int kMode>
static int launch_cluster_p(int cs, const float* params, const float* x, const long long* labels, float* saved,
                            float* part1, float* part2, float* img_loss, int* img_correct, int B, int max_clusters,
                            long long* trace, const unsigned int* ready_ctr, unsigned int ready_expect,
                            const float* packed, const unsigned int* in_flag, unsigned int in_expect, cudaStream_t stream) {
#define DNN_CL_CASE(N)                                                                                              \
  case N:                                                                                                           \
    return launch_cluster_t<N, kMode>(params, x, labels, saved, part1, part2, img_loss, img_correct, B, max_clusters, \
                                    trace, ready_ctr, ready_expect, packed, in_flag, in_expect, stream);
  switch (cs) {
    DNN_CL_CASE(8) DNN_CL_CASE(7) DNN_CL_CASE(6) DNN_CL_CASE(5) DNN_CL_CASE(4) DNN_CL_CASE(3) DNN_CL_CASE(2)
    default: set_last_error("cluster size %d not supported", cs); return kErrArg;
  }
#undef DNN_CL_CASE
}

// mode 0: everything fp32 FFMA (the parity path, bit-comparable in structure with the one-CTA kernel); 1: conv1/conv2
// forward and the conv2 data gradient on tcgen05 (TF32 operands, fp32 accumulate); 2: only the conv2 data gradient
int launch_fwd_bwd_cluster(int cs, int mode, const float* params, const float* x, const long long* labels,
                           float* saved, float* part1, float* part2, float* img_loss, int* img_correct, int B,
                           int max_clusters, long long* trace, const unsigned int* ready_ctr, unsigned int ready_expect,
                           const float* packed, const unsigned int* in_flag, unsigned int in_expect,
                           cudaStream_t stream) {
#define DNN_CL_MODE(M)                                                                                             \
  launch_cluster_p<M>(cs, params, x, labels, saved, part1, part2, img_loss, img_correct, B, max_clusters, trace, \
                      ready_ctr, ready_expect, M != 0 ? packed : nullptr, in_flag, in_expect, stream)
  return mode == 1 ? DNN_CL_MODE(1) : (mode == 2 ? DNN_CL_MODE(2) : DNN_CL_MODE(0));
#undef DNN_CL_MODE
}

}  // namespace dnnb200
