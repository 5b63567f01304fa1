// y[N,M] = act(x[N,K] . W[M,K]^T + b[M]) on the 5th-generation tensor cores, operands fed by TMA
// (reference: models/model.py:24-26, kernels K8-K10 of SURVEY §2.7).
//
// Roles swapped so that the FEATURE dimension is UMMA-M: D[m][n] = sum_k W[m][k] x[n][k]. Both
// operands are K-major exactly as they sit in global memory (row-major W and x), so they are loaded
// with 2-D tensor maps (cuTensorMapEncodeTiled, 128-byte swizzle) straight into the canonical
// SW128 K-major layout that tcgen05.mma reads. Rows beyond M (120/84/10 < 128) and beyond N, and the
// K tail (400 = 12.5 x 32), are zero-filled by the TMA unit -- no padding copies anywhere.
//
//   CTA = one batch tile of BN rows. K loop in blocks of 32 floats (one 128-byte swizzle row),
//   4-stage ring: thread 0 = TMA producer, thread 32 = MMA issuer (4 x tcgen05.mma kind::tf32 per
//   block, K = 8 each, descriptor start address advanced by 32 bytes inside the swizzle atom),
//   tcgen05.commit releases the stage; after the last block the 4 warps read their 32 TMEM lanes
//   (= 32 features) 16 batch columns at a time, add bias, apply ReLU, and store coalesced along M.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "tc_tmap.h"

namespace dnnb200 {

using namespace tc;

constexpr int kLinBK = 32;       // floats per K block = 128 bytes = one swizzle row
constexpr int kLinStages = 4;
constexpr int kLinM = 128;       // UMMA M (features, zero-filled beyond the real M)

template <int BN>
struct LinSmem {
  alignas(1024) float a[kLinStages][kLinM * kLinBK];   // W tile   [128 rows][32 floats], SW128
  alignas(1024) float b[kLinStages][BN * kLinBK];      // x tile   [BN rows][32 floats], SW128
  alignas(8) uint64_t full[kLinStages];
  alignas(8) uint64_t empty[kLinStages];
  alignas(8) uint64_t done;
  uint32_t tmem_slot;
};

template <int BN>
__global__ void __launch_bounds__(128)
tc_linear_fwd_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                     const float* __restrict__ bias, float* __restrict__ y, int N, int M, int K, int relu) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // SW128 tiles need 1024-byte alignment: align by hand (the launch reserves 1 KB of slack)
  unsigned char* smem_al = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  LinSmem<BN>& s = *reinterpret_cast<LinSmem<BN>*>(smem_al);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * BN;
  const int nkb = (K + kLinBK - 1) / kLinBK;
  constexpr int kTmemCols = BN < 32 ? 32 : BN;

  if (tid == 0) {
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_x);
    for (int i = 0; i < kLinStages; ++i) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], 1); }
    mbar_init(&s.done, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(&s.tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_slot;

  if (tid == 0) {
    // ---------------- TMA producer ----------------
    for (int kb = 0; kb < nkb; ++kb) {
      const int st = kb % kLinStages;
      const uint32_t ph = (uint32_t)(kb / kLinStages) & 1u;
      mbar_wait(&s.empty[st], ph ^ 1u);   // fresh barrier: parity 1 passes immediately
      mbar_expect_tx(&s.full[st], (kLinM + BN) * kLinBK * 4);
      tma_load_2d(s.a[st], &tmap_w, kb * kLinBK, 0, &s.full[st]);
      tma_load_2d(s.b[st], &tmap_x, kb * kLinBK, n0, &s.full[st]);
    }
  } else if (tid == 32) {
    // ---------------- MMA issuer ----------------
    constexpr uint32_t idesc = idesc_tf32(kLinM, BN);
    for (int kb = 0; kb < nkb; ++kb) {
      const int st = kb % kLinStages;
      const uint32_t ph = (uint32_t)(kb / kLinStages) & 1u;
      mbar_wait(&s.full[st], ph);
      tc_fence_after();
      const uint32_t a0 = smem_u32(s.a[st]), b0 = smem_u32(s.b[st]);
#pragma unroll
      for (int k = 0; k < kLinBK / 8; ++k) {
        const uint64_t ad = smem_desc_sw128(a0 + k * 32, 1024);
        const uint64_t bd = smem_desc_sw128(b0 + k * 32, 1024);
        mma_tf32(tmem, ad, bd, idesc, (kb | k) != 0);
      }
      mma_commit(&s.empty[st]);           // the stage may be refilled once these MMAs have read it
    }
    mma_commit(&s.done);
  }
  __syncwarp();
  mbar_wait(&s.done, 0);
  tc_fence_after();

  // ---------------- epilogue: TMEM lane = feature m, column = batch row ----------------
  const int m = warp * 32 + lane;
  const float bv = m < M ? __ldg(bias + m) : 0.f;
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 16) {
    float v[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
    if (m < M) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int n = n0 + c0 + j;
        if (n < N) {
          const float t = v[j] + bv;
          y[(size_t)n * M + m] = relu ? fmaxf(t, 0.f) : t;   // lanes = consecutive m: coalesced
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<kTmemCols>(tmem);
}

template <int BN>
static int launch_tc_linear_t(const float* x, const float* w, const float* b, float* y, int N, int M, int K, int relu,
                              cudaStream_t s) {
  static PerDeviceOnce attr_set;
  auto kern = tc_linear_fwd_kernel<BN>;
  const size_t smem = sizeof(LinSmem<BN>) + 1024;
  if (int rc0 = ensure_max_dynamic_smem(attr_set, kern, smem)) return rc0;
  CUtensorMap tw, tx;
  int rc = make_tmap_2d(&tw, w, M, K, kLinM);
  if (rc) return rc;
  rc = make_tmap_2d(&tx, x, N, K, BN);
  if (rc) return rc;
  kern<<<(N + BN - 1) / BN, 128, smem, s>>>(tw, tx, b, y, N, M, K, relu);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_tc_linear_fwd(const float* x, const float* w, const float* b, float* y, int N, int M, int K, int relu,
                         cudaStream_t s) {
  DNN_ARG_CHECK(M >= 1 && M <= kLinM, "tc_linear_fwd: M must be in [1, %d]", kLinM);
  DNN_ARG_CHECK(K >= 4 && K % 4 == 0, "tc_linear_fwd: K must be a positive multiple of 4 (16-byte rows for TMA)");
  DNN_ARG_CHECK((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0,
                "tc_linear_fwd: x and w must be 16-byte aligned");
  if (N <= 32) return launch_tc_linear_t<32>(x, w, b, y, N, M, K, relu, s);
  return launch_tc_linear_t<128>(x, w, b, y, N, M, K, relu, s);
}

}  // namespace dnnb200
