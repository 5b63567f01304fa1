// Backward of y = act(x . W^T + b) on the 5th-generation tensor cores (kernels K13/K14 of SURVEY §2.7;
// reference: autograd of models/model.py:24-26). Three launches:
//
//   prep   dy' = dy * (y > 0) (only when the layer has a ReLU) and db[m] = sum_n dy'[n][m]
//   dX     D[k][n] = sum_m W[m][k]   dy'[n][m]      A = W   MN-major (k contiguous), B = dy' K-major
//   dW     D[m][k] = sum_n dy'[n][m] x[n][k]        A = dy' MN-major,               B = x   MN-major
//
// No operand is ever transposed or padded in memory: every tile is a TMA box of [rows][32 floats] with the
// 128-byte swizzle, taken from the row-major matrices as they are, and the MN-major form of the UMMA
// shared-memory descriptor reads a box "sideways" (the contiguous index is M/N, the box rows are the
// reduction index). For 4-byte operands the tensor core accepts a single MN-major layout, the 128-byte
// swizzle with 32-byte atomicity; TMA produces it with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B (tc_common.cuh).
// Rows/columns outside the matrices are zero-filled by the TMA unit; boxes that lie completely outside are
// not loaded at all (they only feed accumulator rows/columns that are never stored).
//
// dW is split over the batch across CTAs (grid.y); partial tiles are added with red.global.add.f32 into a
// zeroed dw, so its summation order is not fixed (TF32 inputs already bound the accuracy; the fp32 FFMA
// kernels of layers.cu are the deterministic path).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "tc_tmap.h"

namespace dnnb200 {

using namespace tc;

constexpr int kBoxFloats = 32 * 32;          // one TMA box: 32 rows x 32 floats = 4 KB = 8 swizzle groups of 4 rows
constexpr int kBoxBytes = kBoxFloats * 4;
constexpr int kBwdStages = 4;

// ---------------------------------------------------------------------------------------------
// prep: ReLU mask + bias gradient
// ---------------------------------------------------------------------------------------------
constexpr int kPrepRows = 32;                 // rows per CTA: 2 row lanes x 16 rows, all loads independent

__global__ void __launch_bounds__(256)
linear_bwd_prep_kernel(const float* __restrict__ y, const float* __restrict__ dy, float* __restrict__ dyp,
                       float* __restrict__ db, int N, int M, int relu) {
  __shared__ float part[128];
  const int m = threadIdx.x & 127, ty = threadIdx.x >> 7;
  const int r0 = blockIdx.x * kPrepRows + ty;
  float acc = 0.f;
  if (m < M) {
    float d[kPrepRows / 2], v[kPrepRows / 2];
#pragma unroll
    for (int i = 0; i < kPrepRows / 2; ++i) {
      const int r = r0 + 2 * i;
      d[i] = r < N ? dy[(size_t)r * M + m] : 0.f;
      v[i] = (relu && r < N) ? y[(size_t)r * M + m] : 1.f;
    }
#pragma unroll
    for (int i = 0; i < kPrepRows / 2; ++i) {
      const int r = r0 + 2 * i;
      if (!(v[i] > 0.f)) d[i] = 0.f;
      if (relu && r < N) dyp[(size_t)r * M + m] = d[i];
      acc += d[i];
    }
  }
  if (ty == 1) part[m] = acc;
  __syncthreads();
  if (ty == 0 && m < M) atomicAdd(db + m, acc + part[m]);
}

// ---------------------------------------------------------------------------------------------
// dX[n][k] = sum_m dy'[n][m] W[m][k]
// ---------------------------------------------------------------------------------------------
template <int BN>
struct DxSmem {
  alignas(1024) float a[4][4 * kBoxFloats];   // per 32-row block of m: 4 boxes of 32 k each (W, MN-major)
  alignas(1024) float b[4][BN * 32];          // per block of m: dy' tile [BN rows][32 floats of m], K-major
  alignas(8) uint64_t full[4];
  alignas(8) uint64_t done;
  uint32_t tmem_slot;
};

template <int BN>
__global__ void __launch_bounds__(128)
tc_linear_dx_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_dy,
                    float* __restrict__ dx, int N, int M, int K) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem_al = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  DxSmem<BN>& s = *reinterpret_cast<DxSmem<BN>*>(smem_al);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * BN, k0 = blockIdx.y * 128;
  const int nmb = (M + 31) / 32;                          // reduction blocks (<= 4 since M <= 128)
  const int nabox = min(4, (K - k0 + 31) / 32);           // boxes of this k tile that touch the matrix
  constexpr int kTmemCols = BN < 32 ? 32 : BN;

  if (tid == 0) {
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_dy);
    for (int i = 0; i < 4; ++i) mbar_init(&s.full[i], 1);
    mbar_init(&s.done, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(&s.tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_slot;

  if (tid == 0) {
    for (int mb = 0; mb < nmb; ++mb) {
      mbar_expect_tx(&s.full[mb], nabox * kBoxBytes + BN * 128);
      for (int i = 0; i < nabox; ++i) tma_load_2d(s.a[mb] + i * kBoxFloats, &tmap_w, k0 + 32 * i, mb * 32, &s.full[mb]);
      tma_load_2d(s.b[mb], &tmap_dy, mb * 32, n0, &s.full[mb]);
    }
  } else if (tid == 32) {
    constexpr uint32_t idesc = idesc_tf32(128, BN, /*a_mn=*/true, /*b_mn=*/false);
    bool acc = false;
    for (int mb = 0; mb < nmb; ++mb) {
      mbar_wait(&s.full[mb], 0);
      tc_fence_after();
      const uint32_t a0 = smem_u32(s.a[mb]), b0 = smem_u32(s.b[mb]);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        if (mb * 32 + kk * 8 < M) {
          // 8 reduction rows (m) = 1 KB of every box = two 4-row swizzle groups; boxes (32 k each) are LBO apart
          const uint64_t ad = smem_desc_mn_sw128_32b(a0 + kk * 1024, kBoxBytes, 512);
          const uint64_t bd = smem_desc_sw128(b0 + kk * 32, 1024);
          mma_tf32(tmem, ad, bd, idesc, acc);
          acc = true;
        }
      }
    }
    mma_commit(&s.done);
  }
  __syncwarp();
  mbar_wait(&s.done, 0);
  tc_fence_after();

  // TMEM lane = input feature k, column = batch row: lanes store consecutive k (coalesced)
  const int k = k0 + warp * 32 + lane;
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 16) {
    float v[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
    if (k < K) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int n = n0 + c0 + j;
        if (n < N) dx[(size_t)n * K + k] = v[j];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<kTmemCols>(tmem);
}

// ---------------------------------------------------------------------------------------------
// dW[m][k] = sum_n dy'[n][m] x[n][k]   (k tile = blockIdx.x, batch split = blockIdx.y)
// ---------------------------------------------------------------------------------------------
struct DwSmem {
  alignas(1024) float a[kBwdStages][4 * kBoxFloats];   // dy' rows n0..n0+31: 4 boxes of 32 m each
  alignas(1024) float b[kBwdStages][4 * kBoxFloats];   // x   rows n0..n0+31: 4 boxes of 32 k each
  alignas(8) uint64_t full[kBwdStages];
  alignas(8) uint64_t empty[kBwdStages];
  alignas(8) uint64_t done;
  uint32_t tmem_slot;
};

__global__ void __launch_bounds__(128)
tc_linear_dw_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                    float* __restrict__ dw, int N, int M, int K, int rows_per_split, int use_atomics) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem_al = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  DwSmem& s = *reinterpret_cast<DwSmem*>(smem_al);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int k0 = blockIdx.x * 128;
  const int r0 = blockIdx.y * rows_per_split;            // multiple of 32
  const int r1 = min(N, r0 + rows_per_split);
  const int nb = (r1 - r0 + 31) / 32;                    // 32-row reduction blocks of this split (>= 1)
  const int nabox = (M + 31) / 32;
  const int nbbox = min(4, (K - k0 + 31) / 32);

  if (tid == 0) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
    for (int i = 0; i < kBwdStages; ++i) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], 1); }
    mbar_init(&s.done, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<128>(&s.tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_slot;

  if (tid == 0) {
    for (int i = 0; i < nb; ++i) {
      const int st = i % kBwdStages;
      const uint32_t ph = (uint32_t)(i / kBwdStages) & 1u;
      mbar_wait(&s.empty[st], ph ^ 1u);
      mbar_expect_tx(&s.full[st], (nabox + nbbox) * kBoxBytes);
      const int r = r0 + i * 32;
      for (int j = 0; j < nabox; ++j) tma_load_2d(s.a[st] + j * kBoxFloats, &tmap_dy, 32 * j, r, &s.full[st]);
      for (int j = 0; j < nbbox; ++j) tma_load_2d(s.b[st] + j * kBoxFloats, &tmap_x, k0 + 32 * j, r, &s.full[st]);
    }
  } else if (tid == 32) {
    constexpr uint32_t idesc = idesc_tf32(128, 128, /*a_mn=*/true, /*b_mn=*/true);
    for (int i = 0; i < nb; ++i) {
      const int st = i % kBwdStages;
      const uint32_t ph = (uint32_t)(i / kBwdStages) & 1u;
      mbar_wait(&s.full[st], ph);
      tc_fence_after();
      const uint32_t a0 = smem_u32(s.a[st]), b0 = smem_u32(s.b[st]);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t ad = smem_desc_mn_sw128_32b(a0 + kk * 1024, kBoxBytes, 512);
        const uint64_t bd = smem_desc_mn_sw128_32b(b0 + kk * 1024, kBoxBytes, 512);
        mma_tf32(tmem, ad, bd, idesc, (i | kk) != 0);
      }
      mma_commit(&s.empty[st]);
    }
    mma_commit(&s.done);
  }
  __syncwarp();
  mbar_wait(&s.done, 0);
  tc_fence_after();

  // TMEM lane = output feature m (row of dW), column = input feature k
  const int m = warp * 32 + lane;
#pragma unroll 1
  for (int c0 = 0; c0 < 128; c0 += 16) {
    if (k0 + c0 >= K) break;                              // warp-uniform
    float v[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
    if (m < M) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int k = k0 + c0 + j;
        if (k < K) {
          if (use_atomics) atomicAdd(dw + (size_t)m * K + k, v[j]);
          else dw[(size_t)m * K + k] = v[j];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<128>(tmem);
}

// ---------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------
template <int BN>
static int launch_dx_t(const CUtensorMap& tw, const float* dyp, float* dx, int N, int M, int K, cudaStream_t s) {
  static PerDeviceOnce attr_set;
  auto kern = tc_linear_dx_kernel<BN>;
  const size_t smem = sizeof(DxSmem<BN>) + 1024;
  if (int rc0 = ensure_max_dynamic_smem(attr_set, kern, smem)) return rc0;
  CUtensorMap tdy;
  const int rc = make_tmap_2d(&tdy, dyp, N, M, BN);
  if (rc) return rc;
  kern<<<dim3((N + BN - 1) / BN, (K + 127) / 128), 128, smem, s>>>(tw, tdy, dx, N, M, K);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_tc_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dyp, float* dx,
                         float* dw, float* db, int N, int M, int K, int relu, cudaStream_t s) {
  DNN_ARG_CHECK(M >= 4 && M <= 128 && M % 4 == 0, "tc_linear_bwd: M must be a multiple of 4 in [4, 128]");
  DNN_ARG_CHECK(K >= 4 && K % 4 == 0, "tc_linear_bwd: K must be a positive multiple of 4 (16-byte rows for TMA)");
  DNN_ARG_CHECK(!relu || dyp != nullptr, "tc_linear_bwd: a ReLU layer needs the dy' scratch buffer [N, M]");
  const float* dmat = relu ? dyp : dy;
  DNN_ARG_CHECK(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(dmat)) &
                 15) == 0,
                "tc_linear_bwd: x, w and dy must be 16-byte aligned");

  DNN_CUDA_CHECK(cudaMemsetAsync(db, 0, (size_t)M * sizeof(float), s));
  linear_bwd_prep_kernel<<<(N + kPrepRows - 1) / kPrepRows, 256, 0, s>>>(y, dy, dyp, db, N, M, relu);
  DNN_CUDA_CHECK(cudaGetLastError());

  if (dx) {
    CUtensorMap tw;
    int rc = make_tmap_2d(&tw, w, M, K, 32, /*mn_major=*/true);
    if (rc) return rc;
    rc = N <= 32 ? launch_dx_t<32>(tw, dmat, dx, N, M, K, s) : launch_dx_t<128>(tw, dmat, dx, N, M, K, s);
    if (rc) return rc;
  }

  {
    static PerDeviceOnce attr_set;
    const size_t smem = sizeof(DwSmem) + 1024;
    if (int rc0 = ensure_max_dynamic_smem(attr_set, tc_linear_dw_kernel, smem)) return rc0;
    CUtensorMap tdy, tx;
    int rc = make_tmap_2d(&tdy, dmat, N, M, 32, /*mn_major=*/true);
    if (rc) return rc;
    rc = make_tmap_2d(&tx, x, N, K, 32, /*mn_major=*/true);
    if (rc) return rc;
    const int ktiles = (K + 127) / 128;
    const int blocks = (N + 31) / 32;
    // at least 8 reduction blocks (256 rows) per CTA; swept 4..64 (bench/sweep_dw.sh): 4-16 are within 3 % of each
    // other, beyond that the streaming phase loses its parallelism
    int splits = blocks / 8;
    if (splits > 148 / ktiles) splits = 148 / ktiles;
    if (splits < 1) splits = 1;
    const int rows_per_split = ((blocks + splits - 1) / splits) * 32;
    splits = (N + rows_per_split - 1) / rows_per_split;    // drop empty trailing splits
    if (splits > 1) DNN_CUDA_CHECK(cudaMemsetAsync(dw, 0, (size_t)M * K * sizeof(float), s));
    tc_linear_dw_kernel<<<dim3(ktiles, splits), 128, smem, s>>>(tdy, tx, dw, N, M, K, rows_per_split, splits > 1);
    DNN_CUDA_CHECK(cudaGetLastError());
  }
  return kOk;
}

}  // namespace dnnb200
