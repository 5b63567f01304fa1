// Descriptor laboratory (development aid, also a unit test): the host supplies the raw shared-memory images of
// both operands and the descriptors; the kernel issues tcgen05.mma with them and returns D. With a one-hot /
// ramp image on one side and an identity on the other, D shows which shared-memory word the tensor core
// reads for every (row, k) of an operand -- the way the MN-major encodings of tc_linear_bwd.cu were pinned
// down without documentation at hand.
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace dnnb200 {

using namespace tc;

// a_img / b_img are copied to 1024-byte aligned shared memory; the descriptors carry start address 0 and get
// the base added here. nmma MMAs are issued, the start addresses advancing by a_step / b_step bytes each.
__global__ void __launch_bounds__(128)
tc_lab_kernel(const float* __restrict__ a_img, int a_floats, const float* __restrict__ b_img, int b_floats,
              unsigned long long adesc, unsigned long long bdesc, unsigned int idesc, int nmma, int a_step,
              int b_step, float* __restrict__ D, int ncols) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float* sA = reinterpret_cast<float*>(base);
  float* sB = sA + ((a_floats + 255) / 256) * 256;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < a_floats; i += 128) sA[i] = a_img[i];
  for (int i = tid; i < b_floats; i += 128) sB[i] = b_img[i];
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc<256>(&tmem_slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    for (int i = 0; i < nmma; ++i) {
      const uint64_t ad = adesc + (uint64_t)((smem_u32(sA) + (uint32_t)(i * a_step)) >> 4);
      const uint64_t bd = bdesc + (uint64_t)((smem_u32(sB) + (uint32_t)(i * b_step)) >> 4);
      mma_tf32(tmem, ad, bd, idesc, i > 0);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < ncols; c0 += 16) {
    float v[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (c0 + j < ncols) D[(size_t)(warp * 32 + lane) * ncols + c0 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_free<256>(tmem);
}

// Timing twin of the kernel above: how long do `nmma` tiny MMAs take when they accumulate into ONE tile of tensor
// memory (a dependent chain) versus rotating over n_acc tiles, and how does that depend on the operand layout
// (descriptors, steps between the operands of consecutive MMAs)? out[0] = cycles thread 0 spent issuing,
// out[1] = cycles from the first issue until the commit barrier fired. Operand contents do not matter (zeros).
__global__ void __launch_bounds__(128)
tc_lab_time_kernel(int a_bytes, int b_bytes, unsigned long long adesc, unsigned long long bdesc, unsigned int idesc,
                   int nmma, int a_step, int b_step, int n_acc, int acc_cols, long long* __restrict__ out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (a_bytes + b_bytes) / 4; i += 128) reinterpret_cast<float*>(base)[i] = 0.f;
  if (tid == 0) { mbar_init(&bar, n_acc < 0 ? -n_acc : 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc<256>(&tmem_slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t sa = smem_u32(base), sb = sa + (uint32_t)a_bytes;
  long long t0 = 0, t1 = 0;
  // n_acc < 0: |n_acc| ISSUING THREADS (lane 0 of warps 0..|n_acc|-1), each with its own accumulator tile and every
  // |n_acc|-th MMA -- is the cost of an MMA a property of the issuing thread or of the tensor pipe?
  const int n_issuers = n_acc < 0 ? -n_acc : 1;
  if ((tid & 31) == 0 && warp < n_issuers) {
    const int first = n_acc < 0 ? warp : 0, stride = n_issuers;
    uint64_t ad = adesc + (uint64_t)((sa + (uint32_t)(first * a_step)) >> 4);
    uint64_t bd = bdesc + (uint64_t)((sb + (uint32_t)(first * b_step)) >> 4);
    const uint64_t da = (uint64_t)((a_step * stride) >> 4), db = (uint64_t)((b_step * stride) >> 4);
    const uint32_t mask = n_acc < 0 ? 0u : (uint32_t)(n_acc - 1);
    const uint32_t my_acc = tmem + (n_acc < 0 ? (uint32_t)(warp * acc_cols) : 0u);
    const int lead = n_acc < 0 ? 1 : n_acc;
    t0 = clock64();
    int k = 0;
#pragma unroll 4
    for (int i = first; i < nmma; i += stride, ++k) {
      mma_tf32(my_acc + ((uint32_t)k & mask) * (uint32_t)acc_cols, ad, bd, idesc, k >= lead);
      ad += da;
      bd += db;
    }
    mma_commit(&bar);
    t1 = clock64();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  if (tid == 0) { out[0] = t1 - t0; out[1] = clock64() - t0; }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_free<256>(tmem);
}

int launch_tc_lab_time(int a_bytes, int b_bytes, unsigned long long adesc, unsigned long long bdesc, unsigned int idesc,
                       int nmma, int a_step, int b_step, int n_acc, int acc_cols, long long* d_out2, cudaStream_t s) {
  const int n_tiles = n_acc < 0 ? -n_acc : n_acc;
  DNN_ARG_CHECK(a_bytes > 0 && b_bytes > 0 && a_bytes % 1024 == 0 && nmma >= 1 && n_tiles >= 1 && n_tiles <= 4 &&
                    (n_tiles & (n_tiles - 1)) == 0 && n_tiles * acc_cols <= 256,
                "tc_lab_time: bad arguments");
  const size_t smem = (size_t)a_bytes + b_bytes + 1024;
  DNN_ARG_CHECK(smem <= 200 * 1024, "tc_lab_time: operand images exceed 200 KB of shared memory");
  DNN_CUDA_CHECK(cudaFuncSetAttribute(tc_lab_time_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tc_lab_time_kernel<<<1, 128, smem, s>>>(a_bytes, b_bytes, adesc, bdesc, idesc, nmma, a_step, b_step, n_acc, acc_cols,
                                          d_out2);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_tc_lab(const float* a_img, int a_floats, const float* b_img, int b_floats, unsigned long long adesc,
                  unsigned long long bdesc, unsigned int idesc, int nmma, int a_step, int b_step, float* D, int ncols,
                  cudaStream_t s) {
  DNN_ARG_CHECK(a_floats > 0 && b_floats > 0 && ncols >= 8 && ncols <= 256 && nmma >= 1, "tc_lab: bad arguments");
  const size_t smem = (size_t)(((a_floats + 255) / 256) * 256 + b_floats) * 4 + 1024;
  DNN_ARG_CHECK(smem <= 200 * 1024, "tc_lab: operand images exceed 200 KB of shared memory");
  DNN_CUDA_CHECK(cudaFuncSetAttribute(tc_lab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tc_lab_kernel<<<1, 128, smem, s>>>(a_img, a_floats, b_img, b_floats, adesc, bdesc, idesc, nmma, a_step, b_step, D,
                                     ncols);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

}  // namespace dnnb200
