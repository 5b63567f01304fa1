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
