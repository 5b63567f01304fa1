// Host side of TMA: 2-D tensor maps over row-major fp32 matrices, encoded through the driver entry
// point (no link dependency on libcuda). Shared by the tensor-core linear kernels.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "common.cuh"

namespace dnnb200 {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// Row-major fp32 matrix [rows][cols] (cols * 4 must be a multiple of 16 bytes); box = [box_rows][32 floats],
// 128-byte swizzle, elements outside the matrix read as zero. mn_major selects the swizzle with 32-byte
// atomicity (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B), the only layout tcgen05.mma accepts for a 4-byte operand
// whose contiguous index is M/N rather than K.
inline int encode_tmap_2d(CUtensorMap* tm, const float* base, int rows, int cols, int box_rows, bool mn_major) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) { set_last_error("cuTensorMapEncodeTiled is not available from this driver"); return kErrCuda; }
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * sizeof(float)};
  const cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE,
                         mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return kErrCuda; }
  return kOk;
}

// A tensor map depends only on (address, shape, box, swizzle); training loops present the same few matrices
// again and again (weights always, activations whenever the caching allocator returns the same block), so the
// last encodings are kept and a hit costs a 128-byte copy instead of a driver call.
inline int make_tmap_2d(CUtensorMap* tm, const float* base, int rows, int cols, int box_rows, bool mn_major = false) {
  struct Entry { const float* base; int rows, cols, box_rows, mn; CUtensorMap map; };
  constexpr int kEntries = 32;
  static Entry cache[kEntries];
  static int used = 0, next = 0;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < used; ++i) {
    const Entry& e = cache[i];
    if (e.base == base && e.rows == rows && e.cols == cols && e.box_rows == box_rows && e.mn == (int)mn_major) {
      *tm = e.map;
      return kOk;
    }
  }
  const int rc = encode_tmap_2d(tm, base, rows, cols, box_rows, mn_major);
  if (rc) return rc;
  Entry& e = cache[next];
  e.base = base; e.rows = rows; e.cols = cols; e.box_rows = box_rows; e.mn = (int)mn_major; e.map = *tm;
  next = (next + 1) % kEntries;
  if (used < kEntries) ++used;
  return kOk;
}

}  // namespace dnnb200
