// Small device/host helpers shared by all translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace dnnb200 {

// Error codes returned by every extern "C" entry point (0 = ok).
enum : int {
  kOk = 0,
  kErrCuda = 1,      // a CUDA runtime call failed (message via dnnb200_last_error)
  kErrArg = 2,       // invalid argument / unsupported shape
  kErrState = 3,     // object not in the right state
  kErrTimeout = 4,   // a peer did not show up in time
};

void set_last_error(const char* fmt, ...);

#define DNN_CUDA_CHECK(expr)                                                         \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ::dnnb200::set_last_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, \
                                cudaGetErrorString(_e));                             \
      return ::dnnb200::kErrCuda;                                                    \
    }                                                                                \
  } while (0)

#define DNN_ARG_CHECK(cond, ...)               \
  do {                                         \
    if (!(cond)) {                             \
      ::dnnb200::set_last_error(__VA_ARGS__);  \
      return ::dnnb200::kErrArg;               \
    }                                          \
  } while (0)

// Kernel function attributes (cudaFuncSetAttribute) belong to the DEVICE that is current when they are set, so
// "done once per process" is wrong as soon as a process drives a second GPU: its launches with more than 48 KB of
// dynamic shared memory would fail. One of these per kernel remembers which devices have been configured.
struct PerDeviceOnce {
  static constexpr int kMaxDevices = 64;
  volatile unsigned char done[kMaxDevices] = {};
};
template <class Kernel>
inline int ensure_max_dynamic_smem(PerDeviceOnce& once, Kernel kern, size_t bytes) {
  int dev = 0;
  DNN_CUDA_CHECK(cudaGetDevice(&dev));
  const bool tracked = dev >= 0 && dev < PerDeviceOnce::kMaxDevices;
  if (tracked && once.done[dev]) return kOk;
  // (two threads racing here both set the same value: harmless)
  DNN_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  if (tracked) once.done[dev] = 1;
  return kOk;
}

#ifdef __CUDACC__

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- system-scope flag traffic for the peer-memory exchange -------------------
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// gpu-scope pair for flags that only this GPU's kernels exchange (no NVLink round trip in the fence)
__device__ __forceinline__ void st_release_gpu(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_relaxed_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// L2-coherent loads (skip the non-coherent L1) for data a peer GPU wrote.
__device__ __forceinline__ float4 ld_cg_f4(const float4* p) { return __ldcg(p); }
__device__ __forceinline__ float ld_cg_f(const float* p) { return __ldcg(p); }

// Programmatic dependent launch: a kernel launched with the programmatic-stream-serialization attribute may
// start while its predecessor in the stream is still running; it must not touch the predecessor's results
// before pdl_wait(). pdl_trigger() in the predecessor says "my successor may start being scheduled now".
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

#endif  // __CUDACC__

}  // namespace dnnb200
