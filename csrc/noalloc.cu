// dnnb200_batch_grad_noalloc: the batch gradient of the CNN without a single allocation.
//
// Every other entry point that runs the fused forward/backward gets its scratch (per-image FC
// records, per-image conv-gradient partials, per-image losses) from an engine object, and the
// engine gets it from cudaMalloc. Some callers cannot do that:
//   * a process whose device memory is already owned by somebody else's allocator -- a serving
//     process that reserved the whole GPU up front, a framework whose caching allocator sits at
//     its cap -- where a raw cudaMalloc from a library either fails or steals memory that was
//     promised elsewhere;
//   * a latency-critical loop (or a callback from another runtime) where cudaMalloc/cudaFree are
//     off limits because they take driver locks and can synchronise the whole device.
// For them this file carries the scratch inside the library image itself: a file-scope
// `static __device__` array of fixed size. It exists once per GPU that loads the library and stays
// until the process exits (0.92 MiB), whether or not the entry point is ever called.
//
// Consequences the caller has to know (also in include/dnnb200.h):
//   * batch size is limited to DNNB200_NOALLOC_MAX_BATCH (64); larger batches are refused;
//   * all calls on a device share the one array: two calls whose kernels can overlap on the GPU
//     (different streams, no ordering between them) race on it. Serialise them;
//   * the array keeps whatever the previous call left. Nothing in it has to be zero on entry: every
//     word that is read in a call was written earlier in the same call, and the alignment gaps of
//     the flat layout are never read (the kernel writes literal zeros for them);
//   * the first call on each device looks the array's device address up
//     (cudaGetSymbolAddress, which may also load the module). Make that first call during start-up,
//     before any stream capture is begun.
//
// The math is the same as dnnb200_engine_grad_only: fwd_bwd_kernel (one CTA per image, reused
// unchanged through its launcher) followed by a reduction kernel that is the gradient half of
// update_kernel with the exchange and the optimizer removed.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../include/dnnb200.h"
#include "common.cuh"
#include "layout.h"

namespace dnnb200 {

// existing launcher from fused_step.cu (takes its scratch as plain pointers; allocates nothing)
int launch_fwd_bwd(bool train, const float* params, const float* x, const long long* labels, float* saved,
                   float* part1, float* part2, float* img_loss, int* img_correct, float* logits_out, int B,
                   int max_ctas, cudaStream_t stream);

constexpr int kNaMaxBatch = DNNB200_NOALLOC_MAX_BATCH;
constexpr int kPart2Stride = kConvPrefix - kOffW2;  // 2432
// carve-up of the scratch array, in floats
constexpr size_t kNaSaved = 0;                                              // [B][kSavStride]
constexpr size_t kNaPart1 = kNaSaved + (size_t)kNaMaxBatch * kSavStride;    // [B][kOffW2]
constexpr size_t kNaPart2 = kNaPart1 + (size_t)kNaMaxBatch * kOffW2;        // [B][kPart2Stride]
constexpr size_t kNaLoss = kNaPart2 + (size_t)kNaMaxBatch * kPart2Stride;   // [B]
constexpr size_t kNaCorrect = kNaLoss + kNaMaxBatch;                        // [B] (int)
constexpr size_t kNaFloats = kNaCorrect + kNaMaxBatch;                      // 241 792 floats = 967 168 bytes

// The scratch. In HOST code this name is only a handle for the runtime's symbol table, not a
// device address: it must go through cudaGetSymbolAddress before a kernel can be given it.
static __device__ __align__(256) float g_noalloc_scratch[kNaFloats];

static_assert(sizeof(float) * kNaFloats == DNNB200_NOALLOC_SCRATCH_BYTES, "header and scratch size disagree");

// Is flat index i one of the 62 006 real parameters (as opposed to an alignment gap)?
__device__ __forceinline__ bool is_real_param(int i) {
  if (i < kOffB1) return i < kOffW1 + 450;
  if (i < kOffW2) return i < kOffB1 + 6;
  if (i < kOffB2) return i < kOffW2 + 2400;
  if (i < kOffF1W) return i < kOffB2 + 16;
  if (i < kOffF1B) return true;
  if (i < kOffF2W) return i < kOffF1B + kF1;
  if (i < kOffF2B) return true;
  if (i < kOffF3W) return i < kOffF2B + kF2;
  if (i < kOffF3B) return i < kOffF3W + kF3 * kF2;
  return i < kOffF3B + kF3;
}

// d(loss)/d(param[i]) summed over the batch, from the records fwd_bwd_kernel left in the scratch
__device__ __forceinline__ float grad_of(const float* __restrict__ sc, int i, int B) {
  float g = 0.f;
  if (i < kOffW2) {
    const float* p = sc + kNaPart1 + i;
    for (int b = 0; b < B; ++b) g += p[(size_t)b * kOffW2];
    return g;
  }
  if (i < kConvPrefix) {
    const float* p = sc + kNaPart2 + (i - kOffW2);
    for (int b = 0; b < B; ++b) g += p[(size_t)b * kPart2Stride];
    return g;
  }
  int off_d, off_a = -1, j, k = 0;
  if (i < kOffF1B) { const int e = i - kOffF1W; j = e / kFlat; k = e - j * kFlat; off_d = kSavD1; off_a = kSavP2; }
  else if (i < kOffF2W) { j = i - kOffF1B; off_d = kSavD1; }
  else if (i < kOffF2B) { const int e = i - kOffF2W; j = e / kF1; k = e - j * kF1; off_d = kSavD2; off_a = kSavH1; }
  else if (i < kOffF3W) { j = i - kOffF2B; off_d = kSavD2; }
  else if (i < kOffF3B) { const int e = i - kOffF3W; j = e / kF2; k = e - j * kF2; off_d = kSavD3; off_a = kSavH2; }
  else { j = i - kOffF3B; off_d = kSavD3; }
  const float* sv = sc + kNaSaved;
  for (int b = 0; b < B; ++b) {
    const float d = sv[(size_t)b * kSavStride + off_d + j];
    g += off_a >= 0 ? d * sv[(size_t)b * kSavStride + off_a + k] : d;
  }
  return g;
}

__global__ void __launch_bounds__(256)
noalloc_grad_reduce_kernel(const float* __restrict__ scratch, int B, float* __restrict__ grads_out,
                           float* __restrict__ loss_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < kFlatNumel) grads_out[i] = is_real_param(i) ? grad_of(scratch, i, B) : 0.f;
  if (blockIdx.x == 0 && threadIdx.x < 32 && loss_out != nullptr) {
    float a = 0.f;
    for (int b = threadIdx.x; b < B; b += 32) a += scratch[kNaLoss + b];
    a = warp_sum(a);
    if (threadIdx.x == 0) *loss_out = a / (float)B;
  }
}

// device address of the scratch array, looked up once per device
static float* noalloc_scratch_ptr(int device, cudaError_t* err) {
  static std::mutex mu;
  static float* ptrs[64] = {};
  std::lock_guard<std::mutex> lock(mu);
  *err = cudaSuccess;
  if (device < 0 || device >= 64) { *err = cudaErrorInvalidDevice; return nullptr; }
  if (ptrs[device] == nullptr) {
    float* p = nullptr;
    *err = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);  // `device` is current
    if (*err != cudaSuccess) return nullptr;
    ptrs[device] = p;
  }
  return ptrs[device];
}

}  // namespace dnnb200

using namespace dnnb200;

extern "C" int dnnb200_batch_grad_noalloc(const float* params, const float* x, const int64_t* y, int B,
                                          float* grads_out, float* loss_out, void* stream) {
  DNN_ARG_CHECK(params && x && y && grads_out, "batch_grad_noalloc: null argument");
  DNN_ARG_CHECK(B >= 1, "batch_grad_noalloc: batch must be positive, got %d", B);
  DNN_ARG_CHECK(B <= kNaMaxBatch,
                "batch_grad_noalloc: batch %d exceeds the built-in scratch (max %d images, %zu bytes); "
                "split the batch or use dnnb200_engine_grad_only",
                B, kNaMaxBatch, (size_t)DNNB200_NOALLOC_SCRATCH_BYTES);
  int device = 0;
  DNN_CUDA_CHECK(cudaGetDevice(&device));
  cudaError_t err;
  float* sc = noalloc_scratch_ptr(device, &err);
  if (sc == nullptr) {
    set_last_error("batch_grad_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                   cudaGetErrorString(err));
    return kErrCuda;
  }
  cudaStream_t s = (cudaStream_t)stream;
  int rc = launch_fwd_bwd(true, params, x, (const long long*)y, sc + kNaSaved, sc + kNaPart1, sc + kNaPart2,
                          sc + kNaLoss, reinterpret_cast<int*>(sc + kNaCorrect), nullptr, B, kNaMaxBatch, s);
  if (rc) return rc;
  noalloc_grad_reduce_kernel<<<(kFlatNumel + 255) / 256, 256, 0, s>>>(sc, B, grads_out, loss_out);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}
