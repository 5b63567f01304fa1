// Host runtime: peer group (CUDA IPC symmetric buffers), fused step engine
// (buffers, launch sequencing, host-fed input ring) and the extern "C" surface
// declared in include/dnnb200.h.
//
// Reference counterparts: the mpi4py bootstrap + send/recv star
// (data_parallelism_train.py:60-62,118,135,210,227), the child training loop
// (:185-213) and eval (:157-183). The reference has no native code;
// everything here is new.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <vector>

#include "../include/dnnb200.h"
#include "common.cuh"
#include "layout.h"
#include "step_types.h"

namespace dnnb200 {

// ---- error plumbing ---------------------------------------------------------
static thread_local char g_err[512] = "";
void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- launchers implemented in the .cu files -----------------------------------
int launch_fwd_bwd(bool train, const float* params, const float* x, const long long* labels, float* saved,
                   float* part1, float* part2, float* img_loss, int* img_correct, float* logits_out, int B,
                   int max_ctas, cudaStream_t stream);
int launch_fwd_bwd_cluster(int cs, int mode, const float* params, const float* x, const long long* labels,
                           float* saved, float* part1, float* part2, float* img_loss, int* img_correct, int B,
                           int max_clusters, long long* trace, const unsigned int* ready_ctr, unsigned int ready_expect,
                           const float* packed, const unsigned int* in_flag, unsigned int in_expect, cudaStream_t stream);
int max_active_clusters(int cs, int mode);
int launch_pingpong(const RankCtx* d_ctx, int peer, int iters, unsigned int base_seq, long long* d_out_ns,
                    cudaStream_t stream);
int launch_ll128_probe(const RankCtx* d_ctx, int peer, int iters, int n_ctas, unsigned int base_seq,
                       unsigned long long* d_out3, cudaStream_t stream);
int launch_update(const RankCtx* d_ctxs, int n_virtual_ranks, int B, int n_part1, int loss_slot, int flags,
                  float* host_loss, cudaStream_t stream);
int launch_eval_reduce(const float* img_loss, const int* img_correct, int N, int bs, double* out, cudaStream_t stream);
int launch_allreduce(const RankCtx* d_ctxs, int n_virtual_ranks, const float* const* in, float* const* out, int n,
                     int mode, int two_shot, cudaStream_t stream);
int launch_conv_relu_pool_fwd(const float*, const float*, const float*, float*, unsigned char*, int, int, int, int,
                              int, cudaStream_t);
int launch_conv_relu_pool_bwd(const float*, const float*, const float*, const unsigned char*, const float*, float*,
                              float*, float*, int, int, int, int, int, int, cudaStream_t);
int launch_linear_fwd(const float*, const float*, const float*, float*, int, int, int, int, cudaStream_t);
int launch_linear_bwd(const float*, const float*, const float*, const float*, float*, float*, float*, int, int, int,
                      int, cudaStream_t);
int launch_softmax_xent(const float*, const long long*, float*, float*, int, int, cudaStream_t);
int launch_sgd_momentum(float*, const float*, float*, int, float, float, int, float, cudaStream_t);
int launch_tc_conv_relu_pool_fwd(const float*, const float*, const float*, float*, unsigned char*, int, int, int, int,
                                 int, cudaStream_t);
int launch_tc_conv_wgrad(const float*, const float*, const unsigned char*, const float*, float*, float*, int, int, int, int,
                         int, cudaStream_t);
int launch_tc_gemm_probe(const float*, const float*, float*, int, cudaStream_t);
int launch_tc_linear_fwd(const float*, const float*, const float*, float*, int, int, int, int, cudaStream_t);
int launch_tc_lab(const float*, int, const float*, int, unsigned long long, unsigned long long, unsigned int, int, int,
                  int, float*, int, cudaStream_t);
int launch_tc_lab_time(int, int, unsigned long long, unsigned long long, unsigned int, int, int, int, int, int, long long*,
                       cudaStream_t);
int launch_tc_linear_bwd(const float*, const float*, const float*, const float*, float*, float*, float*, float*, int,
                         int, int, int, cudaStream_t);

// ---- tiny control kernels ---------------------------------------------------
__global__ void ctl_set_hparams_kernel(StepCtl* c, float lr, float mu) { c->lr = lr; c->mu = mu; }
__global__ void ctl_set_membership_kernel(StepCtl* c, unsigned int mask, unsigned long long tmo) {
  c->active_mask = mask;
  c->timeout_ns = tmo;
  c->status = 0u;
}
__global__ void ctl_set_divisor_kernel(StepCtl* c, unsigned int d) { c->divisor_override = d; }
__global__ void ctl_set_variant_kernel(StepCtl* c, unsigned int variant) { c->two_shot = variant; }
__global__ void ctl_reset_loss_kernel(StepCtl* c) { c->loss_sum = 0.f; c->loss_steps = 0u; }
__global__ void ctl_set_step_kernel(StepCtl* c, unsigned long long step) { c->step = step; c->ctas_done = 0u; }
// leave marker = first sequence number this rank will not take part in
__global__ void leave_kernel(const RankCtx* cx, int clear) {
  const int p = threadIdx.x;
  if (p >= cx->world || p == cx->rank) return;
  const unsigned int marker = clear ? 0u : (unsigned int)(cx->ctl->step + 1ull);
  __threadfence_system();
  st_release_sys(cx->flags[p] + cx->rank * kFlagRow + kLeaveSlot, marker);
}

// ============================================================================
// Peer group
// ============================================================================
constexpr size_t kFlagBytes = (size_t)2 * kMaxWorld * kFlagRow * sizeof(unsigned int);  // two banks

}  // namespace dnnb200

struct dnnb200_group {
  int rank = 0, world = 1, device = 0;
  void* base = nullptr;  // [flags | recv]
  size_t bytes = 0;
  void* peer_base[dnnb200::kMaxWorld] = {};
  bool ipc_opened[dnnb200::kMaxWorld] = {};
  bool connected = false;
};

struct dnnb200_engine {
  int device = 0, max_batch = 0;
  float *params = nullptr, *momentum = nullptr, *grads = nullptr;
  float *saved = nullptr, *part1 = nullptr, *part2 = nullptr, *img_loss = nullptr;
  float* packed = nullptr;         // conv weights as tensor-core B operands (layout.h), refreshed by every update kernel
  long long* trace = nullptr;      // development aid: per-CTA stage timestamps of the cluster kernel
  unsigned int updates_launched = 0;   // update kernels launched on d_ctx == StepCtl::updates_done once they finish
  int debug_flags = 0;             // OR-ed into the update kernel's flags (bit 1: no early launch of the successor)
  int forced_cluster = 0;          // 0 = pick per batch size
  int max_clusters[9] = {};        // resident clusters of the fwd/bwd kernel per cluster size (index = CS), for `tf32`
  int max_clusters_p[3][9] = {};   // the same per precision mode
  int tf32 = 0;                    // conv stages of the clustered kernel: 0 fp32 FFMA, 1 all three GEMM-shaped conv stages on
                                   // tcgen05 (TF32 operands), 2 fp32 FFMA except the conv2 data gradient (tcgen05)
  int* img_correct = nullptr;
  float* eval_loss = nullptr;
  int* eval_correct = nullptr;
  int eval_cap = 0;
  double* d_eval_out = nullptr;
  dnnb200::StepCtl* ctl = nullptr;
  float* loss_ring = nullptr;
  int loss_ring_mask = 1023;
  dnnb200::RankCtx h_ctx;
  dnnb200::RankCtx* d_ctx = nullptr;
  dnnb200::RankCtx* h_ctx_pinned = nullptr;
  dnnb200_group* group = nullptr;
  unsigned long long launches = 0;
  unsigned long long steps_launched = 0;  // host mirror: update launches (loss ring slot)
  int num_sms = 148;
  // host-fed ring
  static constexpr int kRing = 32;   // device input slots
  static constexpr int kGroup = 16;  // steps per event group (stream events break the kernels' programmatic overlap and
                                     // the chaining of steps, so groups are long; two regions of the ring alternate)
  cudaStream_t copy_stream = nullptr, d2h_stream = nullptr;
  cudaEvent_t slot_ready[kRing] = {}, slot_free[kRing] = {}, step_done[kRing] = {};
  float* ring_x = nullptr;        // [kRing / kGroup] group regions of kGroup * max_batch images; a call packs its steps at its own batch size
  long long* ring_y = nullptr;    // [kRing][max_batch]
  bool ring_ready = false;
  unsigned long long ring_uses = 0;
  // event-free variant of the host-fed path (clustered kernel): "inputs landed" flags written by the copy stream,
  // "region free" known from a progress word the update kernel stores to pinned host memory
  unsigned int* d_in_flags = nullptr;            // [kRing / kGroup] device words
  unsigned int* h_flag_vals = nullptr;           // pinned: sources of the 4-byte flag copies (a ring of 64)
  volatile unsigned int* h_progress = nullptr;   // pinned, written by the update kernel (RankCtx::host_progress)
  unsigned int flag_seq = 0;
  unsigned int region_busy_until[kRing] = {};    // updates_launched after the last step that reads the region
  bool region_used[kRing] = {};
  int flag_group = 2;                            // steps per group of the event-free variant (kRing / flag_group regions)
  unsigned long long flag_uses = 0;              // groups handed out by the event-free variant
  int last_host_path = -1;                       // 1: the event-free variant ran last, 0: the event variant
  unsigned int pingpong_seq = 1000;
};

namespace dnnb200 {

static int push_ctx(dnnb200_engine* e, cudaStream_t s) {
  // RankCtx lives in device memory; stage through pinned memory so the copy is truly async
  *e->h_ctx_pinned = e->h_ctx;
  DNN_CUDA_CHECK(cudaMemcpyAsync(e->d_ctx, e->h_ctx_pinned, sizeof(RankCtx), cudaMemcpyHostToDevice, s));
  // the pinned staging copy is reused: make sure it is consumed before anyone rewrites it
  DNN_CUDA_CHECK(cudaStreamSynchronize(s));
  return kOk;
}

static int group_fill(dnnb200_group* g, int rank, int world, void* h_handle64) {
  g->rank = rank;
  g->world = world;
  DNN_CUDA_CHECK(cudaGetDevice(&g->device));
  g->bytes = kFlagBytes + (size_t)2 * world * 2 * kFlatNumel * sizeof(float);  // LL lines: 2 floats per value
  DNN_CUDA_CHECK(cudaMalloc(&g->base, g->bytes));
  DNN_CUDA_CHECK(cudaMemset(g->base, 0, g->bytes));
  DNN_CUDA_CHECK(cudaDeviceSynchronize());
  g->peer_base[rank] = g->base;
  if (h_handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == DNNB200_IPC_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t h;
    DNN_CUDA_CHECK(cudaIpcGetMemHandle(&h, g->base));
    memcpy(h_handle64, &h, sizeof(h));
  }
  return kOk;
}

static int group_create(int rank, int world, dnnb200_group** out, void* h_handle64) {
  DNN_ARG_CHECK(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "group: bad rank/world %d/%d", rank,
                world);
  DNN_ARG_CHECK(out != nullptr, "group: null output pointer");
  dnnb200_group* g = new (std::nothrow) dnnb200_group();
  DNN_ARG_CHECK(g != nullptr, "group: out of host memory");
  const int rc = group_fill(g, rank, world, h_handle64);
  if (rc) {                       // nothing half-built survives a failure (the error text is already set)
    if (g->base) cudaFree(g->base);
    delete g;
    return rc;
  }
  *out = g;
  return kOk;
}

static int engine_refresh_peers(dnnb200_engine* e, cudaStream_t s) {
  dnnb200_group* g = e->group;
  for (int p = 0; p < kMaxWorld; ++p) { e->h_ctx.recv[p] = nullptr; e->h_ctx.flags[p] = nullptr; }
  if (g) {
    for (int p = 0; p < g->world; ++p) {
      unsigned char* b = static_cast<unsigned char*>(g->peer_base[p]);
      e->h_ctx.flags[p] = reinterpret_cast<unsigned int*>(b);
      e->h_ctx.recv[p] = reinterpret_cast<float*>(b + kFlagBytes);
    }
    e->h_ctx.rank = g->rank;
    e->h_ctx.world = g->world;
  } else {
    e->h_ctx.rank = 0;
    e->h_ctx.world = 1;
  }
  return push_ctx(e, s);
}

static void free_ring(dnnb200_engine* e) {   // tolerates a ring that was only partly created
  for (int i = 0; i < dnnb200_engine::kRing; ++i) {
    if (e->slot_ready[i]) cudaEventDestroy(e->slot_ready[i]);
    if (e->slot_free[i]) cudaEventDestroy(e->slot_free[i]);
    if (e->step_done[i]) cudaEventDestroy(e->step_done[i]);
    e->slot_ready[i] = e->slot_free[i] = e->step_done[i] = nullptr;
  }
  cudaFree(e->ring_x);
  cudaFree(e->ring_y);
  cudaFree(e->d_in_flags);
  if (e->h_flag_vals) cudaFreeHost(e->h_flag_vals);
  if (e->h_progress) cudaFreeHost(const_cast<unsigned int*>(e->h_progress));
  e->ring_x = nullptr;
  e->ring_y = nullptr;
  e->d_in_flags = nullptr;
  e->h_flag_vals = nullptr;
  e->h_progress = nullptr;
  e->h_ctx.host_progress = nullptr;
  if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
  if (e->d2h_stream) cudaStreamDestroy(e->d2h_stream);
  e->copy_stream = e->d2h_stream = nullptr;
  e->ring_ready = false;
}

static int create_ring(dnnb200_engine* e) {
  DNN_CUDA_CHECK(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
  DNN_CUDA_CHECK(cudaStreamCreateWithFlags(&e->d2h_stream, cudaStreamNonBlocking));
  for (int i = 0; i < dnnb200_engine::kRing; ++i) {
    DNN_CUDA_CHECK(cudaEventCreateWithFlags(&e->slot_ready[i], cudaEventDisableTiming));
    DNN_CUDA_CHECK(cudaEventCreateWithFlags(&e->slot_free[i], cudaEventDisableTiming));
    DNN_CUDA_CHECK(cudaEventCreateWithFlags(&e->step_done[i], cudaEventDisableTiming));
  }
  // ONE allocation for all slots: the inputs of consecutive steps are consecutive in the caller's pinned buffer and
  // in the ring, so a whole group of steps is moved by one copy for the images and one for the labels
  DNN_CUDA_CHECK(cudaMalloc(&e->ring_x, (size_t)dnnb200_engine::kRing * e->max_batch * kImgNumel * sizeof(float)));
  DNN_CUDA_CHECK(cudaMalloc(&e->ring_y, (size_t)dnnb200_engine::kRing * e->max_batch * sizeof(long long)));
  constexpr int kRegions = dnnb200_engine::kRing;   // (one flag word per slot: any group size divides the ring)
  DNN_CUDA_CHECK(cudaMalloc(&e->d_in_flags, kRegions * sizeof(unsigned int)));
  DNN_CUDA_CHECK(cudaMemset(e->d_in_flags, 0, kRegions * sizeof(unsigned int)));
  if (const char* g = getenv("DNN_HOST_GROUP")) {   // development: steps per group of the event-free host-fed path
    const int v = atoi(g);
    if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) e->flag_group = v;
  }
  DNN_CUDA_CHECK(cudaMallocHost(&e->h_flag_vals, 64 * sizeof(unsigned int)));
  unsigned int* prog = nullptr;
  DNN_CUDA_CHECK(cudaHostAlloc(&prog, sizeof(unsigned int), cudaHostAllocMapped));
  *prog = e->updates_launched;
  e->h_progress = prog;
  DNN_CUDA_CHECK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&e->h_ctx.host_progress), prog, 0));
  return push_ctx(e, nullptr);   // (synchronises: happens once, when the host-fed path is first used)
}

static int ensure_ring(dnnb200_engine* e) {
  if (e->ring_ready) return kOk;
  const int rc = create_ring(e);
  if (rc) {          // a retry must not find (and leak) the slots of a failed attempt
    free_ring(e);
    cudaGetLastError();
    return rc;
  }
  e->ring_ready = true;
  return kOk;
}

// Cluster size for a batch: the largest one whose B clusters are all resident at once, so the
// whole batch runs as ONE wave over as many SMs as possible; 1 = the one-CTA-per-image kernel.
static int cluster_size_for(const dnnb200_engine* e, int B) {
  if (e->forced_cluster) return e->forced_cluster;
  for (int cs = 8; cs >= 2; --cs)
    if (e->max_clusters[cs] >= B) return cs;
  return 1;
}

// the forward/backward half of a step; returns the number of conv1 partial rows it produced
// chained: the previous operation this engine put on the stream is the update kernel of the previous step and
// nothing else came between (inner iterations of the multi-step calls). Only then may the kernel wait on the
// update counter instead of on the completion of the previous grid (fused_cluster.cu).
static int fwd_bwd_any(dnnb200_engine* e, const float* x, const long long* y, int B, cudaStream_t s, int* n_part1,
                       bool chained = false, const unsigned int* in_flag = nullptr, unsigned int in_expect = 0u) {
  const int cs = cluster_size_for(e, B);
  *n_part1 = B;  // clusters reduce their conv1 partials to one row per image before they write them
  if (cs == 1)
    return launch_fwd_bwd(true, e->params, x, y, e->saved, e->part1, e->part2, e->img_loss, e->img_correct, nullptr,
                          B, e->num_sms, s);
  return launch_fwd_bwd_cluster(cs, e->tf32, e->params, x, y, e->saved, e->part1, e->part2, e->img_loss, e->img_correct, B,
                                e->max_clusters[cs] > 0 ? e->max_clusters[cs] : 1, e->trace,
                                chained && !(e->debug_flags & 8) ? &e->ctl->updates_done : nullptr, e->updates_launched,
                                // ready-made operands are only trusted when the previous operation on the stream was this
                                // engine's own update kernel (which wrote them together with the parameters): the host
                                // may change the parameters between calls
                                chained && !(e->debug_flags & 128) ? e->packed : nullptr, in_flag, in_expect, s);
}

// The counter value a chained step waits for is a kernel ARGUMENT: baked into a CUDA graph it would be stale on
// every replay, so steps recorded by a stream capture keep the grid-completion dependency.
static inline bool stream_capturing(cudaStream_t s) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  return cudaStreamIsCapturing(s, &st) == cudaSuccess && st != cudaStreamCaptureStatusNone;
}

static inline int peers_flag(const dnnb200_engine* e) {
  if (e->h_ctx.world <= 1) return 0;
  return (int)kUpdFlagHasGroup | ((e->debug_flags & 32) ? 0 : (int)kUpdFlagPeers);   // debug bit 32: A/B switch
}

static int step_kernels(dnnb200_engine* e, const float* x, const long long* y, int B, int flags, cudaStream_t s,
                        bool chained = false) {
  int n_part1 = B;
  int rc = fwd_bwd_any(e, x, y, B, s, &n_part1, chained);
  if (rc) return rc;
  rc = launch_update(e->d_ctx, 1, B, n_part1, (int)(e->steps_launched & (unsigned long long)e->loss_ring_mask),
                     flags | e->debug_flags | peers_flag(e), nullptr, s);
  if (rc) return rc;
  e->updates_launched += 1u;
  e->launches += 2;
  return kOk;
}

}  // namespace dnnb200

using namespace dnnb200;

extern "C" {

const char* dnnb200_last_error(void) { return g_err; }
int dnnb200_version(void) { return 100; }

int dnnb200_layout(int* offsets10, int* flat_numel, int* slice, int* num_slices) {
  const int offs[10] = {kOffW1, kOffB1, kOffW2, kOffB2, kOffF1W, kOffF1B, kOffF2W, kOffF2B, kOffF3W, kOffF3B};
  if (offsets10) memcpy(offsets10, offs, sizeof(offs));
  if (flat_numel) *flat_numel = kFlatNumel;
  if (slice) *slice = kSlice;
  if (num_slices) *num_slices = kNumSlices;
  return kOk;
}

// ---- layer kernels ------------------------------------------------------------
int dnnb200_conv_relu_pool_fwd(const float* x, const float* w, const float* b, float* y, uint8_t* idx, int N, int Ci,
                               int H, int W, int Co, void* stream) {
  DNN_ARG_CHECK(N > 0 && Ci > 0 && Co > 0 && Ci * 25 * 4 <= 48 * 1024, "conv_relu_pool_fwd: bad shape");
  return launch_conv_relu_pool_fwd(x, w, b, y, idx, N, Ci, H, W, Co, (cudaStream_t)stream);
}
int dnnb200_conv_relu_pool_bwd(const float* x, const float* w, const float* y, const uint8_t* idx, const float* dy,
                               float* dx, float* dw, float* db, int N, int Ci, int H, int W, int Co, void* stream) {
  DNN_ARG_CHECK(N > 0 && Ci > 0 && Co > 0, "conv_relu_pool_bwd: bad shape");
  return launch_conv_relu_pool_bwd(x, w, y, idx, dy, dx, dw, db, N, Ci, H, W, Co, 0, (cudaStream_t)stream);
}
int dnnb200_conv_relu_pool_bwd_tc(const float* x, const float* w, const float* y, const uint8_t* idx, const float* dy,
                                  float* dx, float* dw, float* db, int N, int Ci, int H, int W, int Co, void* stream) {
  DNN_ARG_CHECK(N > 0 && Ci > 0 && Co > 0, "conv_relu_pool_bwd_tc: bad shape");
  return launch_conv_relu_pool_bwd(x, w, y, idx, dy, dx, dw, db, N, Ci, H, W, Co, 1, (cudaStream_t)stream);
}
int dnnb200_tc_conv_wgrad(const float* x, const float* y, const uint8_t* idx, const float* dy, float* dw, float* db, int N,
                          int Ci, int H, int W, int Co, void* stream) {
  DNN_ARG_CHECK(N > 0 && x && y && idx && dy && dw && db, "tc_conv_wgrad: bad arguments");
  return launch_tc_conv_wgrad(x, y, idx, dy, dw, db, N, Ci, H, W, Co, (cudaStream_t)stream);
}
int dnnb200_tc_conv_relu_pool_fwd(const float* x, const float* w, const float* b, float* y, uint8_t* idx, int N,
                                  int Ci, int H, int W, int Co, void* stream) {
  DNN_ARG_CHECK(N > 0, "tc_conv_relu_pool_fwd: empty batch");
  return launch_tc_conv_relu_pool_fwd(x, w, b, y, idx, N, Ci, H, W, Co, (cudaStream_t)stream);
}
int dnnb200_tc_linear_fwd(const float* x, const float* w, const float* b, float* y, int N, int M, int K, int relu,
                          void* stream) {
  DNN_ARG_CHECK(N > 0, "tc_linear_fwd: empty batch");
  return launch_tc_linear_fwd(x, w, b, y, N, M, K, relu, (cudaStream_t)stream);
}
int dnnb200_tc_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dy_scratch,
                          float* dx, float* dw, float* db, int N, int M, int K, int relu, void* stream) {
  DNN_ARG_CHECK(N > 0, "tc_linear_bwd: empty batch");
  return launch_tc_linear_bwd(x, w, y, dy, dy_scratch, dx, dw, db, N, M, K, relu, (cudaStream_t)stream);
}
int dnnb200_tc_lab(const float* a_img, int a_floats, const float* b_img, int b_floats, unsigned long long adesc,
                   unsigned long long bdesc, unsigned int idesc, int nmma, int a_step, int b_step, float* D, int ncols,
                   void* stream) {
  return launch_tc_lab(a_img, a_floats, b_img, b_floats, adesc, bdesc, idesc, nmma, a_step, b_step, D, ncols,
                       (cudaStream_t)stream);
}
int dnnb200_tc_lab_time(int a_bytes, int b_bytes, unsigned long long adesc, unsigned long long bdesc, unsigned int idesc,
                        int nmma, int a_step, int b_step, int n_acc, int acc_cols, long long* d_out2, void* stream) {
  return launch_tc_lab_time(a_bytes, b_bytes, adesc, bdesc, idesc, nmma, a_step, b_step, n_acc, acc_cols, d_out2,
                            (cudaStream_t)stream);
}
int dnnb200_tc_gemm_probe(const float* A, const float* B, float* D, int K, void* stream) {
  return launch_tc_gemm_probe(A, B, D, K, (cudaStream_t)stream);
}
int dnnb200_linear_fwd(const float* x, const float* w, const float* b, float* y, int N, int M, int K, int relu,
                       void* stream) {
  DNN_ARG_CHECK(N > 0 && M > 0 && K > 0, "linear_fwd: bad shape");
  return launch_linear_fwd(x, w, b, y, N, M, K, relu, (cudaStream_t)stream);
}
int dnnb200_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dx, float* dw,
                       float* db, int N, int M, int K, int relu, void* stream) {
  DNN_ARG_CHECK(N > 0 && M > 0 && K > 0, "linear_bwd: bad shape");
  return launch_linear_bwd(x, w, y, dy, dx, dw, db, N, M, K, relu, (cudaStream_t)stream);
}
int dnnb200_softmax_xent(const float* logits, const int64_t* target, float* loss, float* dlogits, int N, int C,
                         void* stream) {
  DNN_ARG_CHECK(N > 0 && C > 0, "softmax_xent: bad shape");
  return launch_softmax_xent(logits, (const long long*)target, loss, dlogits, N, C, (cudaStream_t)stream);
}
int dnnb200_sgd_momentum(float* p, const float* g, float* buf, int n, float lr, float mu, int first_step, float scale,
                         void* stream) {
  DNN_ARG_CHECK(n > 0, "sgd_momentum: n must be positive");
  return launch_sgd_momentum(p, g, buf, n, lr, mu, first_step, scale, (cudaStream_t)stream);
}

// ---- peer group -----------------------------------------------------------------
int dnnb200_group_create(int rank, int world, dnnb200_group** out, void* h_handle64) {
  return group_create(rank, world, out, h_handle64);
}

int dnnb200_group_connect(dnnb200_group* g, const void* h_all_handles) {
  DNN_ARG_CHECK(g && h_all_handles, "group_connect: null argument");
  const unsigned char* hs = static_cast<const unsigned char*>(h_all_handles);
  for (int p = 0; p < g->world; ++p) {
    if (p == g->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, hs + (size_t)p * DNNB200_IPC_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    DNN_CUDA_CHECK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    g->peer_base[p] = ptr;
    g->ipc_opened[p] = true;
  }
  g->connected = true;
  return kOk;
}

int dnnb200_group_connect_local(dnnb200_group** groups, int n) {
  DNN_ARG_CHECK(groups && n >= 1 && n <= kMaxWorld, "group_connect_local: bad arguments");
  for (int a = 0; a < n; ++a) {
    DNN_ARG_CHECK(groups[a] && groups[a]->world == n && groups[a]->rank == a,
                  "group_connect_local: group %d must have rank %d of world %d", a, a, n);
    for (int b = 0; b < n; ++b) {
      if (groups[a]->device != groups[b]->device) {
        int can = 0;
        DNN_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, groups[a]->device, groups[b]->device));
        DNN_ARG_CHECK(can, "group_connect_local: device %d cannot access device %d", groups[a]->device,
                      groups[b]->device);
        int cur = 0;
        DNN_CUDA_CHECK(cudaGetDevice(&cur));
        DNN_CUDA_CHECK(cudaSetDevice(groups[a]->device));
        const cudaError_t pe = cudaDeviceEnablePeerAccess(groups[b]->device, 0);
        cudaSetDevice(cur);
        if (pe == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else DNN_CUDA_CHECK(pe);
      }
      groups[a]->peer_base[b] = groups[b]->base;
    }
    groups[a]->connected = true;
  }
  return kOk;
}

int dnnb200_group_destroy(dnnb200_group* g) {
  if (!g) return kOk;
  for (int p = 0; p < g->world; ++p)
    if (g->ipc_opened[p] && g->peer_base[p]) cudaIpcCloseMemHandle(g->peer_base[p]);
  if (g->base) cudaFree(g->base);
  delete g;
  return kOk;
}

// ---- engine ---------------------------------------------------------------------
static int engine_fill(dnnb200_engine* e, float* params, float* momentum, float* grads, int max_batch, float lr,
                       float mu) {
  DNN_CUDA_CHECK(cudaGetDevice(&e->device));
  cudaDeviceProp prop;
  DNN_CUDA_CHECK(cudaGetDeviceProperties(&prop, e->device));
  e->num_sms = prop.multiProcessorCount;
  e->max_batch = max_batch;
  e->params = params;
  e->momentum = momentum;
  e->grads = grads;
  DNN_CUDA_CHECK(cudaMalloc(&e->saved, (size_t)max_batch * kSavStride * sizeof(float)));
  // gaps of the flat layout inside the partial buffers are never written: zero them once
  DNN_CUDA_CHECK(cudaMalloc(&e->part1, (size_t)max_batch * 8 * kOffW2 * sizeof(float)));
  DNN_CUDA_CHECK(cudaMemset(e->part1, 0, (size_t)max_batch * 8 * kOffW2 * sizeof(float)));
  DNN_CUDA_CHECK(cudaMalloc(&e->part2, (size_t)max_batch * (kConvPrefix - kOffW2) * sizeof(float)));
  DNN_CUDA_CHECK(cudaMemset(e->part2, 0, (size_t)max_batch * (kConvPrefix - kOffW2) * sizeof(float)));
  for (int cs = 2; cs <= 8; ++cs) {
    for (int m = 0; m < 3; ++m) e->max_clusters_p[m][cs] = max_active_clusters(cs, m);
    e->max_clusters[cs] = e->max_clusters_p[e->tf32][cs];
  }
  DNN_CUDA_CHECK(cudaMemset(e->saved, 0, (size_t)max_batch * kSavStride * sizeof(float)));
  DNN_CUDA_CHECK(cudaMalloc(&e->img_loss, (size_t)max_batch * sizeof(float)));
  DNN_CUDA_CHECK(cudaMalloc(&e->packed, (size_t)kPackNumel * sizeof(float)));
  DNN_CUDA_CHECK(cudaMemset(e->packed, 0, (size_t)kPackNumel * sizeof(float)));   // everything that is not a weight stays zero
  DNN_CUDA_CHECK(cudaMalloc(&e->img_correct, (size_t)max_batch * sizeof(int)));
  DNN_CUDA_CHECK(cudaMalloc(&e->ctl, sizeof(StepCtl)));
  DNN_CUDA_CHECK(cudaMalloc(&e->loss_ring, (size_t)(e->loss_ring_mask + 1) * sizeof(float)));
  DNN_CUDA_CHECK(cudaMemset(e->loss_ring, 0, (size_t)(e->loss_ring_mask + 1) * sizeof(float)));
  DNN_CUDA_CHECK(cudaMalloc(&e->d_ctx, sizeof(RankCtx)));
  DNN_CUDA_CHECK(cudaMalloc(&e->d_eval_out, 2 * sizeof(double)));
  DNN_CUDA_CHECK(cudaMallocHost(&e->h_ctx_pinned, sizeof(RankCtx)));
  StepCtl c;
  memset(&c, 0, sizeof(c));
  c.lr = lr;
  c.mu = mu;
  c.active_mask = 1u;
  c.timeout_ns = 0ull;
  DNN_CUDA_CHECK(cudaMemcpy(e->ctl, &c, sizeof(c), cudaMemcpyHostToDevice));
  memset(&e->h_ctx, 0, sizeof(e->h_ctx));
  e->h_ctx.params = params;
  e->h_ctx.momentum = momentum;
  e->h_ctx.grads = grads;
  e->h_ctx.saved = e->saved;
  e->h_ctx.part1 = e->part1;
  e->h_ctx.part2 = e->part2;
  e->h_ctx.img_loss = e->img_loss;
  e->h_ctx.packed = e->packed;
  e->h_ctx.loss_ring = e->loss_ring;
  e->h_ctx.loss_ring_mask = e->loss_ring_mask;
  e->h_ctx.ctl = e->ctl;
  return engine_refresh_peers(e, nullptr);
}

int dnnb200_engine_destroy(dnnb200_engine* e);

int dnnb200_engine_create(float* params, float* momentum, float* grads, int max_batch, float lr, float mu,
                          dnnb200_engine** out) {
  DNN_ARG_CHECK(params && momentum && out, "engine_create: null argument");
  DNN_ARG_CHECK(max_batch >= 1 && max_batch <= 65536, "engine_create: max_batch out of range");
  dnnb200_engine* e = new (std::nothrow) dnnb200_engine();
  DNN_ARG_CHECK(e != nullptr, "engine_create: out of host memory");
  const int rc = engine_fill(e, params, momentum, grads, max_batch, lr, mu);
  if (rc) {   // e.g. out of device memory at a large max_batch: give back what was allocated so far
    char keep[sizeof(g_err)];
    memcpy(keep, g_err, sizeof(keep));
    dnnb200_engine_destroy(e);
    cudaGetLastError();
    memcpy(g_err, keep, sizeof(keep));
    return rc;
  }
  *out = e;
  return kOk;
}

int dnnb200_engine_destroy(dnnb200_engine* e) {
  if (!e) return kOk;
  cudaDeviceSynchronize();
  cudaFree(e->packed);
  cudaFree(e->saved); cudaFree(e->part1); cudaFree(e->part2); cudaFree(e->img_loss); cudaFree(e->img_correct);
  cudaFree(e->eval_loss); cudaFree(e->eval_correct); cudaFree(e->d_eval_out);
  cudaFree(e->ctl); cudaFree(e->loss_ring); cudaFree(e->d_ctx);
  if (e->h_ctx_pinned) cudaFreeHost(e->h_ctx_pinned);
  free_ring(e);
  delete e;
  return kOk;
}

int dnnb200_engine_attach_group(dnnb200_engine* e, dnnb200_group* g) {
  DNN_ARG_CHECK(e, "engine_attach_group: null engine");
  if (g) {
    DNN_ARG_CHECK(g->connected || g->world == 1, "engine_attach_group: group is not connected yet");
  }
  e->group = g;
  int rc = engine_refresh_peers(e, nullptr);
  if (rc) return rc;
  const unsigned int mask = g ? ((g->world >= 32) ? 0xffffffffu : ((1u << g->world) - 1u)) : 1u;
  ctl_set_membership_kernel<<<1, 1>>>(e->ctl, mask, 0ull);
  DNN_CUDA_CHECK(cudaGetLastError());
  DNN_CUDA_CHECK(cudaDeviceSynchronize());
  return kOk;
}

int dnnb200_engine_set_hparams(dnnb200_engine* e, float lr, float mu, void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  ctl_set_hparams_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(e->ctl, lr, mu);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int dnnb200_engine_set_membership(dnnb200_engine* e, unsigned int active_mask, unsigned long long timeout_ns,
                                  void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  ctl_set_membership_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(e->ctl, active_mask, timeout_ns);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int dnnb200_engine_set_exchange_variant(dnnb200_engine* e, int two_shot, void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  DNN_ARG_CHECK(two_shot >= 0 && two_shot <= 3,
                "exchange variant must be 0 (one-shot), 1 (two-shot), 2 (NVLS) or 3 (one-shot, 128-byte lines)");
  DNN_ARG_CHECK(two_shot != 2 || e->h_ctx.nvls_mc != nullptr, "the NVLS variant needs dnnb200_engine_attach_nvls first");
  DNN_ARG_CHECK(two_shot != 3 || e->h_ctx.world <= 8, "the 128-byte-line variant handles at most 8 ranks, this group has %d",
                e->h_ctx.world);
  ctl_set_variant_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(e->ctl, (unsigned int)two_shot);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int dnnb200_engine_set_divisor(dnnb200_engine* e, unsigned int divisor, void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  ctl_set_divisor_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(e->ctl, divisor);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int dnnb200_engine_zero_momentum(dnnb200_engine* e, void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  DNN_CUDA_CHECK(cudaMemsetAsync(e->momentum, 0, kFlatNumel * sizeof(float), (cudaStream_t)stream));
  return kOk;
}

int dnnb200_engine_reset_loss(dnnb200_engine* e, void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  ctl_reset_loss_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(e->ctl);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int dnnb200_engine_read_ctl(dnnb200_engine* e, dnnb200_ctl* h_out, void* stream) {
  DNN_ARG_CHECK(e && h_out, "null argument");
  static_assert(sizeof(dnnb200_ctl) == sizeof(StepCtl), "dnnb200_ctl must mirror StepCtl");
  DNN_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
  DNN_CUDA_CHECK(cudaMemcpy(h_out, e->ctl, sizeof(StepCtl), cudaMemcpyDeviceToHost));
  return kOk;
}

unsigned long long dnnb200_engine_launches(const dnnb200_engine* e) { return e ? e->launches : 0ull; }

int dnnb200_engine_set_cluster_size(dnnb200_engine* e, int cs) {
  DNN_ARG_CHECK(e, "null engine");
  DNN_ARG_CHECK(cs >= 0 && cs <= 8, "cluster size must be 0 (auto) or 1..8");
  DNN_ARG_CHECK(cs <= 1 || e->max_clusters[cs] > 0, "clusters of %d CTAs cannot be scheduled on this device", cs);
  e->forced_cluster = cs;
  return kOk;
}

int dnnb200_engine_set_precision(dnnb200_engine* e, int tf32) {
  DNN_ARG_CHECK(e, "null engine");
  DNN_ARG_CHECK(tf32 >= 0 && tf32 <= 2,
                "precision must be 0 (fp32 FFMA), 1 (TF32 tensor cores) or 2 (fp32 with the conv2 data gradient on tensor cores)");
  e->tf32 = tf32;
  for (int cs = 2; cs <= 8; ++cs) e->max_clusters[cs] = e->max_clusters_p[tf32][cs];
  if (e->forced_cluster > 1 && e->max_clusters[e->forced_cluster] <= 0) e->forced_cluster = 0;
  return kOk;
}

int dnnb200_engine_set_trace(dnnb200_engine* e, long long* d_trace) {
  DNN_ARG_CHECK(e, "null engine");
  e->trace = d_trace;
  return kOk;
}

int dnnb200_engine_pingpong(dnnb200_engine* e, int peer, int iters, double* h_one_way_us, void* stream) {
  DNN_ARG_CHECK(e && h_one_way_us && e->group && e->group->world > 1, "pingpong needs an attached group of >= 2 ranks");
  DNN_ARG_CHECK(peer >= 0 && peer < e->group->world && peer != e->group->rank && iters > 0, "pingpong: bad peer/iters");
  // every call uses fresh sequence numbers (both ranks call in lockstep, so their engines count alike)
  int rc = launch_pingpong(e->d_ctx, peer, iters, e->pingpong_seq, reinterpret_cast<long long*>(e->d_eval_out),
                           (cudaStream_t)stream);
  e->pingpong_seq += (unsigned int)iters;
  if (rc) return rc;
  long long ns = 0;
  DNN_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
  DNN_CUDA_CHECK(cudaMemcpy(&ns, e->d_eval_out, sizeof(ns), cudaMemcpyDeviceToHost));
  *h_one_way_us = (double)ns / 1e3 / (2.0 * iters);
  e->launches += 1;
  return kOk;
}

int dnnb200_engine_ll128_probe(dnnb200_engine* e, int peer, int iters, int n_ctas, unsigned long long* h_out3,
                               void* stream) {
  DNN_ARG_CHECK(e && h_out3 && e->group && e->group->world > 1, "ll128_probe needs an attached group of >= 2 ranks");
  DNN_ARG_CHECK(peer >= 0 && peer < e->group->world && peer != e->group->rank && iters > 0, "ll128_probe: bad peer/iters");
  cudaStream_t s = (cudaStream_t)stream;
  unsigned long long* d_out = nullptr;
  DNN_CUDA_CHECK(cudaMalloc(&d_out, 3 * sizeof(unsigned long long)));
  // sequence numbers far above anything a training run reaches, so a leftover flag word can never be taken for a step's
  int rc = launch_ll128_probe(e->d_ctx, peer, iters, n_ctas, 0x40000000u + e->pingpong_seq, d_out, s);
  e->pingpong_seq += (unsigned int)iters;
  if (rc == kOk) {
    cudaError_t ce = cudaStreamSynchronize(s);
    if (ce == cudaSuccess) ce = cudaMemcpy(h_out3, d_out, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    if (ce != cudaSuccess) { set_last_error("ll128_probe: %s", cudaGetErrorString(ce)); rc = kErrCuda; }
  }
  cudaFree(d_out);
  if (rc) return rc;
  e->launches += 1;
  return kOk;
}

int dnnb200_engine_attach_nvls(dnnb200_engine* e, float* local_buf, float* multicast_buf) {
  DNN_ARG_CHECK(e, "null engine");
  e->h_ctx.nvls_local = local_buf;
  e->h_ctx.nvls_mc = multicast_buf;
  return push_ctx(e, nullptr);
}

int dnnb200_engine_set_debug_flags(dnnb200_engine* e, int flags) {
  DNN_ARG_CHECK(e, "null engine");
  e->debug_flags = flags;
  return kOk;
}

int dnnb200_engine_set_update_trace(dnnb200_engine* e, long long* d_trace) {
  DNN_ARG_CHECK(e, "null engine");
  e->h_ctx.trace = d_trace;
  return push_ctx(e, nullptr);
}

int dnnb200_engine_cluster_info(const dnnb200_engine* e, int B, int* chosen, int* max_clusters9) {
  DNN_ARG_CHECK(e, "null engine");
  if (chosen) *chosen = cluster_size_for(e, B);
  if (max_clusters9) memcpy(max_clusters9, e->max_clusters, sizeof(e->max_clusters));
  return kOk;
}

int dnnb200_engine_step(dnnb200_engine* e, const float* x, const int64_t* y, int B, void* stream) {
  DNN_ARG_CHECK(e && x && y, "engine_step: null argument");
  DNN_ARG_CHECK(B >= 1 && B <= e->max_batch, "engine_step: batch %d exceeds max_batch %d", B, e->max_batch);
  int rc = step_kernels(e, x, (const long long*)y, B, 0, (cudaStream_t)stream);
  if (rc) return rc;
  e->steps_launched += 1;
  return kOk;
}

int dnnb200_engine_train_device(dnnb200_engine* e, const float* x, const int64_t* y, int B, int nsteps, void* stream) {
  DNN_ARG_CHECK(e && x && y, "engine_train_device: null argument");
  DNN_ARG_CHECK(B >= 1 && B <= e->max_batch, "engine_train_device: batch %d exceeds max_batch %d", B, e->max_batch);
  const bool may_chain = !stream_capturing((cudaStream_t)stream);
  for (int t = 0; t < nsteps; ++t) {
    int rc = step_kernels(e, x + (size_t)t * B * kImgNumel, (const long long*)y + (size_t)t * B, B, 0,
                          (cudaStream_t)stream, /*chained=*/may_chain && t > 0);
    if (rc) return rc;
    e->steps_launched += 1;
  }
  return kOk;
}

int dnnb200_engine_grad_only(dnnb200_engine* e, const float* x, const int64_t* y, int B, float* grads_out,
                             float* h_loss, void* stream) {
  DNN_ARG_CHECK(e && x && y && grads_out, "engine_grad_only: null argument");
  DNN_ARG_CHECK(B >= 1 && B <= e->max_batch, "engine_grad_only: batch %d exceeds max_batch %d", B, e->max_batch);
  cudaStream_t s = (cudaStream_t)stream;
  float* keep = e->h_ctx.grads;
  e->h_ctx.grads = grads_out;
  int rc = push_ctx(e, s);
  if (rc) return rc;
  rc = step_kernels(e, x, (const long long*)y, B, /*grad_only=*/1, s);
  if (rc) return rc;
  const int slot = (int)(e->steps_launched & (unsigned long long)e->loss_ring_mask);
  DNN_CUDA_CHECK(cudaStreamSynchronize(s));
  if (h_loss) DNN_CUDA_CHECK(cudaMemcpy(h_loss, e->loss_ring + slot, sizeof(float), cudaMemcpyDeviceToHost));
  e->h_ctx.grads = keep;
  return push_ctx(e, s);
}

// Host-fed steps without a single stream event on the compute stream. A group's inputs go out on the copy stream
// (images, labels, then a 4-byte sequence number into the region's flag word); the first forward/backward kernel that
// reads the region waits for that number itself (fused_cluster.cu), so every step of the call stays chained to its
// predecessor and programmatic dependent launch is never interrupted. (With an event pair per group, the two steps
// after every group boundary took 28 and 47 us instead of 22 at four GPUs -- bench/step_jitter.py with DNN_HOST=1 --
// and a boundary on ANY rank stalls all of them.) A region is overwritten only when the update kernel of the last
// step that read it has reported in through the pinned progress word, so the call runs at most two groups ahead of
// the GPU: it returns when everything is enqueued, which for long calls is shortly before the end.
static int train_host_flags(dnnb200_engine* e, const float* h_x, const long long* h_y, int B, int nsteps, float* d_losses,
                            cudaStream_t s) {
  // Groups are SMALL here (a boundary costs nothing): a 3 MB copy burst per 16 steps delays the GPU's fetch of the
  // kernel launches that share PCIe with it by tens of microseconds, and at N GPUs a late launch on any rank stalls
  // all of them. Measured at 4 GPUs, us per step host-fed vs device-resident 21.75: groups of 16 steps 23.67, of 4
  // 21.80, of 2 21.80, of 1 21.89; 20-step calls 27.3 / 26.2 / 24.8 / 26.1 (profiles/host_fed_groups_n4.txt).
  const int G = e->flag_group, kRegions = dnnb200_engine::kRing / G;
  const size_t xb = (size_t)B * kImgNumel * sizeof(float), yb = (size_t)B * sizeof(long long);
  for (int t0 = 0; t0 < nsteps;) {
    const int n = nsteps - t0 < G ? nsteps - t0 : G;
    const int region = (int)(e->flag_uses % (unsigned long long)kRegions);
    if (e->region_used[region]) {      // its last readers must be through (they were enqueued two groups ago)
      const unsigned int need = e->region_busy_until[region];
      unsigned long long spins = 0;
      while ((int)(*e->h_progress - need) < 0) {
        if ((++spins & 0xfffffull) == 0 && cudaStreamQuery(s) == cudaSuccess) break;   // everything has run: also fine
      }
    }
    float* gx = e->ring_x + (size_t)region * G * e->max_batch * kImgNumel;
    long long* gy = e->ring_y + (size_t)region * G * e->max_batch;
    const unsigned int seq = ++e->flag_seq;
    unsigned int* src = e->h_flag_vals + (seq & 63u);
    *src = seq;
    DNN_CUDA_CHECK(cudaMemcpyAsync(gx, h_x + (size_t)t0 * B * kImgNumel, (size_t)n * xb, cudaMemcpyHostToDevice, e->copy_stream));
    DNN_CUDA_CHECK(cudaMemcpyAsync(gy, h_y + (size_t)t0 * B, (size_t)n * yb, cudaMemcpyHostToDevice, e->copy_stream));
    DNN_CUDA_CHECK(cudaMemcpyAsync(e->d_in_flags + region, src, sizeof(unsigned int), cudaMemcpyHostToDevice, e->copy_stream));
    for (int k = 0; k < n; ++k) {
      const int t = t0 + k;
      int n_part1 = B;
      // (only the group's first kernel has to look at the flag: the steps behind it are ordered after it)
      int rc = fwd_bwd_any(e, gx + (size_t)k * B * kImgNumel, gy + (size_t)k * B, B, s, &n_part1, /*chained=*/t > 0,
                           k == 0 ? e->d_in_flags + region : nullptr, seq);
      if (rc) return rc;
      const int lslot = (int)(e->steps_launched & (unsigned long long)e->loss_ring_mask);
      rc = launch_update(e->d_ctx, 1, B, n_part1, lslot, e->debug_flags | peers_flag(e), d_losses ? d_losses + t : nullptr, s);
      if (rc) return rc;
      e->updates_launched += 1u;
      e->launches += 2;
      e->steps_launched += 1;
    }
    e->region_busy_until[region] = e->updates_launched;
    e->region_used[region] = true;
    e->flag_uses += 1;
    t0 += n;
  }
  return kOk;
}

int dnnb200_engine_train_host(dnnb200_engine* e, const float* h_x, const int64_t* h_y, int B, int nsteps,
                              float* h_losses, void* stream) {
  DNN_ARG_CHECK(e && h_x && h_y, "engine_train_host: null argument");
  DNN_ARG_CHECK(B >= 1 && B <= e->max_batch, "engine_train_host: batch %d exceeds max_batch %d", B, e->max_batch);
  DNN_ARG_CHECK(nsteps >= 1, "engine_train_host: nsteps must be positive");
  int rc = ensure_ring(e);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  constexpr int R = dnnb200_engine::kRing;
  const size_t xb = (size_t)B * kImgNumel * sizeof(float), yb = (size_t)B * sizeof(long long);
  // pinned host memory is addressable from the device (unified addressing): resolve the device view once
  float* d_losses = nullptr;
  if (h_losses) DNN_CUDA_CHECK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d_losses), h_losses, 0));
  // The clustered kernel can wait for its inputs itself: no stream events (train_host_flags above). The one-CTA kernel
  // of large batches, and calls recorded by a stream capture (the host must not wait inside one), keep the events.
  const int use_flags = (cluster_size_for(e, B) > 1 && !stream_capturing(s) && !(e->debug_flags & 256)) ? 1 : 0;
  if (e->last_host_path >= 0 && e->last_host_path != use_flags) {   // the two variants guard the ring differently
    DNN_CUDA_CHECK(cudaStreamSynchronize(s));
    DNN_CUDA_CHECK(cudaStreamSynchronize(e->copy_stream));
    for (bool& u : e->region_used) u = false;
    e->ring_uses = 0;
    e->flag_uses = 0;
  }
  e->last_host_path = use_flags;
  if (use_flags) return train_host_flags(e, h_x, (const long long*)h_y, B, nsteps, d_losses, s);
  // Steps are enqueued in groups of up to kGroup: the inputs of a group's steps travel in one H2D copy for the
  // images and one for the labels (every step's inputs are still copied, inside the call, just not one DMA
  // descriptor per step), and the cross-stream events (inputs ready / slots free) are recorded once per group, so
  // that inside a group nothing sits between the kernels in the stream and their programmatic dependent launches
  // overlap. The first group of a call holds 4 steps, the others 16: the GPU starts on step 0 after four steps' worth
  // of PCIe time (4 us per step of batch 16), every later copy hides under the kernels of the group before it, and a
  // 20-step call has one group boundary (each boundary costs the first step of the next group its chaining).
  constexpr int G = dnnb200_engine::kGroup;
  const bool may_chain = !stream_capturing(s);   // same rule as train_device: a captured counter value would be stale
  for (int t0 = 0; t0 < nsteps;) {
    const int cap = t0 == 0 ? 4 : G;   // a short first group (the GPU starts after 4 steps' worth of PCIe time), then full ones
    const int n = nsteps - t0 < cap ? nsteps - t0 : cap;
    const int gslot = (int)((e->ring_uses / G) % (R / G));        // group index in the ring
    if (e->ring_uses >= (unsigned long long)R) DNN_CUDA_CHECK(cudaStreamWaitEvent(e->copy_stream, e->slot_free[gslot], 0));
    // a group's region has a fixed place (so the slot_free event of that place guards it whatever batch sizes the
    // calls use); inside it the steps are packed at this call's batch size
    float* gx = e->ring_x + (size_t)gslot * G * e->max_batch * kImgNumel;
    long long* gy = e->ring_y + (size_t)gslot * G * e->max_batch;
    DNN_CUDA_CHECK(cudaMemcpyAsync(gx, h_x + (size_t)t0 * B * kImgNumel, (size_t)n * xb, cudaMemcpyHostToDevice, e->copy_stream));
    DNN_CUDA_CHECK(cudaMemcpyAsync(gy, h_y + (size_t)t0 * B, (size_t)n * yb, cudaMemcpyHostToDevice, e->copy_stream));
    DNN_CUDA_CHECK(cudaEventRecord(e->slot_ready[gslot], e->copy_stream));
    DNN_CUDA_CHECK(cudaStreamWaitEvent(s, e->slot_ready[gslot], 0));
    for (int k = 0; k < n; ++k) {
      const int t = t0 + k;
      int n_part1 = B;
      rc = fwd_bwd_any(e, gx + (size_t)k * B * kImgNumel, gy + (size_t)k * B, B, s, &n_part1, /*chained=*/may_chain && k > 0);
      if (rc) return rc;
      // The step's mean loss is written by the update kernel itself into h_losses[t] (device->host, 4 bytes).
      const int lslot = (int)(e->steps_launched & (unsigned long long)e->loss_ring_mask);
      rc = launch_update(e->d_ctx, 1, B, n_part1, lslot, e->debug_flags | peers_flag(e), d_losses ? d_losses + t : nullptr, s);
      if (rc) return rc;
      e->updates_launched += 1u;
      e->launches += 2;
      e->steps_launched += 1;
    }
    DNN_CUDA_CHECK(cudaEventRecord(e->slot_free[gslot], s));
    e->ring_uses += G;   // a short group still consumes a whole group of slots
    t0 += n;
  }
  return kOk;
}

int dnnb200_engine_eval(dnnb200_engine* e, const float* x, const int64_t* y, int N, int bs, float* logits,
                        double* h_out2, void* stream) {
  DNN_ARG_CHECK(e && x && y && h_out2, "engine_eval: null argument");
  DNN_ARG_CHECK(N >= 1 && bs >= 1, "engine_eval: N and bs must be positive");
  cudaStream_t s = (cudaStream_t)stream;
  if (N > e->eval_cap) {
    DNN_CUDA_CHECK(cudaStreamSynchronize(s));
    cudaFree(e->eval_loss); cudaFree(e->eval_correct);
    e->eval_loss = nullptr; e->eval_correct = nullptr; e->eval_cap = 0;
    DNN_CUDA_CHECK(cudaMalloc(&e->eval_loss, (size_t)N * sizeof(float)));
    DNN_CUDA_CHECK(cudaMalloc(&e->eval_correct, (size_t)N * sizeof(int)));
    e->eval_cap = N;
  }
  int rc = launch_fwd_bwd(false, e->params, x, (const long long*)y, nullptr, nullptr, nullptr, e->eval_loss,
                          e->eval_correct, logits, N, e->num_sms * 2, s);
  if (rc) return rc;
  rc = launch_eval_reduce(e->eval_loss, e->eval_correct, N, bs, e->d_eval_out, s);
  if (rc) return rc;
  e->launches += 2;
  DNN_CUDA_CHECK(cudaMemcpyAsync(h_out2, e->d_eval_out, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
  DNN_CUDA_CHECK(cudaStreamSynchronize(s));
  return kOk;
}

int dnnb200_engine_allreduce(dnnb200_engine* e, const float* in, float* out, int n, int mode, int two_shot,
                             void* stream) {
  DNN_ARG_CHECK(e && in, "engine_allreduce: null argument");
  DNN_ARG_CHECK(two_shot != 2 || (e->h_ctx.nvls_mc != nullptr && e->h_ctx.nvls_local != nullptr),
                "engine_allreduce: the NVLS variant needs dnnb200_engine_attach_nvls first");
  DNN_ARG_CHECK(two_shot != 3 || e->h_ctx.world <= 8, "engine_allreduce: the 128-byte-line variant handles at most 8 ranks");
  const float* ins[1] = {in};
  float* outs[1] = {out};
  int rc = launch_allreduce(e->d_ctx, 1, ins, outs, n, mode, two_shot, (cudaStream_t)stream);
  if (rc) return rc;
  e->launches += 1;
  return kOk;
}

int dnnb200_engine_allreduce_timed(dnnb200_engine* e, const float* in, float* out, int n, int two_shot, int iters,
                                   double* h_us_per_call, void* stream) {
  DNN_ARG_CHECK(e && in && out && h_us_per_call && iters >= 1, "engine_allreduce_timed: bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  // Launched from here, not from an interpreter: 2-3 us of host work per launch against >= 8 us per kernel, so the
  // stream never runs dry and the events bracket device time. The first calls align the ranks (the collective is
  // its own barrier) and are not timed.
  for (int i = 0; i < 8; ++i) {
    const int rc = dnnb200_engine_allreduce(e, in, out, n, 0, two_shot, stream);
    if (rc) return rc;
  }
  cudaEvent_t e0, e1;
  DNN_CUDA_CHECK(cudaEventCreate(&e0));
  DNN_CUDA_CHECK(cudaEventCreate(&e1));
  int rc = kOk;
  cudaEventRecord(e0, s);
  for (int i = 0; i < iters && rc == kOk; ++i) rc = dnnb200_engine_allreduce(e, in, out, n, 0, two_shot, stream);
  cudaEventRecord(e1, s);
  const cudaError_t se = cudaEventSynchronize(e1);
  float ms = 0.f;
  if (se == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (rc) return rc;
  DNN_CUDA_CHECK(se);
  *h_us_per_call = (double)ms * 1e3 / iters;
  return kOk;
}

int dnnb200_engine_leave(dnnb200_engine* e, void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  if (!e->group || e->group->world == 1) return kOk;
  leave_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(e->d_ctx, 0);
  DNN_CUDA_CHECK(cudaGetLastError());
  e->launches += 1;
  return kOk;
}

int dnnb200_engine_rejoin(dnnb200_engine* e, unsigned long long step, void* stream) {
  DNN_ARG_CHECK(e, "null engine");
  cudaStream_t s = (cudaStream_t)stream;
  ctl_set_step_kernel<<<1, 1, 0, s>>>(e->ctl, step);
  DNN_CUDA_CHECK(cudaGetLastError());
  if (e->group && e->group->world > 1) {
    leave_kernel<<<1, 32, 0, s>>>(e->d_ctx, 1);
    DNN_CUDA_CHECK(cudaGetLastError());
  }
  e->launches += 2;
  return kOk;
}

// ---- ranks emulated on one device -----------------------------------------------
static int gather_ctxs(dnnb200_engine** engines, int n, RankCtx** d_out) {
  std::vector<RankCtx> h((size_t)n);
  for (int r = 0; r < n; ++r) h[(size_t)r] = engines[r]->h_ctx;
  RankCtx* d = nullptr;
  DNN_CUDA_CHECK(cudaMalloc(&d, sizeof(RankCtx) * (size_t)n));
  DNN_CUDA_CHECK(cudaMemcpy(d, h.data(), sizeof(RankCtx) * (size_t)n, cudaMemcpyHostToDevice));
  *d_out = d;
  return kOk;
}

int dnnb200_vr_step(dnnb200_engine** engines, int n, const float* const* xs, const int64_t* const* ys, int B,
                    void* stream) {
  DNN_ARG_CHECK(engines && xs && ys && n >= 1 && n <= kMaxWorld, "vr_step: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  int n_part1 = B;
  for (int r = 0; r < n; ++r) {
    dnnb200_engine* e = engines[r];
    if (!xs[r]) continue;  // this virtual rank sits the step out (it has left)
    DNN_ARG_CHECK(B >= 1 && B <= e->max_batch, "vr_step: batch exceeds max_batch");
    int np1 = B;
    int rc = fwd_bwd_any(e, xs[r], (const long long*)ys[r], B, s, &np1);
    if (rc) return rc;
    n_part1 = np1;
    e->launches += 1;
  }
  // compact the participating ranks into one cooperative launch
  std::vector<dnnb200_engine*> live;
  for (int r = 0; r < n; ++r) if (xs[r]) live.push_back(engines[r]);
  if (live.empty()) return kOk;
  RankCtx* d = nullptr;
  int rc = gather_ctxs(live.data(), (int)live.size(), &d);
  if (rc) return rc;
  const int lslot = (int)(live[0]->steps_launched & (unsigned long long)live[0]->loss_ring_mask);
  // (a single live rank of a larger group still runs the exchange code: it waits for, and gives up on, the others)
  rc = launch_update(d, (int)live.size(), B, n_part1, lslot, live[0]->h_ctx.world > 1 ? (int)kUpdFlagHasGroup : 0, nullptr, s);
  cudaStreamSynchronize(s);
  cudaFree(d);
  if (rc) return rc;
  for (auto* e : live) { e->launches += 1; e->steps_launched += 1; e->updates_launched += 1u; }
  return kOk;
}

int dnnb200_vr_allreduce(dnnb200_engine** engines, int n, const float* const* in, float* const* out, int count,
                         int mode, int two_shot, void* stream) {
  DNN_ARG_CHECK(engines && in && out && n >= 1 && n <= kMaxWorld, "vr_allreduce: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  std::vector<dnnb200_engine*> live;
  std::vector<const float*> lin;
  std::vector<float*> lout;
  for (int r = 0; r < n; ++r)
    if (in[r]) { live.push_back(engines[r]); lin.push_back(in[r]); lout.push_back(out[r]); }
  if (live.empty()) return kOk;
  RankCtx* d = nullptr;
  int rc = gather_ctxs(live.data(), (int)live.size(), &d);
  if (rc) return rc;
  rc = launch_allreduce(d, (int)live.size(), lin.data(), lout.data(), count, mode, two_shot, s);
  cudaStreamSynchronize(s);
  cudaFree(d);
  if (rc) return rc;
  for (auto* e : live) e->launches += 1;
  return kOk;
}

}  // extern "C"
