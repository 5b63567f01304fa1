/* C ABI of the native runtime (libdnnb200.so), loaded from Python with ctypes
 * (distributed_neural_network_b200/ops/native.py) and usable from C/C++.
 *
 * Conventions: every function returns 0 on success or a DNNB200_ERR_* code;
 * dnnb200_last_error() gives the message of the last failure on the calling
 * thread. Pointers are device pointers unless the name starts with h_ (host,
 * pinned where noted). All work is enqueued on the given cudaStream_t (passed as
 * void*; NULL = legacy default stream) and nothing synchronises unless stated.
 * Tensors are fp32, contiguous, NCHW; labels are int64.
 */
#ifndef DNNB200_H_
#define DNNB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DNNB200_OK 0
#define DNNB200_ERR_CUDA 1
#define DNNB200_ERR_ARG 2
#define DNNB200_ERR_STATE 3
#define DNNB200_ERR_TIMEOUT 4

#define DNNB200_FLAT_NUMEL 62144 /* floats in the flat parameter vector (62006 real + alignment gaps) */
#define DNNB200_IPC_HANDLE_BYTES 64

const char* dnnb200_last_error(void);
int dnnb200_version(void);
/* layout constants, so Python can assert it agrees with csrc/layout.h */
int dnnb200_layout(int* offsets10, int* flat_numel, int* slice, int* num_slices);

/* ---- layer kernels (reference: models/model.py:20-27) ---------------- */
/* y = maxpool2x2(relu(conv5x5_valid(x, w) + b)); idx = 2-bit argmax per pooled pixel (K1-K6) */
int dnnb200_conv_relu_pool_fwd(const float* x, const float* w, const float* b, float* y, uint8_t* idx,
                               int N, int Ci, int H, int W, int Co, void* stream);
/* backward of the above; dx may be NULL (first layer). ReLU'/pool' are folded in (K14-K17) */
int dnnb200_conv_relu_pool_bwd(const float* x, const float* w, const float* y, const uint8_t* idx,
                               const float* dy, float* dx, float* dw, float* db,
                               int N, int Ci, int H, int W, int Co, void* stream);
/* as above, with the data gradient of the 6x14x14 -> 16 layer computed on the tensor cores (tcgen05 implicit
 * GEMM over the zero-padded gradient planes with the flipped filter; TF32). Other shapes fall back to FFMA. */
int dnnb200_conv_relu_pool_bwd_tc(const float* x, const float* w, const float* y, const uint8_t* idx,
                                  const float* dy, float* dx, float* dw, float* db,
                                  int N, int Ci, int H, int W, int Co, void* stream);
/* weight and bias gradient of conv5x5 + ReLU + maxpool2x2 on the tensor cores (tcgen05.mma kind::tf32): the reduction
 * runs over output pixels, so the transposed im2col operand is gathered into shared memory block by block and the
 * accumulators stay in tensor memory over all images of a CTA (csrc/tc_conv_wgrad.cu). dw, db are overwritten.
 * Only the CNN's two conv shapes. (reference: convolution_backward of data_parallelism_train.py:198) */
int dnnb200_tc_conv_wgrad(const float* x, const float* y, const uint8_t* idx, const float* dy, float* dw, float* db, int N,
                          int Ci, int H, int W, int Co, void* stream);
/* the same forward on the tensor cores: implicit GEMM, tcgen05.mma kind::tf32 with TMEM accumulators,
 * the shared-memory descriptor performs the im2col (csrc/tc_conv.cu). TF32 operands, fp32 accumulate.
 * Instantiated for the two conv layers of the CNN: (Ci,H,W,Co) = (3,32,32,6) and (6,14,14,16). */
int dnnb200_tc_conv_relu_pool_fwd(const float* x, const float* w, const float* b, float* y, uint8_t* idx,
                                  int N, int Ci, int H, int W, int Co, void* stream);
/* descriptor laboratory (csrc/tc_lab.cu): raw shared-memory images + raw descriptors (start address 0) in,
 * D[128, ncols] out; nmma MMAs with the start addresses advancing by a_step / b_step bytes */
int dnnb200_tc_lab(const float* a_img, int a_floats, const float* b_img, int b_floats, unsigned long long adesc,
                   unsigned long long bdesc, unsigned int idesc, int nmma, int a_step, int b_step, float* D,
                   int ncols, void* stream);
/* timing twin: nmma MMAs over zeroed operands of a_bytes / b_bytes, accumulating round-robin into n_acc tiles of
 * acc_cols tensor-memory columns; d_out2 (device) = {cycles spent issuing, cycles until the commit barrier fired} */
int dnnb200_tc_lab_time(int a_bytes, int b_bytes, unsigned long long adesc, unsigned long long bdesc, unsigned int idesc,
                        int nmma, int a_step, int b_step, int n_acc, int acc_cols, long long* d_out2, void* stream);
/* D[128,16] = A[128,K] B[16,K]^T through the same descriptors (unit test of the encodings) */
int dnnb200_tc_gemm_probe(const float* A, const float* B, float* D, int K, void* stream);
/* y[N,M] = act(x[N,K] w[M,K]^T + b[M]) (K8-K10) */
int dnnb200_linear_fwd(const float* x, const float* w, const float* b, float* y, int N, int M, int K,
                       int relu, void* stream);
/* the same forward on the tensor cores: tcgen05.mma kind::tf32, TMEM accumulators, both operands loaded by
 * TMA through 2-D tensor maps with 128-byte swizzle (csrc/tc_linear.cu). M <= 128, K % 4 == 0. */
int dnnb200_tc_linear_fwd(const float* x, const float* w, const float* b, float* y, int N, int M, int K,
                          int relu, void* stream);
/* backward of the same layer on the tensor cores (csrc/tc_linear_bwd.cu): dX and dW are tcgen05 GEMMs whose
 * operands are TMA boxes of the row-major matrices read through MN-major shared-memory descriptors (no
 * transposes). M % 4 == 0, M <= 128, K % 4 == 0. dy_scratch [N, M] is needed when relu != 0 (masked dy);
 * dx may be NULL. dW is accumulated with fp32 atomics when the batch is split over CTAs. */
int dnnb200_tc_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dy_scratch,
                          float* dx, float* dw, float* db, int N, int M, int K, int relu, void* stream);
/* dx may be NULL; y is the layer OUTPUT (its sign gives ReLU') (K13, K14) */
int dnnb200_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dx,
                       float* dw, float* db, int N, int M, int K, int relu, void* stream);
/* mean cross-entropy of logits[N,C] and d(loss)/d(logits) in one pass (K11, K12;
 * reference: data_parallelism_train.py:29,197) */
int dnnb200_softmax_xent(const float* logits, const int64_t* target, float* loss, float* dlogits,
                         int N, int C, void* stream);
/* buf = mu*buf + g*scale (buf = g*scale if first_step); p -= lr*buf (K19;
 * reference: data_parallelism_train.py:187,199) */
int dnnb200_sgd_momentum(float* p, const float* g, float* buf, int n, float lr, float mu,
                         int first_step, float scale, void* stream);

/* ---- peer group: symmetric receive/flag buffers over CUDA IPC (reference: the mpi4py
 *      send/recv star of data_parallelism_train.py:118,135,210,227 -- SURVEY 2.5) ---------- */
typedef struct dnnb200_group dnnb200_group;
/* allocates this rank's receive + flag area on the current device and exports its IPC handle */
int dnnb200_group_create(int rank, int world, dnnb200_group** out, void* h_handle64);
/* h_all_handles: world x 64 bytes, gathered by the bootstrap (mpi4py / torch.distributed / file) */
int dnnb200_group_connect(dnnb200_group* g, const void* h_all_handles);
/* single-process variant: the n groups live in this process (tests; ranks emulated on one GPU) */
int dnnb200_group_connect_local(dnnb200_group** groups, int n);
int dnnb200_group_destroy(dnnb200_group* g);

/* ---- fused step engine ------------------------------------------------------------------- */
typedef struct dnnb200_engine dnnb200_engine;
typedef struct dnnb200_ctl {
  unsigned long long step;
  unsigned long long timeout_ns;
  float lr, mu;
  float loss_sum;
  unsigned int loss_steps;
  unsigned int ctas_done;
  unsigned int status;
  unsigned int active_mask;
  unsigned int two_shot;
  unsigned int updates_done;
  unsigned int divisor_override;
} dnnb200_ctl;

/* params/momentum (and optional grads, may be NULL) are caller-owned flat buffers of
 * DNNB200_FLAT_NUMEL floats, e.g. Network._flat: the engine updates them in place */
int dnnb200_engine_create(float* params, float* momentum, float* grads, int max_batch, float lr, float mu,
                          dnnb200_engine** out);
int dnnb200_engine_destroy(dnnb200_engine* e);
int dnnb200_engine_attach_group(dnnb200_engine* e, dnnb200_group* g);
int dnnb200_engine_set_hparams(dnnb200_engine* e, float lr, float mu, void* stream);
int dnnb200_engine_set_membership(dnnb200_engine* e, unsigned int active_mask, unsigned long long timeout_ns,
                                  void* stream);
/* variant of the exchange fused into the update kernel: 0 = one-shot (every rank pushes to every rank; tolerates
 * ranks leaving mid-epoch), 1 = two-shot (slice owners reduce and broadcast; static membership only),
 * 2 = NVLS (multimem.st flags + multimem.ld_reduce in the NVSwitch; static membership, needs attach_nvls),
 * 3 = one-shot with 128-byte lines (31 values + 1 flag word per warp-wide store instead of 2 + 2 per 16-byte
 * line: half the NVLink bytes, same single hop, same membership rules as 0).
 * A peer that stays silent longer than timeout_ns: 0 and 3 leave it out and divide by the ranks present; 1 the
 * owner leaves it out and rescales, a rank with a silent owner keeps its own gradient for that slice; 2 the
 * slice falls back to this rank's own gradient. Always flagged in dnnb200_ctl::status. */
int dnnb200_engine_set_exchange_variant(dnnb200_engine* e, int two_shot, void* stream);
/* 0 = divide the exchanged sum by the number of ranks present (default); n = divide by n */
int dnnb200_engine_set_divisor(dnnb200_engine* e, unsigned int divisor, void* stream);
/* zero the momentum buffer (the reference re-creates its optimizer every epoch:
 * data_parallelism_train.py:187) */
int dnnb200_engine_zero_momentum(dnnb200_engine* e, void* stream);
int dnnb200_engine_reset_loss(dnnb200_engine* e, void* stream);
/* synchronises the stream and copies the control block to the host */
int dnnb200_engine_read_ctl(dnnb200_engine* e, dnnb200_ctl* h_out, void* stream);
/* number of kernels this engine has launched so far */
unsigned long long dnnb200_engine_launches(const dnnb200_engine* e);
/* thread-block cluster size of the forward/backward kernel: each image is spread over `cs` CTAs
 * (distributed shared memory between them). 0 = choose per batch size so the batch is one wave. */
int dnnb200_engine_set_cluster_size(dnnb200_engine* e, int cs);
int dnnb200_engine_cluster_info(const dnnb200_engine* e, int B, int* chosen, int* max_clusters9);
/* arithmetic of the convolution stages of the clustered forward/backward kernel (conv1 and conv2 forward, conv2
 * data gradient): 1 = implicit GEMMs on the tensor cores (tcgen05.mma kind::tf32: operands rounded to TF32, fp32
 * accumulation in tensor memory); 0 (default of a new engine) = fp32 FFMA throughout, the path that matches PyTorch
 * fp32 to 2e-5 and that measures faster at batch 16 (profiles/step_tc_ab_v1.md).
 * Everything else (FC layers, weight gradients, softmax, SGD, master weights) is fp32 either way. Batches that
 * run the one-CTA-per-image kernel (cluster size 1) are always fp32. */
/* 2 = fp32 FFMA everywhere except the conv2 data gradient, which runs on the tensor cores (TF32 operands): the forward
 * pass, the loss and every gradient except conv1.weight / conv1.bias are bit-for-bit those of mode 0. */
int dnnb200_engine_set_precision(dnnb200_engine* e, int tf32);
/* development aid: d_trace[cta*16 + k] receives the SM clock at stage boundary k of the clustered
 * forward/backward kernel (NULL switches it off) */
int dnnb200_engine_set_trace(dnnb200_engine* e, long long* d_trace);
/* same for the update kernel: d_trace[slice*8 + k] = globaltimer (ns) at k = 0 entry, 1 after the dependency wait,
 * 2 after the local gradient, 3 after the exchange, 4 after the SGD stores, 5 after the fence, 6 after the
 * last-CTA counter, and k = 7 the number of ranks that entered the slice's sum; [122*8] words, or [1024*122*8] with debug flag 4 (ring over the last 1024 steps) */
int dnnb200_engine_set_update_trace(dnnb200_engine* e, long long* d_trace);
/* development aid / A-B switches, OR-ed: 2 = the update kernel does not trigger the early launch of the next
 * kernel; 4 = the update trace is a ring over the last 1024 steps; 8 = chained steps wait for the completion of
 * the previous update GRID instead of its update counter; 32 = with peers, use the unconstrained update kernel
 * instead of the one that fits three CTAs per SM; 128 = chained steps of the tensor-core
 * instantiations pack their B operands in the kernel instead of fetching them ready-made; 256 = the host-fed path
 * uses the event variant even with the clustered kernel */
int dnnb200_engine_set_debug_flags(dnnb200_engine* e, int flags);

/* one optimizer step on a device-resident batch: fwd/bwd kernel + update kernel
 * (reference: the loop body data_parallelism_train.py:195-203) */
int dnnb200_engine_step(dnnb200_engine* e, const float* x, const int64_t* y, int B, void* stream);
/* nsteps steps over device-resident batches x[nsteps][B][3][32][32], y[nsteps][B]: one native loop,
 * no per-step Python (the whole dataset fits in HBM many times over) */
int dnnb200_engine_train_device(dnnb200_engine* e, const float* x, const int64_t* y, int B, int nsteps, void* stream);
/* nsteps steps fed from PINNED host memory: every step's inputs are copied host->device inside the call (copy stream,
 * 32-slot device ring) and run through the two kernels; the update kernel writes the step's mean loss straight into
 * h_losses[t] (pinned host memory, 4 bytes device->host per step). With the clustered kernel (batches up to a few
 * dozen images) there is no stream event anywhere: the copy stream writes a sequence number behind every two steps'
 * inputs, the kernel that reads them waits for it, and a ring region is overwritten once the update kernel of its
 * last reader has reported in through a pinned progress word -- so the call runs at most 32 steps ahead of the GPU
 * and returns when everything is enqueued. Larger batches (one-CTA kernel) and calls under stream capture use an
 * event pair per group of 16 steps instead. The caller synchronises `stream` before reading h_losses or reusing the
 * host buffers. */
int dnnb200_engine_train_host(dnnb200_engine* e, const float* h_x, const int64_t* h_y, int B, int nsteps,
                              float* h_losses, void* stream);
/* forward only over N device-resident images; h_out2 = {mean over batches of the batch-mean
 * loss, number of correct predictions}; logits may be NULL. Synchronises.
 * (reference: data_parallelism_train.py:157-183) */
int dnnb200_engine_eval(dnnb200_engine* e, const float* x, const int64_t* y, int N, int bs, float* logits,
                        double* h_out2, void* stream);
/* forward + backward only (no update): fills the per-image records; grads of the local batch
 * are written to `grads_out` (flat, summed over the batch, already scaled by 1/B). For tests. */
int dnnb200_engine_grad_only(dnnb200_engine* e, const float* x, const int64_t* y, int B, float* grads_out,
                             float* h_loss, void* stream);

/* (No counterpart in the reference: it has no native code and no notion of a memory budget.)
 * Batch gradient WITHOUT any allocation, for callers that may not call cudaMalloc: processes whose
 * device memory is fully owned by another allocator (a server that reserved the GPU up front, a
 * framework's caching allocator at its cap), and latency-critical loops or foreign-runtime callbacks
 * where cudaMalloc/cudaFree are ruled out because they take driver locks and can synchronise the device.
 * Same math as dnnb200_engine_grad_only, but the per-image scratch lives in a fixed array inside the
 * library image (DNNB200_NOALLOC_SCRATCH_BYTES on every GPU that loads the library, until exit).
 *   params [DNNB200_FLAT_NUMEL], x [B,3,32,32], y [B] int64: inputs;  grads_out [DNNB200_FLAT_NUMEL]:
 *   d(mean loss)/d(params), zeros in the alignment gaps;  loss_out [1]: mean loss (may be NULL).
 * Limits and rules: 1 <= B <= DNNB200_NOALLOC_MAX_BATCH, otherwise DNNB200_ERR_ARG and nothing is
 * launched. All calls on one device share the array, so calls that can overlap on the GPU race:
 * keep them on one stream or order them yourself. Nothing in the array needs clearing between calls.
 * The FIRST call on each device resolves the array's device address (cudaGetSymbolAddress); make that
 * call at start-up, before any stream capture begins. Calls cudaMalloc/cudaFree/torch never. */
#define DNNB200_NOALLOC_MAX_BATCH 64
#define DNNB200_NOALLOC_SCRATCH_BYTES 967168
int dnnb200_batch_grad_noalloc(const float* params, const float* x, const int64_t* y, int B, float* grads_out,
                               float* loss_out, void* stream);

/* peer-memory all-reduce of a flat vector (n % 4 == 0, n <= FLAT_NUMEL).
 * mode 0: out = mean(in); mode 1: fused SGD-momentum on the engine's params with g = mean(in)
 * (out may be NULL). variant (argument `two_shot`): 0 one-shot push, 1 two-shot (owner-reduce + broadcast),
 * 2 NVLS (in-switch reduction through a multicast mapping; needs dnnb200_engine_attach_nvls),
 * 3 one-shot push with 128-byte lines. */
int dnnb200_engine_allreduce(dnnb200_engine* e, const float* in, float* out, int n, int mode, int two_shot,
                             void* stream);

/* `iters` back-to-back mean all-reduces (after 8 untimed ones) between two events on `stream`, enqueued from C so
 * that the stream never runs dry: device time per call in microseconds. Collective; synchronises. */
int dnnb200_engine_allreduce_timed(dnnb200_engine* e, const float* in, float* out, int n, int two_shot, int iters,
                                   double* h_us_per_call, void* stream);

/* NVLS: give the engine a symmetric buffer of DNNB200_NVLS_FLOATS floats that every rank allocated and mapped
 * into one NVLink multicast object (dnnb200_nvls_open / dnnb200_nvls_map below): `local_buf` is this rank's
 * copy, `multicast_buf` the multicast address. The buffer must be zero-initialised. */
#define DNNB200_NVLS_FLOATS (2 * DNNB200_FLAT_NUMEL + 2 * 16 * 128)
int dnnb200_engine_attach_nvls(dnnb200_engine* e, float* local_buf, float* multicast_buf);

/* ---- NVLink multicast group (csrc/nvls_group.cu): the symmetric buffer for dnnb200_engine_attach_nvls, made with the
 * CUDA virtual-memory and multicast driver calls; rank 0 hands the multicast object to the other processes as a file
 * descriptor over an abstract Unix socket named after `name` (SCM_RIGHTS), which must be unique to the job.
 *   every rank: dnnb200_nvls_open(...)           -> creates / imports the object, adds this rank's device
 *   BARRIER of the caller's control plane        (memory may only be bound once every device is in)
 *   every rank: dnnb200_nvls_map(...)            -> allocates, binds, maps; `local` is zero-filled
 *   BARRIER before the first use                 (a multicast store to a member that has not bound yet is lost)
 * All ranks must pass the same `bytes` (rounded up to the granularity the driver recommends; `*bytes_out`).
 * Errors: DNNB200 code 3 when the driver or device has no multicast support, 4 when a rank did not show up within
 * `timeout_ms`. dnnb200_nvls_close unmaps and releases; detach the engine (or destroy it) first. */
typedef struct dnnb200_nvls dnnb200_nvls;
int dnnb200_nvls_supported(int device);
int dnnb200_nvls_open(dnnb200_nvls** out, int rank, int world, int device, size_t bytes, const char* name,
                      int timeout_ms);
int dnnb200_nvls_map(dnnb200_nvls* g, float** local, float** multicast, size_t* bytes_out);
int dnnb200_nvls_close(dnnb200_nvls* g);
/* the rendezvous underneath, usable on its own: rank 0 passes (*word, *fd) to every other rank of the same host, which
 * receive a duplicate of the descriptor in *fd (theirs to close) and the word in *word. */
int dnnb200_share_fd(int rank, int world, const char* name, unsigned long long* word, int* fd, int timeout_ms);

/* latency probe: `iters` round trips of one 16-byte LL line with rank `peer` (which must make the matching call);
 * returns the one-way store->visible latency over peer memory in microseconds. Synchronises. */
int dnnb200_engine_pingpong(dnnb200_engine* e, int peer, int iters, double* h_one_way_us, void* stream);

/* probe of the 128-byte-line assumption of exchange variant 3 with rank `peer` (which must make the matching
 * call): n_ctas x 8 warps play ping-pong for `iters` rounds with one warp-wide 128-byte store per line and check
 * every line that arrives. h_out3 = {words that did not belong to the flag they arrived with (must be 0), ns for
 * all rounds, 1 if a wait was given up after 2 s}. Synchronises; allocates 24 bytes for the duration of the call. */
int dnnb200_engine_ll128_probe(dnnb200_engine* e, int peer, int iters, int n_ctas, unsigned long long* h_out3,
                               void* stream);

/* membership epoch: announce that this rank stops taking part after the steps
 * already enqueued; peers divide by the survivors from then on */
int dnnb200_engine_leave(dnnb200_engine* e, void* stream);
/* come back: clear the marker on every peer and adopt the group's sequence number */
int dnnb200_engine_rejoin(dnnb200_engine* e, unsigned long long step, void* stream);

/* ranks emulated on ONE device (tests, and boxes with fewer GPUs than ranks): one cooperative
 * launch runs the update (or all-reduce) of all n engines so that they can wait on each other */
int dnnb200_vr_step(dnnb200_engine** engines, int n, const float* const* xs, const int64_t* const* ys, int B,
                    void* stream);
int dnnb200_vr_allreduce(dnnb200_engine** engines, int n, const float* const* in, float* const* out, int count,
                         int mode, int two_shot, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DNNB200_H_ */
