// Stand-alone peer-memory all-reduce of a flat fp32 vector, optionally fused with
// the 1/W scale and the SGD-momentum update.
//
// Used (a) by the layer-at-a-time / autograd path, where gradients already sit in
// Network._flat_grad (one launch for the whole flat vector, after backward); (b) in the
// reference-parity "epoch-average" and "replication" modes, where the vector is
// the PARAMETER buffer and the op is a plain mean (reference:
// data_parallelism_train.py:238-240, model_replication_train.py:134-136 -- R1);
// (c) by the latency benchmark that picks the variant per message size.
//
// Variants
//   one-shot : every rank pushes its slice to every peer, then sums locally.
//              (W-1) x bytes out per rank, ONE flag round trip. Wins while the
//              message is latency-bound (248 KB gradient of this CNN).
//   two-shot : slice s is owned by the (s mod n)-th configured rank; others push
//              only to the owner, the owner reduces and pushes the result back.
//              2(W-1)/W x bytes, TWO flag round trips. Wins for large vectors.
//              Ownership needs a static membership, so ranks must not leave
//              between the host's membership updates while this variant is used.
//
// The fused step path (fused_step.cu: update_kernel) embeds the same variants
// (exchange.cuh) directly behind its weight-gradient reduction.
#include <cuda_runtime.h>

#include "common.cuh"
#include "exchange.cuh"
#include "layout.h"
#include "step_types.h"

namespace dnnb200 {

constexpr int kArThreads = kSlice / 4;  // 128 threads, one float4 each

enum ArMode : int {
  kArMean = 0,  // out = mean(in)
  kArSgd = 1,   // g = mean(in); momentum = mu*momentum + g; params -= lr*momentum   (out optional: g)
};

struct ArParams {
  const float* in[kMaxWorld];  // per virtual rank
  float* out[kMaxWorld];
  int n;         // elements (multiple of 4, <= kFlatNumel)
  int mode;      // ArMode
  int two_shot;  // variant: 0 one-shot, 1 two-shot, 2 NVLS (multimem.ld_reduce in the NVSwitch), 3 one-shot with 128-byte lines
};

__device__ __forceinline__ void finish(const RankCtx& cx, const ArParams& prm, int vr, int i, const float (&v)[4],
                                       int n) {
  const unsigned int dov = cx.ctl->divisor_override;
  const float inv = 1.f / (float)((dov && n > 0) ? (int)dov : (n > 0 ? n : 1));
  const float4 g = make_float4(v[0] * inv, v[1] * inv, v[2] * inv, v[3] * inv);
  if (prm.mode == kArSgd) {
    const float lr = cx.ctl->lr, mu = cx.ctl->mu;
    float4 m = *reinterpret_cast<float4*>(cx.momentum + i);
    float4 p = *reinterpret_cast<float4*>(cx.params + i);
    m.x = mu * m.x + g.x; m.y = mu * m.y + g.y; m.z = mu * m.z + g.z; m.w = mu * m.w + g.w;
    p.x -= lr * m.x; p.y -= lr * m.y; p.z -= lr * m.z; p.w -= lr * m.w;
    *reinterpret_cast<float4*>(cx.momentum + i) = m;
    *reinterpret_cast<float4*>(cx.params + i) = p;
  }
  if (prm.out[vr]) *reinterpret_cast<float4*>(prm.out[vr] + i) = g;
}

template <bool kEmulated>  // see update_kernel: emulated ranks need 8 x 122 co-resident CTAs
__global__ void __launch_bounds__(kArThreads, kEmulated ? 8 : 1)
allreduce_kernel(const RankCtx* __restrict__ ctxs, const ArParams prm) {
  __shared__ ExchangeShared xsh;
  const int vr = blockIdx.y;
  const RankCtx& cx = ctxs[vr];
  StepCtl* ctl = cx.ctl;
  const int slice = blockIdx.x, tid = threadIdx.x;
  const int i = slice * kSlice + 4 * tid;
  const bool in_range = i < prm.n;
  const unsigned long long step = *reinterpret_cast<volatile unsigned long long*>(&ctl->step);
  const unsigned int seq = (unsigned int)(step + 1ull);
  const int world = cx.world;

  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (in_range) {
    const float4 t = *reinterpret_cast<const float4*>(prm.in[vr] + i);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  int n = 1;
  const ExchangeCfg cfg = exchange_cfg(cx, ctl, seq);
  if (world > 1 && !prm.two_shot) {
    exchange_slice<4>(cx, cfg, &xsh, slice, seq, i, in_range, v, &n);
    if (in_range) finish(cx, prm, vr, i, v, n);
  } else if (world > 1 && prm.two_shot == 3) {
    const int left = prm.n - slice * kSlice;
    exchange_slice_ll128<4>(cx, cfg, &xsh, slice, seq, left < kSlice ? left : kSlice, v, &n);  // one-shot, 128-byte lines
    if (in_range) finish(cx, prm, vr, i, v, n);
  } else if (world > 1 && prm.two_shot == 2) {
    exchange_slice_nvls<4>(cx, cfg, &xsh, slice, seq, i, in_range, v, &n);   // in-switch reduction (exchange.cuh)
    if (in_range) finish(cx, prm, vr, i, v, n);
  } else if (world > 1) {
    exchange_slice_two_shot<4>(cx, cfg, &xsh, slice, seq, i, in_range, v, &n);
    if (in_range) finish(cx, prm, vr, i, v, n);
  } else {
    if (in_range) finish(cx, prm, vr, i, v, 1);
  }

  __syncthreads();
  if (tid == 0) {
    __threadfence();
    const unsigned int done = atomicAdd(&ctl->ctas_done, 1u);
    if (done == gridDim.x - 1) {
      ctl->ctas_done = 0u;
      __threadfence();
      *reinterpret_cast<volatile unsigned long long*>(&ctl->step) = step + 1ull;
    }
  }
}

// ---- latency probe: ping-pong of one 16-byte LL line between two ranks over peer memory -------------
// (mailbox = word 124.. of the second flag bank, which no collective uses)
__global__ void pingpong_kernel(const RankCtx* __restrict__ cxp, int peer, int iters, unsigned int base_seq,
                                long long* __restrict__ out_ns) {
  const RankCtx& cx = *cxp;
  if (threadIdx.x != 0) return;
  float* mine = reinterpret_cast<float*>(cx.flags[cx.rank] + (kMaxWorld + peer) * kFlagRow + 124);
  float* theirs = reinterpret_cast<float*>(cx.flags[peer] + (kMaxWorld + cx.rank) * kFlagRow + 124);
  const bool first = cx.rank < peer;
  const unsigned long long t0 = globaltimer_ns();
  float a, b;
  for (int i = 0; i < iters; ++i) {
    const unsigned int seq = base_seq + (unsigned int)i;
    if (first) {
      st_ll(theirs, 1.f, 2.f, seq);
      while (!ld_ll(mine, seq, &a, &b)) {}
    } else {
      while (!ld_ll(mine, seq, &a, &b)) {}
      st_ll(theirs, 1.f, 2.f, seq);
    }
  }
  *out_ns = (long long)(globaltimer_ns() - t0);
}

// ---- 128-byte line probe: is a warp-wide 128-byte store ever seen torn by the peer's warp-wide load? -------------
// Both ranks run this kernel at the same time. Every warp owns one line slot of the receive area (slice =
// blockIdx.x, line = warp) and plays ping-pong with the same warp of the peer: 31 words derived from
// (sequence, slice, line, lane) + the sequence number as flag word, written with ONE st per lane. The receiver
// spins until the flag word carries the sequence number and then checks that the 31 values returned by that
// very load belong to it too. A line delivered in pieces shows up as values of the previous round under the
// new flag. out[0] = mismatching words, out[1] = ns for all rounds, out[2] = 1 if a wait was given up (2 s).
__device__ __forceinline__ unsigned int line_pattern(unsigned int seq, int slice, int line, int lane) {
  return (seq * 2654435761u) ^ (unsigned int)(slice * 7919 + line * 131 + lane + 1);
}
__global__ void __launch_bounds__(256)
ll128_probe_kernel(const RankCtx* __restrict__ cxp, int peer, int iters, unsigned int base_seq,
                   unsigned long long* __restrict__ out) {
  const RankCtx& cx = *cxp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, slice = blockIdx.x;
  const bool first = cx.rank < peer;
  const unsigned long long t_start = globaltimer_ns();
  unsigned long long bad = 0ull;
  bool gave_up = false;
  for (int it = 0; it < iters && !gave_up; ++it) {
    const unsigned int seq = base_seq + (unsigned int)it, parity = seq & 1u;
    const size_t slot = (size_t)slice * kSlice128 + (size_t)warp * 32 + lane;
    float* theirs = cx.recv[peer] + (size_t)(parity * cx.world + cx.rank) * kLLStride + slot;
    const float* mine = cx.recv[cx.rank] + (size_t)(parity * cx.world + peer) * kLLStride + slot;
    const unsigned int word = lane == kLineData ? seq : line_pattern(seq, slice, warp, lane);
    if (first) st_line_word(theirs, word);
    unsigned int got;
    const unsigned long long t0 = globaltimer_ns();
    unsigned int spins = 0;
    for (;;) {
      got = ld_line_word(mine);
      if (__shfl_sync(0xffffffffu, got, kLineData) == seq) break;
      if ((++spins & 0xfffu) == 0u && __shfl_sync(0xffffffffu, globaltimer_ns() - t0 > 2000000000ull ? 1 : 0, 0)) {
        gave_up = true;
        break;
      }
    }
    if (!gave_up && lane < kLineData && got != line_pattern(seq, slice, warp, lane)) ++bad;
    if (!first) st_line_word(theirs, word);
  }
  if (bad) atomicAdd(out, bad);
  if (gave_up && lane == 0) out[2] = 1ull;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[1] = globaltimer_ns() - t_start;
}

int launch_ll128_probe(const RankCtx* d_ctx, int peer, int iters, int n_ctas, unsigned int base_seq,
                       unsigned long long* d_out3, cudaStream_t stream) {
  DNN_ARG_CHECK(n_ctas >= 1 && n_ctas <= kNumSlices, "ll128_probe: 1..%d CTAs", kNumSlices);
  DNN_CUDA_CHECK(cudaMemsetAsync(d_out3, 0, 3 * sizeof(unsigned long long), stream));
  ll128_probe_kernel<<<n_ctas, 256, 0, stream>>>(d_ctx, peer, iters, base_seq, d_out3);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_pingpong(const RankCtx* d_ctx, int peer, int iters, unsigned int base_seq, long long* d_out_ns,
                    cudaStream_t stream) {
  pingpong_kernel<<<1, 32, 0, stream>>>(d_ctx, peer, iters, base_seq, d_out_ns);
  DNN_CUDA_CHECK(cudaGetLastError());
  return kOk;
}

int launch_allreduce(const RankCtx* d_ctxs, int n_virtual_ranks, const float* const* in, float* const* out,
                     int n, int mode, int two_shot, cudaStream_t stream) {
  DNN_ARG_CHECK(n > 0 && n <= kFlatNumel && (n % 4) == 0, "allreduce: n must be a multiple of 4 and <= %d", kFlatNumel);
  DNN_ARG_CHECK(n_virtual_ranks >= 1 && n_virtual_ranks <= kMaxWorld, "allreduce: bad virtual rank count");
  ArParams prm;
  for (int r = 0; r < kMaxWorld; ++r) { prm.in[r] = nullptr; prm.out[r] = nullptr; }
  for (int r = 0; r < n_virtual_ranks; ++r) { prm.in[r] = in[r]; prm.out[r] = out[r]; }
  prm.n = n; prm.mode = mode; prm.two_shot = two_shot;
  dim3 grid((n + kSlice - 1) / kSlice, n_virtual_ranks);
  if (n_virtual_ranks > 1) {
    void* args[] = {(void*)&d_ctxs, (void*)&prm};
    DNN_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)allreduce_kernel<true>, grid, dim3(kArThreads), args, 0, stream));
  } else {
    allreduce_kernel<false><<<grid, kArThreads, 0, stream>>>(d_ctxs, prm);
    DNN_CUDA_CHECK(cudaGetLastError());
  }
  return kOk;
}

}  // namespace dnnb200
