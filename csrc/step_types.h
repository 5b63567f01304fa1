// Device-visible structs shared by the fused step kernels, the stand-alone
// all-reduce kernel and the host runtime (runtime.cu).
#pragma once
#include <stdint.h>
#include "layout.h"

namespace dnnb200 {

// Flags area of one rank: one row per source rank, kFlagRow words per row.
//   [src][0 .. kNumSlices)  sequence number of the newest slice src has pushed here
//   [src][kLeaveSlot]       "leave marker": first sequence number src will NOT take part in (0 = present)
constexpr int kFlagRow = 128;
constexpr int kLeaveSlot = kNumSlices;  // 122
static_assert(kLeaveSlot < kFlagRow, "flag row too small");

// Receive area of one rank: [2 parities][world][kFlatNumel] values stored as 16-byte LL lines
// {value, seq, value, seq} (exchange.cuh), i.e. 2*kFlatNumel floats per (parity, source).
// Parity = seq & 1, so the pushes of step t+1 never land on data a slow reader still sums for step t.

// flags argument of the update kernel (bits 0..2: kernel behaviour; bit 4: host-side kernel selection)
enum UpdateFlags : int {
  kUpdFlagGradOnly = 1,    // leave the local batch gradient in RankCtx::grads, touch nothing else (tests)
  kUpdFlagNoTrigger = 2,   // do not let the successor start early
  kUpdFlagTraceRing = 4,   // trace buffer keeps the last 1024 steps
  kUpdFlagPeers = 16,      // this rank exchanges with real peers: use the instantiation that fits beside the
                           // forward/backward kernel (see update_kernel)
  kUpdFlagHasGroup = 64,   // a group of more than one rank is attached (set even when the A/B switch clears bit 16)
};

enum StepStatus : unsigned int {
  kStatusOk = 0,
  kStatusTimeout = 1u,  // bit 0: some peer was waited for longer than timeout_ns
  // bits 8..23: bitmask of the ranks that timed out
};

struct StepCtl {
  unsigned long long step;        // optimizer steps completed; exchange sequence of the next step = step + 1
  unsigned long long timeout_ns;  // spin bound while waiting for a peer's slice (0 = wait forever)
  float lr;
  float mu;
  float loss_sum;           // sum of per-step mean losses since the host last cleared it
  unsigned int loss_steps;  // number of steps in loss_sum
  unsigned int ctas_done;   // last-CTA-out counter of the update kernel
  unsigned int status;      // StepStatus bits, sticky until the host clears them
  unsigned int active_mask; // ranks configured into the exchange (bit r = rank r)
  unsigned int two_shot;          // fused exchange variant: 0 one-shot with 16-byte lines, 3 one-shot with 128-byte lines
                                  // (both membership-safe), 1 two-shot, 2 NVLS (both need a static membership)
  unsigned int updates_done;      // update kernels completed on this rank, ever (never reset): the next step's
                                  // forward/backward kernel spins on it instead of waiting for grid completion
  unsigned int divisor_override;  // 0: divide by the ranks present; else by this (parent/children parity mode:
                                  // the non-training parent contributes zeros and the divisor is W-1)
};

struct RankCtx {
  float* params;          // [kFlatNumel] master weights
  float* momentum;        // [kFlatNumel]
  float* grads;           // [kFlatNumel] optional: averaged gradient written back (nullptr = skip)
  const float* saved;     // [B][kSavStride]   per-image FC inputs/deltas from the fwd/bwd kernel
  const float* part1;     // [B*CS][kOffW2]  conv1 weight/bias gradient partials (one per CTA of an image's cluster)
  const float* part2;     // [B][kConvPrefix-kOffW2]  per-image conv2 weight/bias gradients
  const float* img_loss;  // [B]
  float* loss_ring;       // [loss_ring_mask + 1] per-step mean loss, indexed by step & mask
  StepCtl* ctl;
  int loss_ring_mask;
  int rank;
  int world;
  int pad_;
  float* nvls_local;      // NVLS variant: this rank's symmetric buffer [2][kFlatNumel] floats + [2][kMaxWorld][kFlagRow] flags
  float* nvls_mc;         //   and the MULTICAST address of the same buffer on all ranks (nullptr: not available)
  long long* trace;       // development aid: update-kernel timestamps (globaltimer ns), nullptr in production
  float* packed;          // [kPackNumel] conv weights as tensor-core B operands (layout.h), refreshed with every update
  unsigned int* host_progress;   // pinned host word (device view) that receives StepCtl::updates_done after every update
                                 // kernel: how the host-fed path knows, without a stream event, that a ring region is free
  float* recv[kMaxWorld];          // recv[p]  = rank p's receive area as mapped in THIS process
  unsigned int* flags[kMaxWorld];  // flags[p] = rank p's flag area as mapped in THIS process
};

}  // namespace dnnb200
