// Device side of the peer-memory gradient exchange (one-shot "push" all-reduce).
//
// Replaces the reference's rank-0-routed weight averaging
// (reference: data_parallelism_train.py:118,135,210,226-244 -- C1..C4 + R1 in
// SURVEY §2.5) with direct NVLink stores into every peer's receive area:
//
//   for every slice s of the flat vector, the CTA that owns s on rank r
//     1. stores its local values into recv[p][parity][r][s] of every peer p as 16-byte
//        LL lines {value, seq, value, seq}                                   (remote st over NVLink)
//     2. polls its OWN recv[r][parity][p][s] until the lines of every peer p carry `seq`
//        (local ld.volatile; the data IS the flag: no fence, no second round trip)
//     3. sums the W contributions in rank order 0..W-1 (its own from registers)
//        so every rank gets bit-identical results
//
// There is no barrier across slices or CTAs: slice s is exchanged while other
// CTAs still compute their wgrad tiles. Two parities of the receive area make the
// pushes of step t+1 safe against a slow reader of step t; sequence numbers only
// grow and nothing is ever reset. (A first version used plain stores + fence.sys +
// a release flag per slice: ~9 us per step at 2 GPUs; see profiles/.)
//
// Membership (north-star item 3): a rank that leaves writes a
// "leave marker" L (the first sequence number it will not take part in) into
// every peer's flag row. A waiter treats rank p as absent for seq >= L, and the
// divisor becomes the number of ranks present. Because p either pushed all
// slices of a step or none, every survivor takes the same decision for the same
// step. A peer that neither pushes nor leaves within ctl->timeout_ns is treated
// as absent for this slice and StepCtl::status records it; the host then runs a
// membership rebuild + parameter re-sync.
#pragma once
#include "common.cuh"
#include "step_types.h"

namespace dnnb200 {

// 128-byte lines (one-shot-128): a slice of kSlice floats travels as kLines128 lines of 31 values + 1 flag word
constexpr int kLineData = 31;
constexpr int kLines128 = (kSlice + kLineData - 1) / kLineData;   // 17
constexpr int kSlice128 = kLines128 * 32;                         // 544 words per (slice, source) in the receive area

constexpr int kMaxPeers128 = 7;                                   // one NVSwitch domain: 8 GPUs

struct ExchangeShared {
  unsigned int present;  // bitmask of ranks whose contribution enters the sum
  alignas(16) float in[kLines128 * kLineData + 1];                 // 128-byte-line variant: my slice, and ...
  alignas(16) float rx[kMaxPeers128][kLines128 * kLineData + 1];   //   ... the lines of every peer as they arrive
};

// The two StepCtl fields every exchange needs, read by the caller (the update kernel reads them BEFORE it waits for
// its predecessor, so that no dependent global load sits between "gradient ready" and the first remote store).
struct ExchangeCfg {
  unsigned int configured;        // StepCtl::active_mask
  unsigned long long timeout_ns;  // StepCtl::timeout_ns
  unsigned int left;              // ranks whose leave marker already covers this step (leave_mask below): known to be
                                  // absent before the exchange starts, so nobody spins on them step after step
};
// Which configured ranks have announced that they take no part in sequence number `seq` any more? Their markers sit
// in MY flag rows. Read early (the update kernel does it before its dependency wait): a marker that appears later is
// still found by the waiters themselves, once, the slow way.
__device__ __forceinline__ unsigned int leave_mask(const RankCtx& cx, unsigned int configured, unsigned int seq) {
  unsigned int m = 0u;
  for (int p = 0; p < cx.world; ++p) {
    if (p == cx.rank || !((configured >> p) & 1u)) continue;
    const unsigned int l = ld_relaxed_sys(cx.flags[cx.rank] + p * kFlagRow + kLeaveSlot);
    if (l != 0u && (int)(seq - l) >= 0) m |= 1u << p;
  }
  return m;
}
__device__ __forceinline__ ExchangeCfg exchange_cfg(const RankCtx& cx, const StepCtl* ctl, unsigned int seq) {
  ExchangeCfg c;
  c.configured = ctl->active_mask;
  c.timeout_ns = ctl->timeout_ns;
  c.left = cx.world > 1 ? leave_mask(cx, c.configured, seq) : 0u;
  return c;
}

__device__ __forceinline__ bool seq_reached(unsigned int v, unsigned int seq) {
  return v != 0u && (int)(v - seq) >= 0;
}

// ---- LL ("low latency") lines: 16 bytes = {value0, seq, value1, seq} -------------------------
// A 16-byte store is delivered as two atomic 8-byte halves, so a reader that sees BOTH sequence
// words equal to `seq` also sees both values. Data and "it has arrived" travel in the same store:
// no fence and no separate flag round trip on the critical path (the idea of NCCL's LL protocol).
__device__ __forceinline__ void st_ll(float* line, float a, float b, unsigned int seq) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(line), "r"(__float_as_uint(a)), "r"(seq),
               "r"(__float_as_uint(b)), "r"(seq)
               : "memory");
}
__device__ __forceinline__ bool ld_ll(const float* line, unsigned int seq, float* a, float* b) {
  unsigned int x, f0, y, f1;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(f0), "=r"(y), "=r"(f1) : "l"(line)
               : "memory");
  *a = __uint_as_float(x);
  *b = __uint_as_float(y);
  return f0 == seq && f1 == seq;
}

// Receive area of one rank, LL layout: [2 parities][world][kFlatNumel] VALUES, each pair of values
// occupying one 16-byte line => 2 * kFlatNumel floats per (parity, source).
constexpr size_t kLLStride = 2 * (size_t)kFlatNumel;

// Values: each thread owns kVec consecutive floats at flat index `i` (kVec = 2 or 4, i % kVec == 0).
// Returns the sum over present ranks in rank order; *n_present gets the divisor.
template <int kVec>
__device__ __forceinline__ void exchange_slice(const RankCtx& cx, const ExchangeCfg cfg, ExchangeShared* sh,
                                               int slice, unsigned int seq, int i, bool in_range,
                                               float (&v)[kVec], int* n_present) {
  static_assert(kVec == 2 || kVec == 4, "vector width");
  const int world = cx.world, rank = cx.rank;
  const int tid = threadIdx.x;
  const unsigned int parity = seq & 1u;
  const unsigned int configured = cfg.configured & ~(cfg.left & ~(1u << rank));   // (ranks already known to be gone)
  const unsigned int* my_flags = cx.flags[rank];

  // 1. push my values to every configured peer: one 16-byte LL line per pair of floats.
  //    (Departed peers keep their memory mapped, so a late store to them is harmless.)
  if (in_range) {
    for (int p = 0; p < world; ++p) {
      if (p == rank || !((configured >> p) & 1u)) continue;
      float* dst = cx.recv[p] + ((size_t)(parity * world + rank) * kLLStride + 2 * (size_t)i);
#pragma unroll
      for (int k = 0; k < kVec; k += 2) st_ll(dst + 2 * k, v[k], v[k + 1], seq);
    }
  }

  // 2. gather: poll my own receive area until each peer's lines carry this step's sequence number.
  //    All pending peers are polled in ONE batch of independent loads per round (8 peers at a time): polling
  //    them one after the other costs one L2 round trip per peer even when everything has already arrived.
  if (tid == 0) sh->present = configured;
  __syncthreads();
  const float* my_recv = cx.recv[rank];
  const unsigned long long tmo = cfg.timeout_ns;
  float acc[kVec];
#pragma unroll
  for (int k = 0; k < kVec; ++k) acc[k] = 0.f;
  unsigned int absent = 0u;
  for (int g0 = 0; g0 < world; g0 += 8) {   // groups of 8 ranks (one group unless world > 8)
    float t[8][kVec];
    unsigned int pending = in_range ? ((configured >> g0) & 0xffu) : 0u;
    if (rank >= g0 && rank < g0 + 8) pending &= ~(1u << (rank - g0));
    if (world - g0 < 8) pending &= (1u << (world - g0)) - 1u;
    unsigned int spins = 0;
    unsigned long long t0 = 0;
    while (pending) {
      unsigned int arrived = 0u;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if ((pending >> j) & 1u) {
          const float* src = my_recv + ((size_t)(parity * world + g0 + j) * kLLStride + 2 * (size_t)i);
          bool ok = true;
#pragma unroll
          for (int k = 0; k < kVec; k += 2) ok = ld_ll(src + 2 * k, seq, &t[j][k], &t[j][k + 1]) && ok;
          if (ok) arrived |= 1u << j;
        }
      }
      pending &= ~arrived;
      if (!pending) break;
      if ((++spins & 0x3fu) != 0u) continue;
      // something is late: has it left, or is it time to give up?
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (!((pending >> j) & 1u)) continue;
        const unsigned int l = ld_acquire_sys(my_flags + (g0 + j) * kFlagRow + kLeaveSlot);
        if (l != 0u && (int)(seq - l) >= 0) {
          // a peer that left pushed every line of its last step BEFORE it wrote the marker: one more look settles it
          const float* src = my_recv + ((size_t)(parity * world + g0 + j) * kLLStride + 2 * (size_t)i);
          bool ok = true;
#pragma unroll
          for (int k = 0; k < kVec; k += 2) ok = ld_ll(src + 2 * k, seq, &t[j][k], &t[j][k + 1]) && ok;
          if (!ok) absent |= 1u << (g0 + j);
          pending &= ~(1u << j);
        }
      }
      if (pending && tmo != 0ull) {
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0ull) t0 = now;
        else if (now - t0 > tmo) {
          atomicOr(&cx.ctl->status, kStatusTimeout | (pending << (8 + g0)));
          absent |= pending << g0;
          pending = 0u;
        }
      }
    }
    // rank-ordered sum of this group (my own contribution from registers at its rank position)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int pr = g0 + j;
      if (pr >= world || !((configured >> pr) & 1u)) continue;
      if (pr == rank) {
#pragma unroll
        for (int k = 0; k < kVec; ++k) acc[k] += v[k];
      } else if (in_range && !((absent >> pr) & 1u)) {
#pragma unroll
        for (int k = 0; k < kVec; ++k) acc[k] += t[j][k];
      }
    }
  }
  // every thread of the CTA must agree on who is present (a rank pushes all lines of a step or none,
  // so disagreement can only come from a timeout racing with late data: take the union of "absent")
  if (absent) atomicAnd(&sh->present, ~absent);
  __syncthreads();
  const unsigned int present = sh->present;
  if (present == (configured & ~absent)) {  // the normal case: my view is the CTA's view
#pragma unroll
    for (int k = 0; k < kVec; ++k) v[k] = acc[k];
  } else if (in_range) {
    // another thread of this CTA declared a peer absent that I had received: redo the sum without it
#pragma unroll
    for (int k = 0; k < kVec; ++k) acc[k] = 0.f;
    for (int p = 0; p < world; ++p) {
      if (!((present >> p) & 1u)) continue;
      if (p == rank) {
#pragma unroll
        for (int k = 0; k < kVec; ++k) acc[k] += v[k];
      } else {
        const float* src = my_recv + ((size_t)(parity * world + p) * kLLStride + 2 * (size_t)i);
        float t2[kVec];
#pragma unroll
        for (int k = 0; k < kVec; k += 2) (void)ld_ll(src + 2 * k, seq, &t2[k], &t2[k + 1]);
#pragma unroll
        for (int k = 0; k < kVec; ++k) acc[k] += t2[k];
      }
    }
#pragma unroll
    for (int k = 0; k < kVec; ++k) v[k] = acc[k];
  }
  *n_present = __popc(present);
}

// ---- one-shot with 128-byte lines ------------------------------------------------------------------
// The 16-byte LL lines above spend half of every NVLink byte on sequence numbers; at 8 GPUs the one-shot push
// then moves 7 x 497 KB per GPU and loses to two-shot. Here a line is 128 bytes: 31 values + ONE flag word,
// written by one warp-wide store instruction (32 lanes x 4 bytes, one fully covered 128-byte transaction) and
// read back by one warp-wide load. A 128-byte aligned write travels over NVLink and lands in the peer's L2 as
// one unit -- the property NCCL's LL128 protocol is built on -- so a reader whose load returns the flag word
// of this step also holds the 31 values of this step. 1/32 of the bytes is overhead instead of 1/2:
// 7 x 265 KB per GPU at 8 GPUs, still ONE latency hop, and membership-safe like the 16-byte one-shot
// (leave markers, timeout, the CTA agrees on one set of contributors, rank-ordered sum => identical bits).
// The thread-owned values are picked up line by line from shared memory (sh->in); arriving lines are parked there (sh->rx).
// tests/test_gpu_multiproc.py::test_ll128_lines_are_never_torn hammers the atomicity assumption over NVLink.
__device__ __forceinline__ void st_line_word(float* p, unsigned int w) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(w) : "memory");
}
__device__ __forceinline__ unsigned int ld_line_word(const float* p) {
  unsigned int w;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(w) : "l"(p) : "memory");
  return w;
}

// n_valid = how many of the slice's kSlice values exist (the last slice of a vector is short): only the lines
// that carry one of them travel. At most kMaxPeers128 peers (one NVSwitch domain).
//
// Gather = independent jobs. Every (line, peer) pair is one job; the jobs are dealt round-robin to the warps and
// a warp keeps all of its pending jobs in flight at once (one batch of loads per polling round), parks each line
// that has arrived in shared memory (sh->rx[peer slot]) and drops it from its pending mask. Nothing waits for a
// particular line, so a late peer costs one latency, not one per line. After one block barrier every thread adds,
// in rank order, its own values and the parked values of the ranks the CTA agreed on.
template <int kVec>
__device__ __forceinline__ void exchange_slice_ll128(const RankCtx& cx, const ExchangeCfg cfg, ExchangeShared* sh,
                                                     int slice, unsigned int seq, int n_valid, float (&v)[kVec],
                                                     int* n_present) {
  static_assert(kVec == 2 || kVec == 4, "vector width");
  constexpr int kWarps = kSlice / kVec / 32;
  constexpr int kMaxJobs = (kLines128 * kMaxPeers128 + kWarps - 1) / kWarps;   // 15 (8 warps) / 30 (4 warps)
  static_assert(kMaxJobs <= 32, "a warp tracks its jobs in one 32-bit mask");
  const int world = cx.world, rank = cx.rank;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const unsigned int parity = seq & 1u;
  const unsigned int configured = cfg.configured & ~(cfg.left & ~(1u << rank));   // (ranks already known to be gone)
  const unsigned long long tmo = cfg.timeout_ns;
  const unsigned int* my_flags = cx.flags[rank];
  const int n_lines = (n_valid + kLineData - 1) / kLineData;
  const size_t area = (size_t)slice * kSlice128;   // word offset of this slice inside a (parity, source) block

  // 1. my values where a warp can pick them up line by line (word w of the slice at in[w])
#pragma unroll
  for (int k = 0; k < kVec; ++k) sh->in[kVec * tid + k] = v[k];
  if (tid == 0) sh->present = configured;
  // the configured peers, compacted: slot q <-> rank (same on every thread: `configured` is uniform)
  int n_peers = 0;
  unsigned int peer_of = 0u;   // 4 bits per slot
  for (int p = 0; p < world; ++p)
    if (p != rank && ((configured >> p) & 1u) && n_peers < kMaxPeers128) peer_of |= (unsigned int)p << (4 * n_peers++);
  __syncthreads();

  // 2. push: one warp-wide store per (line, peer)
  for (int L = warp; L < n_lines; L += kWarps) {
    const int w = L * kLineData + lane;
    const unsigned int word = lane == kLineData ? seq : (w < kSlice ? __float_as_uint(sh->in[w]) : 0u);
    for (int q = 0; q < n_peers; ++q) {
      const int p = (int)((peer_of >> (4 * q)) & 15u);
      st_line_word(cx.recv[p] + ((size_t)(parity * world + rank) * kLLStride + area + (size_t)L * 32 + lane), word);
    }
  }

  // 3. gather: job j = line * kMaxPeers128 + peer slot; warp w owns the jobs j = w + kWarps * i
  const float* my_recv = cx.recv[rank];
  unsigned int pending = 0u;
#pragma unroll
  for (int i = 0; i < kMaxJobs; ++i) {
    const int j = warp + kWarps * i, L = j / kMaxPeers128, q = j - L * kMaxPeers128;
    if (L < n_lines && q < n_peers) pending |= 1u << i;
  }
  unsigned int absent = 0u;    // ranks this warp gave up on (warp-uniform)
  unsigned int spins = 0;
  unsigned long long t0 = 0;
  while (pending) {
#pragma unroll
    for (int i0 = 0; i0 < kMaxJobs; i0 += 8) {   // up to 8 independent loads in flight, THEN the checks
      unsigned int got[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u;
        if (i < kMaxJobs && ((pending >> i) & 1u)) {
          const int j = warp + kWarps * i, L = j / kMaxPeers128, q = j - L * kMaxPeers128;
          const int p = (int)((peer_of >> (4 * q)) & 15u);
          got[u] = ld_line_word(my_recv + ((size_t)(parity * world + p) * kLLStride + area + (size_t)L * 32 + lane));
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u;
        if (i < kMaxJobs && ((pending >> i) & 1u)) {
          const int j = warp + kWarps * i, L = j / kMaxPeers128, q = j - L * kMaxPeers128;
          if (__shfl_sync(0xffffffffu, got[u], kLineData) == seq) {
            if (lane < kLineData) sh->rx[q][L * kLineData + lane] = __uint_as_float(got[u]);
            pending &= ~(1u << i);
          }
        }
      }
    }
    if (!pending) break;
    if ((++spins & 0x3fu) != 0u) continue;
    // something is late: has its sender left (announced), or is it time to give up? (all lanes take the same path)
    bool expired = false;
    if (tmo != 0ull) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0ull) t0 = now;
      expired = __shfl_sync(0xffffffffu, (now - t0 > tmo) ? 1 : 0, 0) != 0;
    }
    for (int i = 0; i < kMaxJobs; ++i) {
      if (!((pending >> i) & 1u)) continue;
      const int j = warp + kWarps * i, L = j / kMaxPeers128, q = j - L * kMaxPeers128;
      const int p = (int)((peer_of >> (4 * q)) & 15u);
      const unsigned int l = ld_acquire_sys(my_flags + p * kFlagRow + kLeaveSlot);
      const bool left = l != 0u && (int)(seq - l) >= 0;
      if (!left && !expired) continue;
      if (left) {
        // a peer that left pushed every line of its last step BEFORE it wrote the marker: one more look settles it
        const unsigned int got = ld_line_word(my_recv + ((size_t)(parity * world + p) * kLLStride + area + (size_t)L * 32 + lane));
        if (__shfl_sync(0xffffffffu, got, kLineData) == seq) {
          if (lane < kLineData) sh->rx[q][L * kLineData + lane] = __uint_as_float(got);
          pending &= ~(1u << i);
          continue;
        }
      } else if (lane == 0) {
        atomicOr(&cx.ctl->status, kStatusTimeout | (1u << (8 + p)));
      }
      absent |= 1u << p;
      pending &= ~(1u << i);
    }
  }
  // 4. the CTA agrees on ONE set of contributors (a rank pushes all lines of a step or none, so views only differ
  //    when a timeout races with late data): everybody's "absent" is taken out
  if (absent && lane == 0) atomicAnd(&sh->present, ~absent);
  __syncthreads();
  const unsigned int present = sh->present;
  // 5. rank-ordered sum, my own contribution (registers) at its rank position
  float acc[kVec];
#pragma unroll
  for (int k = 0; k < kVec; ++k) acc[k] = 0.f;
  {
    int q = 0;
    for (int p = 0; p < world; ++p) {
      if (p == rank) {
        if ((configured >> p) & 1u) {
#pragma unroll
          for (int k = 0; k < kVec; ++k) acc[k] += v[k];
        }
        continue;
      }
      if (!((configured >> p) & 1u) || q >= n_peers) continue;
      if ((present >> p) & 1u) {
#pragma unroll
        for (int k = 0; k < kVec; ++k) acc[k] += sh->rx[q][kVec * tid + k];
      }
      ++q;
    }
  }
  if (kVec * tid < n_valid) {
#pragma unroll
    for (int k = 0; k < kVec; ++k) v[k] = acc[k];
  }
  *n_present = __popc(present);
}

// ---- two-shot variant ------------------------------------------------------------------------
// Slice s is owned by the (s mod n)-th configured rank. Everybody else sends its values for s only to
// the owner (LL lines, src slot = sender); the owner adds the n contributions in rank order and sends
// the SUM to everybody (src slot = owner, which nobody else uses for this slice). Two latency hops, but
// 2(n-1)/n x the vector per GPU instead of (n-1) x: wins from about 4 GPUs on (measured table in
// profiles/). All ranks end up with the same bits because all use the owner's sum.
// Ownership is static: this variant is only valid while no configured rank leaves (the host picks the
// one-shot variant whenever the failure drill can drop ranks).
// What a stall does (timeout_ns != 0), so that nobody ever applies garbage:
//   * a contribution that does not reach the OWNER in time is left out, for the whole slice (the CTA agrees on
//     one set, like the one-shot path); the owner scales the sum by n_configured / n_arrived before it sends it,
//     so every rank still divides by n_configured and all of them apply the same mean over the ranks that arrived;
//   * a rank whose OWNER stays silent keeps its own local gradient for the slice (divisor 1: a local SGD step).
//     Its replica then differs from the others in that slice until the host re-averages (Trainer._resync_state).
// Either way StepCtl::status carries kStatusTimeout and the mask of the ranks given up on.
__device__ __forceinline__ bool ll_poll(const float* line, unsigned int seq, float* a, float* b,
                                        unsigned long long tmo) {
  unsigned int spins = 0;
  unsigned long long t0 = 0;
  while (!ld_ll(line, seq, a, b)) {
    if (tmo != 0ull && (++spins & 0x3ffu) == 0u) {
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0ull) t0 = now;
      else if (now - t0 > tmo) return false;
    }
  }
  return true;
}

template <int kVec>
__device__ __forceinline__ void exchange_slice_two_shot(const RankCtx& cx, const ExchangeCfg cfg, ExchangeShared* sh,
                                                        int slice, unsigned int seq, int i, bool in_range,
                                                        float (&v)[kVec], int* n_present) {
  static_assert(kVec == 2 || kVec == 4, "vector width");
  const int world = cx.world, rank = cx.rank;
  const unsigned int parity = seq & 1u;
  const unsigned int configured = cfg.configured;
  const int n_cfg = __popc(configured);
  const unsigned long long tmo = cfg.timeout_ns;
  *n_present = n_cfg;
  int owner = 0;
  {
    int k = slice % n_cfg;
    for (int p = 0; p < world; ++p)
      if ((configured >> p) & 1u) { if (k == 0) { owner = p; break; } --k; }
  }
  if (!((configured >> rank) & 1u)) return;   // (whole CTA: rank is uniform)
  const float* my_recv = cx.recv[rank];
  if (rank != owner) {
    if (!in_range) return;
    float* dst = cx.recv[owner] + ((size_t)(parity * world + rank) * kLLStride + 2 * (size_t)i);
#pragma unroll
    for (int k = 0; k < kVec; k += 2) st_ll(dst + 2 * k, v[k], v[k + 1], seq);
    const float* src = my_recv + ((size_t)(parity * world + owner) * kLLStride + 2 * (size_t)i);
    float got[kVec];
    bool ok = true;
#pragma unroll
    for (int k = 0; k < kVec; k += 2) ok = ll_poll(src + 2 * k, seq, &got[k], &got[k + 1], tmo) && ok;
    if (ok) {
#pragma unroll
      for (int k = 0; k < kVec; ++k) v[k] = got[k];
    } else {   // silent owner: keep my own gradient for these elements, as a local step
      atomicOr(&cx.ctl->status, kStatusTimeout | (1u << (8 + owner)));
      *n_present = 1;
    }
  } else {
    if (threadIdx.x == 0) sh->present = configured;
    __syncthreads();
    unsigned int absent = 0u;
    float acc[kVec];
#pragma unroll
    for (int k = 0; k < kVec; ++k) acc[k] = 0.f;
    for (int g0 = 0; g0 < world; g0 += 8) {   // all pending contributions are polled as one batch of loads
      float t[8][kVec];
      unsigned int pending = in_range ? ((configured >> g0) & 0xffu) : 0u;
      if (rank >= g0 && rank < g0 + 8) pending &= ~(1u << (rank - g0));
      if (world - g0 < 8) pending &= (1u << (world - g0)) - 1u;
      unsigned int spins = 0;
      unsigned long long t0 = 0;
      while (pending) {
        unsigned int arrived = 0u;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if ((pending >> j) & 1u) {
            const float* src = my_recv + ((size_t)(parity * world + g0 + j) * kLLStride + 2 * (size_t)i);
            bool ok = true;
#pragma unroll
            for (int k = 0; k < kVec; k += 2) ok = ld_ll(src + 2 * k, seq, &t[j][k], &t[j][k + 1]) && ok;
            if (ok) arrived |= 1u << j;
          }
        }
        pending &= ~arrived;
        if (pending && tmo != 0ull && (++spins & 0x3ffu) == 0u) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0ull) t0 = now;
          else if (now - t0 > tmo) {
            atomicOr(&cx.ctl->status, kStatusTimeout | (pending << (8 + g0)));
            absent |= pending << g0;
            pending = 0u;
          }
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int pr = g0 + j;
        if (pr >= world || !((configured >> pr) & 1u)) continue;
        if (pr == rank) {
#pragma unroll
          for (int k = 0; k < kVec; ++k) acc[k] += v[k];
        } else if (in_range && !((absent >> pr) & 1u)) {
#pragma unroll
          for (int k = 0; k < kVec; ++k) acc[k] += t[j][k];
        }
      }
    }
    // one set of contributors for the whole slice: the union of what any thread gave up on (normally nothing)
    if (absent) atomicAnd(&sh->present, ~absent);
    __syncthreads();
    const unsigned int present = sh->present;
    if (!in_range) return;
    if (present != (configured & ~absent)) {   // somebody else dropped a rank whose line I did receive: redo the sum
#pragma unroll
      for (int k = 0; k < kVec; ++k) acc[k] = 0.f;
      for (int p = 0; p < world; ++p) {
        if (!((present >> p) & 1u)) continue;
        if (p == rank) {
#pragma unroll
          for (int k = 0; k < kVec; ++k) acc[k] += v[k];
        } else {
          const float* src = my_recv + ((size_t)(parity * world + p) * kLLStride + 2 * (size_t)i);
          float t2[kVec];
#pragma unroll
          for (int k = 0; k < kVec; k += 2) (void)ld_ll(src + 2 * k, seq, &t2[k], &t2[k + 1]);
#pragma unroll
          for (int k = 0; k < kVec; ++k) acc[k] += t2[k];
        }
      }
    }
    if (present != configured) {   // everyone divides by n_cfg: make that the mean over the ranks that arrived
      const float fix = (float)n_cfg / (float)__popc(present);
#pragma unroll
      for (int k = 0; k < kVec; ++k) acc[k] *= fix;
    }
    for (int p = 0; p < world; ++p) {
      if (p == rank || !((configured >> p) & 1u)) continue;
      float* dst = cx.recv[p] + ((size_t)(parity * world + rank) * kLLStride + 2 * (size_t)i);
#pragma unroll
      for (int k = 0; k < kVec; k += 2) st_ll(dst + 2 * k, acc[k], acc[k + 1], seq);
    }
#pragma unroll
    for (int k = 0; k < kVec; ++k) v[k] = acc[k];
  }
}

// ---- NVLS variant (in-switch reduction) ------------------------------------------------------------
// Each rank parks its values in its own symmetric buffer; ONE multimem.st publishes "slice s of rank r is
// there" to every rank; multimem.ld_reduce returns the switch's sum of all W copies. Static membership (the
// switch adds every member of the multicast group). Used when the host exposes NVLink multicast.
// A member that does not publish its slice in time cannot be left out of the switch's sum, so the whole slice
// falls back to this rank's own gradient (divisor 1, a local step) and StepCtl::status says so.
template <int kVec>
__device__ __forceinline__ void exchange_slice_nvls(const RankCtx& cx, const ExchangeCfg cfg, ExchangeShared* sh,
                                                    int slice, unsigned int seq, int i, bool in_range,
                                                    float (&v)[kVec], int* n_present) {
  static_assert(kVec == 2 || kVec == 4, "vector width");
  const int world = cx.world, rank = cx.rank, tid = threadIdx.x;
  const unsigned int parity = seq & 1u;
  const unsigned int configured = cfg.configured;
  float* loc = cx.nvls_local;
  float* mc = cx.nvls_mc;
  *n_present = world;
  if (tid == 0) sh->present = 1u;   // here: "every member published in time"
  if (in_range) {
    float* dst = loc + (size_t)parity * kFlatNumel + i;
    if (kVec == 4) *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    else *reinterpret_cast<float2*>(dst) = make_float2(v[0], v[1]);
  }
  __syncthreads();
  const unsigned int* lflags = reinterpret_cast<const unsigned int*>(loc + 2 * (size_t)kFlatNumel);
  unsigned int* mflags = reinterpret_cast<unsigned int*>(mc + 2 * (size_t)kFlatNumel);
  if (tid == 0) {
    __threadfence_system();  // the CTA's stores, cumulatively, before the flag
    asm volatile("multimem.st.relaxed.sys.global.b32 [%0], %1;" ::"l"(mflags + (parity * kMaxWorld + rank) * kFlagRow + slice),
                 "r"(seq)
                 : "memory");
  }
  if (tid < world && ((configured >> tid) & 1u)) {
    const unsigned int* f = lflags + (parity * kMaxWorld + tid) * kFlagRow + slice;
    const unsigned long long tmo = cfg.timeout_ns;
    unsigned long long t0 = 0;
    unsigned int spins = 0;
    while (ld_acquire_sys(f) != seq) {
      if (tmo != 0ull && (++spins & 0x3ffu) == 0u) {
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0ull) t0 = now;
        else if (now - t0 > tmo) {
          atomicOr(&cx.ctl->status, kStatusTimeout | (1u << (8 + tid)));
          sh->present = 0u;
          break;
        }
      }
    }
  }
  __syncthreads();
  if (sh->present == 0u) {
    *n_present = 1;
    return;
  }
  if (in_range) {
    const float* src = mc + (size_t)parity * kFlatNumel + i;
    if (kVec == 4) {
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3])
                   : "l"(src)
                   : "memory");
    } else {
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v2.f32 {%0, %1}, [%2];"
                   : "=f"(v[0]), "=f"(v[1])
                   : "l"(src)
                   : "memory");
    }
  }
}

}  // namespace dnnb200
