// Device side of the peer-memory gradient exchange (one-shot "push" all-reduce).
//
// Replaces the reference's rank-0-routed weight averaging
// (reference: data_parallelism_train.py:118,135,210,226-244 -- C1..C4 + R1 in
// SURVEY §2.5) with direct NVLink stores into every peer's receive area:
//
//   for every slice s of the flat vector, the CTA that owns s on rank r
//     1. stores its local values into recv[p][parity][r][s] of every peer p  (remote st over NVLink)
//     2. __syncthreads, fence.sys, st.release.sys flags[p][r][s] = seq        (publish)
//     3. spins on its OWN flags[r][p][s] >= seq for every peer p              (local ld.acquire.sys)
//     4. sums the W contributions in rank order 0..W-1 (its own from registers)
//        so every rank gets bit-identical results
//
// There is no barrier across slices or CTAs: slice s is exchanged while other
// CTAs still compute their wgrad tiles. Two parities of the receive area make the
// pushes of step t+1 safe against a slow reader of step t; flags carry a
// monotonically increasing sequence number and are never reset.
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

struct ExchangeShared {
  unsigned int present;  // bitmask of ranks whose contribution enters the sum
};

__device__ __forceinline__ bool seq_reached(unsigned int v, unsigned int seq) {
  return v != 0u && (int)(v - seq) >= 0;
}

// Values: each thread owns kVec consecutive floats at flat index `i` (16B/8B aligned).
// Returns the sum over present ranks in rank order; *n_present gets the divisor.
template <int kVec>
__device__ __forceinline__ void exchange_slice(const RankCtx& cx, ExchangeShared* sh, int slice,
                                               unsigned int seq, int i, bool in_range,
                                               float (&v)[kVec], int* n_present) {
  static_assert(kVec == 2 || kVec == 4, "vector width");
  const int world = cx.world, rank = cx.rank;
  const int tid = threadIdx.x;
  const unsigned int parity = seq & 1u;
  const unsigned int configured = cx.ctl->active_mask;
  unsigned int* my_flags = cx.flags[rank];

  // 1. push my values to every configured peer (departed peers keep their memory mapped,
  //    so a late store to them is harmless and saves a flag read here)
  if (in_range) {
    for (int p = 0; p < world; ++p) {
      if (p == rank || !((configured >> p) & 1u)) continue;
      float* dst = cx.recv[p] + ((size_t)(parity * world + rank) * kFlatNumel + i);
      if (kVec == 4) {
        *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
        *reinterpret_cast<float2*>(dst) = make_float2(v[0], v[1]);
      }
    }
  }
  __syncthreads();  // all stores of this CTA are issued ...
  if (tid < world && tid != rank && ((configured >> tid) & 1u)) {
    __threadfence_system();  // ... and ordered before the flag, cumulatively, at system scope
    st_release_sys(cx.flags[tid] + rank * kFlagRow + slice, seq);
  }

  // 2. wait for every configured peer's slice (or its leave marker)
  if (tid == 0) sh->present = 0u;
  __syncthreads();
  if (tid < world && ((configured >> tid) & 1u)) {
    bool here = false;
    if (tid == rank) {
      here = true;
    } else {
      const unsigned int* f = my_flags + tid * kFlagRow + slice;
      const unsigned int* lv = my_flags + tid * kFlagRow + kLeaveSlot;
      const unsigned long long tmo = cx.ctl->timeout_ns;
      unsigned long long t0 = 0;
      unsigned int spins = 0;
      for (;;) {
        if (seq_reached(ld_acquire_sys(f), seq)) { here = true; break; }
        const unsigned int l = ld_acquire_sys(lv);
        if (l != 0u && (int)(seq - l) >= 0) {
          // the marker may have overtaken nothing: p pushes every slice of its last step
          // before it writes the marker, so one more look at the flag settles it
          here = seq_reached(ld_acquire_sys(f), seq);
          break;
        }
        if (tmo != 0ull && (++spins & 0x3ffu) == 0u) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0ull) t0 = now;
          else if (now - t0 > tmo) {
            atomicOr(&cx.ctl->status, kStatusTimeout | (1u << (8 + tid)));
            break;
          }
        }
      }
    }
    if (here) atomicOr(&sh->present, 1u << tid);
  }
  __syncthreads();
  const unsigned int present = sh->present;

  // 3. rank-ordered sum; peers' data was written by another GPU -> read through L2
  float acc[kVec];
#pragma unroll
  for (int k = 0; k < kVec; ++k) acc[k] = 0.f;
  if (in_range) {
    const float* my_recv = cx.recv[rank];
    for (int p = 0; p < world; ++p) {
      if (!((present >> p) & 1u)) continue;
      if (p == rank) {
#pragma unroll
        for (int k = 0; k < kVec; ++k) acc[k] += v[k];
      } else {
        const float* src = my_recv + ((size_t)(parity * world + p) * kFlatNumel + i);
        if (kVec == 4) {
          const float4 t = ld_cg_f4(reinterpret_cast<const float4*>(src));
          acc[0] += t.x; acc[1] += t.y; acc[2] += t.z; acc[3] += t.w;
        } else {
          const float2 t = __ldcg(reinterpret_cast<const float2*>(src));
          acc[0] += t.x; acc[1] += t.y;
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < kVec; ++k) v[k] = acc[k];
  *n_present = __popc(present);
}

}  // namespace dnnb200
