// lookback.cuh -- single-pass "decoupled look-back" prefix machinery (Merrill & Garland) shared by the scan, the
// duplicate-collapse pass of the CSR build and (per digit) the onesweep sort.
#pragma once
#include "common.cuh"

namespace qdsp {

// descriptor word: status in the top 2 bits, value in the low 62.  cudaMemsetAsync(0) == all tiles INVALID.
constexpr unsigned long long LB_INVALID = 0ull;
constexpr unsigned long long LB_AGGREGATE = 1ull << 62;
constexpr unsigned long long LB_PREFIX = 2ull << 62;
constexpr unsigned long long LB_STATUS_MASK = 3ull << 62;
constexpr unsigned long long LB_VALUE_MASK = ~LB_STATUS_MASK;

#ifdef __CUDACC__
// Tiles take their index from an atomic ticket, so every tile with a smaller index has already started running:
// the spin below cannot deadlock whatever order the hardware schedules CTAs in.
__device__ __forceinline__ unsigned take_ticket(unsigned *counter, unsigned *smem_slot) {
  if (threadIdx.x == 0) *smem_slot = atomicAdd(counter, 1u);
  __syncthreads();
  return *smem_slot;
}

// Called by ONE full warp of the CTA.  Publishes this tile's aggregate, walks back over the predecessors' descriptors
// 32 at a time and returns (to every lane) the exclusive prefix of the tile; finally publishes the inclusive prefix.
// Values are summed modulo 2^62 (callers truncate to their width, so i32 wrap-around is preserved).
__device__ __forceinline__ unsigned long long lookback_exclusive(unsigned long long *desc, unsigned tile,
                                                                 unsigned long long aggregate, int shift = 0,
                                                                 unsigned backoff_ns = 0) {
  // shift: descriptors are 8 << shift bytes apart (padding spreads the hot descriptors over more L2 slices)
  const unsigned lane = lane_id();
  aggregate &= LB_VALUE_MASK;
  if (tile == 0) {
    if (lane == 0) st_relaxed_u64(&desc[0], LB_PREFIX | aggregate);
    return 0ull;
  }
  if (lane == 0) st_relaxed_u64(&desc[static_cast<size_t>(tile) << shift], LB_AGGREGATE | aggregate);
  unsigned long long exclusive = 0ull;
  long long idx = static_cast<long long>(tile) - 1;
  while (true) {
    const long long t = idx - lane;
    // wait only for the predecessors that matter: those nearer than the nearest one that already knows its inclusive
    // prefix (a straggler further back than that is irrelevant -- waiting for all 32 made every tile pay for the
    // slowest of its 32 predecessors)
    unsigned long long d;
    unsigned pmask;
    while (true) {
      d = (t >= 0) ? ld_relaxed_u64(&desc[static_cast<size_t>(t) << shift]) : LB_PREFIX;  // before tile 0: a prefix of zero
      const unsigned inv = __ballot_sync(FULL, (d & LB_STATUS_MASK) == LB_INVALID);
      pmask = __ballot_sync(FULL, (d & LB_STATUS_MASK) == LB_PREFIX);
      const unsigned nearer = pmask ? ((1u << (__ffs(pmask) - 1)) - 1u) : FULL;
      if ((inv & nearer) == 0u) break;
      if (backoff_ns) __nanosleep(backoff_ns);
    }
    unsigned long long v = d & LB_VALUE_MASK;
    if (pmask) {
      const unsigned first = __ffs(pmask) - 1;  // nearest predecessor that already knows its inclusive prefix
      if (lane > first) v = 0ull;
      exclusive += warp_sum(v);
      break;
    }
    exclusive += warp_sum(v);
    idx -= 32;
  }
  exclusive &= LB_VALUE_MASK;
  if (lane == 0) st_relaxed_u64(&desc[static_cast<size_t>(tile) << shift], LB_PREFIX | ((exclusive + aggregate) & LB_VALUE_MASK));
  return exclusive;
}

#endif

}  // namespace qdsp
