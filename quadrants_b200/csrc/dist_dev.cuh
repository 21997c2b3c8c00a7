// dist_dev.cuh -- device side of the multi-GPU path: what a CG / SpMV kernel needs to talk to the other ranks over
// NVLink peer memory, with no collective library in the loop.
//
// The reference has nothing here (single device, quadrants/rhi/cuda/cuda_context.cpp:16-26); the design is SURVEY 8(e):
// row-block partition, ghost entries of x written by their owner straight into the consumer's vector (peer stores
// through NVSwitch), CG scalars all-reduced in-kernel.
//
// Every rank owns one "comm buffer" (cudaMalloc, exported with cudaIpcGetMemHandle, mapped by every peer):
//   [0, 2048)      reduction slots   slots[2 parity][8 ranks][8 values][2 words]   8-byte words {payload32 | seq32}
//   [2048, 2112)   halo flags        flags[8 ranks]                                 8-byte, monotone sequence numbers
//   [2112, 2176)   z-halo flags      the same for the second halo channel (below)
//   [2176, 2240)   consumed flags    stand-alone SpMV: flags[q] = calls whose ghosts rank q has finished reading (dist.cu)
//   [4096, ...)    the operand vector of SpMV: n_loc owned entries, the ghost entries (from ghost_base), and from zbase a
//                  second ghost-sized segment for the z channel
// Protocol:
//   * halo: after a rank has produced its owned part of the vector it stores the entries each neighbour needs into
//     that neighbour's ghost segment, fences at system scope, and stores its push count into the neighbour's
//     flags[rank].  A consumer kernel spins (ld.acquire.sys) until flags[q] >= its own push count for every
//     neighbour q it receives from -- ranks run the same kernel sequence, so "my push count" is the expected value.
//   * second channel ("z"), used by CG: instead of pushing the new direction p' = z' + beta p at the END of an iteration
//     (the neighbour's next SpMV then waits for it, and the push sits on the critical path), the owner pushes
//     z' = D^-1 r' as soon as r' exists -- in the residual kernel, whose own reduction + all-reduce hides the NVLink
//     latency -- and every rank updates its ghost copies of p itself, p_ghost = fma(beta, p_ghost, z_ghost): the same
//     operands and the same fma as the owner's, hence the same bits.  The SpMV never waits after the first iteration.
//   * all-reduce of up to 8 doubles (NCCL "LL" style): each 64-bit value travels as two 8-byte words carrying 32 data
//     bits + the 32-bit sequence number; an aligned 8-byte store is single-copy atomic, so data and flag arrive
//     together and one NVLink latency is the whole cost.  Slots alternate by sequence parity: a rank can be at most
//     one reduction ahead of a peer (it needs the peer's value to finish the current one).  Every rank sums the 8
//     contributions in rank order, so all ranks compute bit-identical scalars and take identical decisions.
//   * a spin that lasts longer than ~4 s raises *error and stops: a dead peer cannot hang the GPU.
#pragma once
#include "common.cuh"

namespace qdsp {

constexpr int DIST_MAX_RANKS = 8;
constexpr int DIST_MAX_VALUES = 8;
static_assert(2 * DIST_MAX_RANKS * DIST_MAX_VALUES * 2 * 8 <= 2048, "reduction slots overlap the halo flags");
constexpr size_t DIST_SLOTS_OFF = 0;
constexpr size_t DIST_FLAGS_OFF = 2048;
constexpr size_t DIST_VECTOR_OFF = 4096;
constexpr long long DIST_SPIN_LIMIT = 8000000000LL;  // clock64 ticks (~4 s)

struct DistDev {
  int rank = 0, world = 1;
  unsigned char *self = nullptr;                 // my comm buffer
  unsigned char *peer[DIST_MAX_RANKS] = {};      // peer[q] = rank q's comm buffer mapped here (peer[rank] == self)
  // halo: what I send
  int n_send = 0;
  int send_peer[DIST_MAX_RANKS] = {};
  const int *send_list[DIST_MAX_RANKS] = {};     // device, sorted local row ids
  int send_count[DIST_MAX_RANKS] = {};
  int send_dst[DIST_MAX_RANKS] = {};             // element offset inside the peer's operand vector (its ghost slot)
  int send_dst_z[DIST_MAX_RANKS] = {};           // the same for the z channel (the peer's zbase + its owner offset)
  int zmode = 0;                                 // 1: every neighbour registered a z slot -> CG uses the z channel
  int ghost_base = 0, n_ghost = 0, zbase = 0;    // my own vector: first ghost, ghost count, first z-ghost
  int send_first[DIST_MAX_RANKS] = {};           // smallest / largest row of the list, and whether it is one contiguous
  int send_last[DIST_MAX_RANKS] = {};            // run (plane halos): lets a CTA skip or index the list without searching
  int send_contig[DIST_MAX_RANKS] = {};
  // halo: whom I receive from
  int n_recv = 0;
  int recv_peer[DIST_MAX_RANKS] = {};
  // which SpMV row blocks (of ghost_block_rows rows) read a ghost column: only those wait for the halo
  const unsigned char *ghost_blocks = nullptr;
  int ghost_block_rows = 0;
  // the same information as a work order: ids of the row blocks that read no ghost (ascending), then of those that do.
  // Every CTA walks a contiguous run of interior blocks first and takes its share of the ghost blocks last, so the
  // NVLink latency of the halo hides behind interior work.
  const int *block_order = nullptr;
  int n_interior = 0;
  const int *block_order_all = nullptr;          // the same list, always there (block_order is set only when it is to be used)
  // amode (the CG's A.p kernel, cg.cu): the kernel that READS the operand also sends its halo -- the CTAs that own a halo
  // row push it before their first block, everybody reads the interior blocks first.  Then the number of pushes the
  // flags must show is *want_ptr + 1 (want_ptr: a device word holding the push count at the start of the kernel;
  // counters[1] itself advances while the kernel runs).
  int amode = 0;
  const unsigned *want_ptr = nullptr;
  unsigned *counters = nullptr;                  // device, local: [0] reduction seq, [1] halo push count, [2] ticket,
                                                 // [3] z push count, [4] z ticket, [5] stand-alone SpMV calls
  int *error = nullptr;                          // device, local: != 0 after a timeout
};

#ifdef __CUDACC__
__device__ __forceinline__ void st_sys_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long *dist_slot(unsigned char *buf, unsigned parity, int from_rank, int j) {
  return reinterpret_cast<unsigned long long *>(buf + DIST_SLOTS_OFF) +
         ((static_cast<size_t>(parity) * DIST_MAX_RANKS + from_rank) * DIST_MAX_VALUES + j) * 2;
}
__device__ __forceinline__ unsigned long long *dist_flag(unsigned char *buf, int from_rank) {
  return reinterpret_cast<unsigned long long *>(buf + DIST_FLAGS_OFF) + from_rank;
}

__device__ __forceinline__ unsigned long long *dist_flag_z(unsigned char *buf, int from_rank) {
  return reinterpret_cast<unsigned long long *>(buf + DIST_FLAGS_OFF + 64) + from_rank;
}

// All-reduce (sum) of v[0..NV) across ranks.  Called by ALL 32 lanes of ONE warp of ONE CTA per rank; every lane
// returns the totals.  world == 1: identity.
template <int NV>
__device__ __forceinline__ void dist_allreduce(const DistDev &d, double (&v)[NV]) {
  static_assert(NV <= DIST_MAX_VALUES, "too many values");
  if (d.world <= 1) return;
  const int lane = lane_id();
  unsigned seq = 0;
  if (lane == 0) seq = atomicAdd(&d.counters[0], 1u) + 1u;  // one round trip (a load + store pair would be two)
  seq = __shfl_sync(FULL, seq, 0);
  const unsigned parity = seq & 1u;
  // lane q talks to rank q
  if (lane < d.world && lane != d.rank) {
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const unsigned long long bits = static_cast<unsigned long long>(__double_as_longlong(v[j]));
      unsigned long long *dst = dist_slot(d.peer[lane], parity, d.rank, j);
      st_sys_u64(dst, ((bits & 0xffffffffull) << 32) | seq);
      st_sys_u64(dst + 1, ((bits >> 32) << 32) | seq);
    }
  }
  double got[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) got[j] = v[j];
  if (lane < d.world && lane != d.rank) {
    const long long t0 = clock64();
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const unsigned long long *src = dist_slot(d.self, parity, lane, j);
      unsigned long long lo, hi;
      while (true) {
        lo = ld_sys_u64(src);
        hi = ld_sys_u64(src + 1);
        if (static_cast<unsigned>(lo) == seq && static_cast<unsigned>(hi) == seq) break;
        if (clock64() - t0 > DIST_SPIN_LIMIT) {
          atomicExch(d.error, 1);
          lo = hi = 0;
          break;
        }
      }
      got[j] = __longlong_as_double(static_cast<long long>((lo >> 32) | ((hi >> 32) << 32)));
    }
  }
  // fixed rank order: identical bits on every rank
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    double s = 0.0;
    for (int q = 0; q < d.world; ++q) s += __shfl_sync(FULL, got[j], q);
    v[j] = s;
  }
}

__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

// Stores the entries of vec[lo, hi) that neighbours need into their ghost segments and signals them.  Called by all
// threads of every CTA of a kernel that has just written vec[lo, hi); the chunks [lo, hi) of the CTAs are the
// cta-chunk partition of [0, n): CTA b owns [b * per, (b + 1) * per), per = chunk size (a multiple of 1024).
// Only the CTAs that own a halo row take part: each stores its part (16-byte stores where the run is contiguous and
// aligned), fences ONCE at system scope (the wait for the NVLink write acknowledgements) and takes a ticket; the last
// of them publishes the new push count to the neighbours.  The other CTAs -- all but a few dozen -- do nothing, and
// the signal leaves as soon as the halo rows are done instead of at the end of the kernel (measured at 8 ranks: the
// direction kernel took 25 us with an all-CTA ticket against 10 us without any push, profiles/r01h_8gpu_trace.txt).
// Z = false: channel p, pushes vec[i].  Z = true: channel z, pushes vec[i] * mul[i] (r and D^-1).
// NT = 0: called by every thread of the CTA.  NT > 0: called by threads [0, NT) only (the consumer warps of the persistent
// CG kernel; the producer warp is elsewhere) -- they synchronise on named barrier 1 and stride by NT.
template <int NT>
__device__ __forceinline__ void dist_push_bar() {
  if constexpr (NT == 0) __syncthreads();
  else asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
}
// owners_override > 0: the caller's CTAs are not the uniform cta-chunk partition; that many CTAs own a halo row.
template <typename T, bool Z, int NT = 0>
__device__ __forceinline__ void dist_halo_push_impl(const DistDev &d, const T *vec, const T *mul, int lo, int hi, int per,
                                                    int owners_override = 0) {
  if (d.world <= 1) return;
  const int nthreads = NT == 0 ? static_cast<int>(blockDim.x) : NT;
  bool mine = false;
  for (int s = 0; s < d.n_send; ++s) mine = mine || (lo < hi && hi > d.send_first[s] && lo <= d.send_last[s]);
  if (!mine) return;
  dist_push_bar<NT>();  // the CTA's own writes of vec[lo, hi) are visible to all its threads
  for (int s = 0; s < d.n_send; ++s) {
    if (hi <= d.send_first[s] || lo > d.send_last[s]) continue;
    const int *list = d.send_list[s];
    const int cnt = d.send_count[s];
    T *dst = reinterpret_cast<T *>(d.peer[d.send_peer[s]] + DIST_VECTOR_OFF) + (Z ? d.send_dst_z[s] : d.send_dst[s]);
    if (d.send_contig[s]) {
      const int first = d.send_first[s];
      const int jlo = max(lo, first) - first, jhi = min(hi, d.send_last[s] + 1) - first;
      constexpr int V = 16 / sizeof(T);
      const bool vec_ok = ((reinterpret_cast<uintptr_t>(dst + jlo) | reinterpret_cast<uintptr_t>(vec + first + jlo)) & 15u) == 0;
      if (vec_ok && !Z) {
        const int nv = (jhi - jlo) / V;
        const int4 *src4 = reinterpret_cast<const int4 *>(vec + first + jlo);
        int4 *dst4 = reinterpret_cast<int4 *>(dst + jlo);
        for (int j = threadIdx.x; j < nv; j += nthreads) dst4[j] = src4[j];
        for (int j = jlo + nv * V + threadIdx.x; j < jhi; j += nthreads) dst[j] = vec[first + j];
      } else if (vec_ok && ((reinterpret_cast<uintptr_t>(mul + first + jlo) & 15u) == 0)) {
        const int nv = (jhi - jlo) / V;
        for (int j = threadIdx.x; j < nv; j += nthreads) {
          T a[V], m[V];
          *reinterpret_cast<int4 *>(a) = reinterpret_cast<const int4 *>(vec + first + jlo)[j];
          *reinterpret_cast<int4 *>(m) = reinterpret_cast<const int4 *>(mul + first + jlo)[j];
#pragma unroll
          for (int e = 0; e < V; ++e) a[e] = m[e] * a[e];
          reinterpret_cast<int4 *>(dst + jlo)[j] = *reinterpret_cast<int4 *>(a);
        }
        for (int j = jlo + nv * V + threadIdx.x; j < jhi; j += nthreads) dst[j] = mul[first + j] * vec[first + j];
      } else {
        for (int j = jlo + threadIdx.x; j < jhi; j += nthreads) dst[j] = Z ? mul[first + j] * vec[first + j] : vec[first + j];
      }
    } else {
      int a = 0, b = cnt;
      while (a < b) { const int m = (a + b) >> 1; if (list[m] < lo) a = m + 1; else b = m; }
      const int jlo = a;
      b = cnt;
      while (a < b) { const int m = (a + b) >> 1; if (list[m] < hi) a = m + 1; else b = m; }
      const int jhi = a;
      for (int j = jlo + threadIdx.x; j < jhi; j += nthreads) dst[j] = Z ? mul[list[j]] * vec[list[j]] : vec[list[j]];
    }
  }
  dist_push_bar<NT>();
  if (threadIdx.x == 0) {
    // number of CTAs that own a halo row: chunks [c * per, (c + 1) * per) meeting any [send_first, send_last]
    int owners = owners_override, covered_to = -1;  // send ranges are visited in ascending order of their first row
    if (owners_override <= 0) {
      int order[DIST_MAX_RANKS];
      for (int s = 0; s < d.n_send; ++s) order[s] = s;
      for (int i = 1; i < d.n_send; ++i)
        for (int j = i; j > 0 && d.send_first[order[j]] < d.send_first[order[j - 1]]; --j) {
          const int t = order[j]; order[j] = order[j - 1]; order[j - 1] = t;
        }
      for (int i = 0; i < d.n_send; ++i) {
        const int s = order[i];
        const int c0 = max(d.send_first[s] / per, covered_to + 1), c1 = d.send_last[s] / per;
        if (c1 >= c0) { owners += c1 - c0 + 1; covered_to = c1; }
      }
    }
    fence_acq_rel_sys();  // my peer stores (and, cumulatively, the CTA's) are performed before the ticket is visible
    const unsigned t = atomicAdd(&d.counters[Z ? 4 : 2], 1u);
    if (t == static_cast<unsigned>(owners) - 1u) {
      d.counters[Z ? 4 : 2] = 0u;
      const unsigned count = ++d.counters[Z ? 3 : 1];
      fence_acq_rel_sys();  // the other owners' tickets (hence their stores) are ordered before the flags
      for (int s = 0; s < d.n_send; ++s) {
        unsigned char *peer = d.peer[d.send_peer[s]];
        st_sys_u64(Z ? dist_flag_z(peer, d.rank) : dist_flag(peer, d.rank), count);
      }
    }
  }
}

template <typename T, int NT = 0>
__device__ __forceinline__ void dist_halo_push_signal(const DistDev &d, const T *vec, int lo, int hi, int per) {
  dist_halo_push_impl<T, false, NT>(d, vec, nullptr, lo, hi, per);
}
// z channel: pushes D^-1[i] * r[i] of the halo rows in [lo, hi)
template <typename T>
__device__ __forceinline__ void dist_halo_push_signal_z(const DistDev &d, const T *r, const T *dinv, int lo, int hi, int per,
                                                        int owners_override = 0) {
  dist_halo_push_impl<T, true>(d, r, dinv, lo, hi, per, owners_override);
}

// Called by thread 0 of a CTA before the CTA's first read of a ghost entry; follow with __syncthreads().
__device__ __forceinline__ void dist_halo_wait(const DistDev &d) {
  if (d.world <= 1) return;
  const unsigned want = d.want_ptr ? *reinterpret_cast<const volatile unsigned *>(d.want_ptr) + 1u
                                   : *reinterpret_cast<volatile unsigned *>(&d.counters[1]);
  const long long t0 = clock64();
  for (int r = 0; r < d.n_recv; ++r) {
    // (relaxed polls first: every acquire invalidates the SM's L1, which the other CTAs of the SM are gathering x
    // through, so acquiring polls would do that once per spin)
    const unsigned long long *f = dist_flag(d.self, d.recv_peer[r]);
    while (ld_sys_u64(f) < want) {
      if (clock64() - t0 > DIST_SPIN_LIMIT) {
        atomicExch(d.error, 2);
        break;
      }
    }
  }
  // The polls above are relaxed; the reads of the ghost entries have to be ordered after EVERY peer's release store, and an
  // acquire load of one flag only synchronises with the peer that wrote that flag: so one acquire load per peer (the flags
  // are already at their final values, these loads hit at once).  Measured at 4 GPUs on one box, C2 CG (profiles/
  // r03i_halo_wait_ab.txt): one acquire on peer 0 only (round 2, formally unordered for the other peers) 12 008 it/s, one
  // per peer 11 852 (-1.3 %), a fence.acq_rel.sys after the polls instead 11 191 (-6.8 %).
#if defined(QDSP_HALO_WAIT_ONE_ACQUIRE)   // round-2 behaviour, kept as a build switch for A/B timing only
  if (d.n_recv > 0) (void)ld_acquire_sys_u64(dist_flag(d.self, d.recv_peer[0]));
#else
  for (int r = 0; r < d.n_recv; ++r) (void)ld_acquire_sys_u64(dist_flag(d.self, d.recv_peer[r]));
#endif
}
// The same for the z channel (thread 0 of a CTA, before its first read of a z-ghost; follow with __syncthreads()).
__device__ __forceinline__ void dist_halo_wait_z(const DistDev &d) {
  if (d.world <= 1) return;
  const unsigned want = *reinterpret_cast<volatile unsigned *>(&d.counters[3]);
  const long long t0 = clock64();
  for (int r = 0; r < d.n_recv; ++r) {
    const unsigned long long *f = dist_flag_z(d.self, d.recv_peer[r]);
    while (ld_sys_u64(f) < want) {
      if (clock64() - t0 > DIST_SPIN_LIMIT) {
        atomicExch(d.error, 2);
        break;
      }
    }
  }
#if defined(QDSP_HALO_WAIT_ONE_ACQUIRE)
  if (d.n_recv > 0) (void)ld_acquire_sys_u64(dist_flag_z(d.self, d.recv_peer[0]));
#else
  for (int r = 0; r < d.n_recv; ++r) (void)ld_acquire_sys_u64(dist_flag_z(d.self, d.recv_peer[r]));  // see dist_halo_wait
#endif
}
#endif  // __CUDACC__

}  // namespace qdsp
