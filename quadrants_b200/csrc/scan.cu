// scan.cu -- single-pass inclusive prefix sum with cross-tile descriptor look-back (i32, wrapping), the kernel behind
// qd.algorithms.PrefixSumExecutor.run (python/quadrants/algorithms/_algorithms.py:109-163, which runs 2 blits +
// 5 scan levels + 4 uniform-add levels of 64-thread blocks, python/quadrants/_kernels.py:298-352).
//
// HBM-bound: algorithmic traffic 2*N*4 bytes (read once, write once), in place.
//
// What was measured on B200 (2^28 elements; profiles/r02b_kernel_scan.txt, r02i/j/k_scan_variants.txt):
//   * one CTA per tile, ticket order, Merrill-Garland walk by warp 0 (round 1): 0.54 of the measured HBM peak whatever the
//     tile size (4096 / 16384) and the walk width (32 / 128 / 1024 descriptors per L2 round trip).  ncu: DRAM 45 %, issue
//     slots 15 %, 70 % of the stall samples in the CTA barrier behind the walk.  Two reasons: a CTA has loads in flight
//     only during the first third of its life, and the walk advances ~32 tiles per round trip while ~90 tiles are
//     issued per microsecond -- every poll hits the same two dozen L2 lines that hold the in-flight descriptors.
//   * descriptors padded to 64 bytes (one L2 sector pair each): +36 %.
// Now: PERSISTENT, warp-specialised, round-structured.
//   * G = 2 x SM count CTAs, launched cooperatively (co-resident by contract).  In round r CTA c owns tile r G + c.
//   * a producer thread streams the CTA's tiles into a 3-stage shared-memory ring with TMA bulk copies (cp.async.bulk +
//     mbarrier complete_tx), two tiles ahead: loads are in flight all the time, whatever the consumers wait for.
//   * eight consumer warps copy a landed tile into registers, release the stage, REDUCE it (one shuffle tree) and publish
//     the tile aggregate at once; then they issue the look-back loads and do the full warp scans while those are in flight.
//   * look-back: exclusive prefix = base_r + sum of the aggregates of tiles (r, 0..c-1).  Only AGGREGATES are ever waited
//     for (each published right after a tile is reduced); base_r, the inclusive prefix of round r-1's last tile, is
//     published once per round by the CTA that owns that tile.  No chain of inclusive prefixes: one coalesced read of
//     <= c descriptors by the whole CTA, repeated only while a predecessor of the SAME round has not reduced its tile.
// Fallback (cooperative launch unavailable): the same kernel with ticket order and the Merrill-Garland walk.
#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <string>

#include "lookback.cuh"
#include "tma.cuh"

namespace qdsp {

constexpr int SCAN_CONSUMERS = 256;
constexpr int SCAN_THREADS = SCAN_CONSUMERS + 32;             // + the producer warp
constexpr int SCAN_VEC = 8;                                    // int4 chunks per lane
constexpr int SCAN_TILE = SCAN_CONSUMERS * SCAN_VEC * 4;       // 8192 elements = 32 KB
constexpr int SCAN_WARP_SPAN = 32 * SCAN_VEC * 4;              // 1024 elements per warp
constexpr int SCAN_STAGES = 3;
constexpr int SCAN_MAX_GRID = 2 * SCAN_CONSUMERS;              // round look-back: <= 2 descriptors per consumer thread
constexpr int SCAN_DESC_SHIFT = 3;                             // descriptors 64 bytes apart
constexpr size_t SCAN_HEADER_BYTES = 128;                      // ticket counter, padded
constexpr size_t SCAN_SMEM_BYTES = SCAN_STAGES * SCAN_TILE * sizeof(int) + 128 /*alignment slack*/;
constexpr unsigned SCAN_NO_TILE = 0xffffffffu;

__device__ __forceinline__ void scan_consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(SCAN_CONSUMERS) : "memory"); }
__device__ __forceinline__ unsigned long long *scan_desc(unsigned long long *desc, unsigned tile) {
  return desc + (static_cast<size_t>(tile) << SCAN_DESC_SHIFT);
}

// `data` is the caller's pointer rounded DOWN to 16 bytes; the first head_skip (0..3) slots are not the caller's: they are
// never written and count as zeros (n includes them).  ROUNDS: see the header.
template <bool ROUNDS>
__global__ void __launch_bounds__(SCAN_THREADS, 2)
scan_inclusive_i32_kernel(int *data, long long n, int head_skip, unsigned n_tiles, unsigned *__restrict__ counter,
                          unsigned long long *__restrict__ desc) {
  extern __shared__ unsigned char smem_raw[];
  __shared__ unsigned long long s_full[SCAN_STAGES], s_empty[SCAN_STAGES];
  __shared__ unsigned s_tile[SCAN_STAGES];
  __shared__ unsigned s_warp_total[SCAN_CONSUMERS / 32];
  __shared__ unsigned s_lb_sum[SCAN_CONSUMERS / 32], s_lb_ok[SCAN_CONSUMERS / 32];
  __shared__ unsigned s_prefix;
  int *ring = reinterpret_cast<int *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const int t = threadIdx.x;
  if (t == 0) {
    for (int s = 0; s < SCAN_STAGES; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&s_empty[s], SCAN_CONSUMERS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (t >= SCAN_CONSUMERS) {
    // ---- producer: one thread ----
    if (t != SCAN_CONSUMERS) return;
    const unsigned long long pol = l2_evict_first_policy();
    for (unsigned it = 0;; ++it) {
      const int stage = it % SCAN_STAGES;
      const unsigned phase = (it / SCAN_STAGES) & 1;
      mbar_wait(&s_empty[stage], phase ^ 1u);
      // rounds: tile r G + c.  Fallback: in-order tickets, taken only when a stage is free; a CTA works through its tickets
      // in ascending order, so the tile with the smallest unfinished index can always proceed (no deadlock).
      const unsigned tile = ROUNDS ? it * gridDim.x + blockIdx.x : atomicAdd(counter, 1u);
      if (tile >= n_tiles) {
        s_tile[stage] = SCAN_NO_TILE;
        mbar_arrive(&s_full[stage]);  // release: the consumers see the sentinel
        return;
      }
      s_tile[stage] = tile;
      const long long base = static_cast<long long>(tile) * SCAN_TILE;
      if (base + SCAN_TILE <= n) {
        mbar_expect_tx(&s_full[stage], SCAN_TILE * sizeof(int));
        tma_bulk_g2s(ring + stage * SCAN_TILE, data + base, SCAN_TILE * sizeof(int), &s_full[stage], pol);
      } else {
        mbar_arrive(&s_full[stage]);  // the ragged last tile is read by the consumers themselves
      }
    }
  }

  // ---- consumers ----
  const unsigned lane = t & 31, warp = t >> 5;
  unsigned long long *rbase = scan_desc(desc, n_tiles);  // round bases, 64 bytes apart, behind the tile descriptors
  for (unsigned it = 0;; ++it) {
    const int stage = it % SCAN_STAGES;
    const unsigned phase = (it / SCAN_STAGES) & 1;
    mbar_wait(&s_full[stage], phase);
    const unsigned tile = s_tile[stage];
    if (tile == SCAN_NO_TILE) break;
    const long long warp_base = static_cast<long long>(tile) * SCAN_TILE + warp * SCAN_WARP_SPAN;
    const bool whole = static_cast<long long>(tile + 1) * SCAN_TILE <= n;
    const bool vec_store = whole && !(tile == 0 && head_skip != 0);
    unsigned v[SCAN_VEC][4];
    if (whole) {
      const int *src = ring + stage * SCAN_TILE + warp * SCAN_WARP_SPAN;
#pragma unroll
      for (int j = 0; j < SCAN_VEC; ++j) {
        const int4 q = *reinterpret_cast<const int4 *>(src + j * 128 + lane * 4);
        v[j][0] = q.x; v[j][1] = q.y; v[j][2] = q.z; v[j][3] = q.w;
      }
      if (tile == 0 && head_skip != 0 && t == 0) {  // the foreign slots in front of the caller's array count as zero
#pragma unroll
        for (int e = 0; e < 3; ++e)
          if (e < head_skip) v[0][e] = 0u;
      }
    } else {
#pragma unroll
      for (int j = 0; j < SCAN_VEC; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const long long i = warp_base + j * 128 + lane * 4 + e;
          v[j][e] = (i >= head_skip && i < n) ? static_cast<unsigned>(data[i]) : 0u;
        }
    }
    // the tile is in registers: hand the stage back to the producer
    __syncwarp();
    if (lane == 0) mbar_arrive(&s_empty[stage]);

    // 1. reduce: warp totals -> tile aggregate, published before anything else happens
    {
      unsigned ts = 0u;
#pragma unroll
      for (int j = 0; j < SCAN_VEC; ++j) ts += (v[j][0] + v[j][1]) + (v[j][2] + v[j][3]);
      ts = warp_sum(ts);
      if (lane == 0) s_warp_total[warp] = ts;
    }
    scan_consumer_bar();
    unsigned warp_prefix = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_CONSUMERS / 32; ++w) {
      const unsigned x = s_warp_total[w];
      if (w < static_cast<int>(warp)) warp_prefix += x;
      tile_total += x;
    }
    unsigned exclusive = 0u;
    if constexpr (ROUNDS) {
      const unsigned G = gridDim.x, c = blockIdx.x, r = it;
      if (t == 0) st_relaxed_u64(scan_desc(desc, tile), LB_AGGREGATE | tile_total);
      // 2. look-back loads in flight ...
      unsigned long long d0 = LB_AGGREGATE, d1 = LB_AGGREGATE, db = LB_AGGREGATE;
      auto poll = [&]() {
        if (static_cast<unsigned>(t) < c) d0 = ld_relaxed_u64(scan_desc(desc, r * G + t));
        if (static_cast<unsigned>(t) + SCAN_CONSUMERS < c) d1 = ld_relaxed_u64(scan_desc(desc, r * G + t + SCAN_CONSUMERS));
        if (r > 0 && t == SCAN_CONSUMERS - 1) db = ld_relaxed_u64(rbase + (static_cast<size_t>(r) << 3));
      };
      poll();
      // 3. ... while the warp scans run
      unsigned running = 0;
#pragma unroll
      for (int j = 0; j < SCAN_VEC; ++j) {
        v[j][1] += v[j][0]; v[j][2] += v[j][1]; v[j][3] += v[j][2];
        const unsigned incl = warp_inclusive_scan(v[j][3]);
        const unsigned excl = incl - v[j][3] + running;
#pragma unroll
        for (int e = 0; e < 4; ++e) v[j][e] += excl;
        running += __shfl_sync(FULL, incl, 31);
      }
      // 4. combine; poll again only while an aggregate of this round (or the round base) is still missing
      while (true) {
        const bool ok = (d0 & LB_STATUS_MASK) != LB_INVALID && (d1 & LB_STATUS_MASK) != LB_INVALID &&
                        (db & LB_STATUS_MASK) != LB_INVALID;
        const unsigned sc = warp_sum(static_cast<unsigned>(d0) + static_cast<unsigned>(d1));
        const bool wok = __all_sync(FULL, ok);
        if (lane == 0) s_lb_sum[warp] = sc;
        if (lane == 31) s_lb_ok[warp] = wok ? 1u : 0u;
        if (t == SCAN_CONSUMERS - 1) s_prefix = static_cast<unsigned>(db);
        scan_consumer_bar();
        bool all_ok = true;
        unsigned cur = 0u;
#pragma unroll
        for (int w = 0; w < SCAN_CONSUMERS / 32; ++w) {
          all_ok = all_ok && s_lb_ok[w] != 0u;
          cur += s_lb_sum[w];
        }
        exclusive = s_prefix + cur;
        if (all_ok) break;
        scan_consumer_bar();  // the slots are rewritten by the next poll
        poll();
      }
      if (c == G - 1 && t == 0)  // the last tile of a full round: its inclusive prefix is the next round's base
        st_relaxed_u64(rbase + (static_cast<size_t>(r + 1) << 3), LB_AGGREGATE | static_cast<unsigned>(exclusive + tile_total));
    } else {
      if (warp == 0) {
        const unsigned long long ex = lookback_exclusive(desc, tile, tile_total, SCAN_DESC_SHIFT);
        if (lane == 0) s_prefix = static_cast<unsigned>(ex);
      }
      unsigned running = 0;
#pragma unroll
      for (int j = 0; j < SCAN_VEC; ++j) {
        v[j][1] += v[j][0]; v[j][2] += v[j][1]; v[j][3] += v[j][2];
        const unsigned incl = warp_inclusive_scan(v[j][3]);
        const unsigned excl = incl - v[j][3] + running;
#pragma unroll
        for (int e = 0; e < 4; ++e) v[j][e] += excl;
        running += __shfl_sync(FULL, incl, 31);
      }
      scan_consumer_bar();
      exclusive = s_prefix;
    }
    const unsigned add = exclusive + warp_prefix;
    if (vec_store) {
#pragma unroll
      for (int j = 0; j < SCAN_VEC; ++j) {
        int4 q;
        q.x = v[j][0] + add; q.y = v[j][1] + add; q.z = v[j][2] + add; q.w = v[j][3] + add;
        *reinterpret_cast<int4 *>(data + warp_base + j * 128 + lane * 4) = q;
      }
    } else {
#pragma unroll
      for (int j = 0; j < SCAN_VEC; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const long long i = warp_base + j * 128 + lane * 4 + e;
          if (i >= head_skip && i < n) data[i] = static_cast<int>(v[j][e] + add);
        }
    }
    // (the shared slots are rewritten only after a later consumer barrier of the next tile)
  }
}

size_t scan_tmp_bytes(int64_t n) {
  const int64_t tiles = n > 0 ? (n + 3 + SCAN_TILE - 1) / SCAN_TILE : 1;  // + 3: an unaligned pointer shifts the tiling
  // per tile: one descriptor + room for one round-base word, each 64 bytes apart (rounds <= tiles)
  return SCAN_HEADER_BYTES + (2 * static_cast<size_t>(tiles) + 2) * (sizeof(unsigned long long) << SCAN_DESC_SHIFT);
}

void inclusive_scan_i32(int32_t *inout, int64_t n, void *tmp, size_t tmp_bytes, cudaStream_t stream) {
  QDSP_REQUIRE(n >= 0, QDSP_ERR_INVALID, "scan: negative length");
  if (n == 0) return;
  QDSP_REQUIRE(inout != nullptr && tmp != nullptr, QDSP_ERR_INVALID, "scan: null pointer");
  QDSP_REQUIRE((reinterpret_cast<uintptr_t>(inout) & 3u) == 0, QDSP_ERR_INVALID, "scan: data must be 4-byte aligned");
  // any i32 field is accepted (the reference scans whatever field it is given): a pointer that is not 16-byte aligned is
  // rounded down and the 1-3 foreign slots in front are skipped by tile 0
  int head_skip = static_cast<int>((reinterpret_cast<uintptr_t>(inout) & 15u) / 4u);
  const size_t need = scan_tmp_bytes(n);  // sized for the worst shift
  inout -= head_skip;
  n += head_skip;
  QDSP_REQUIRE(tmp_bytes >= need, QDSP_ERR_INVALID, "scan: tmp buffer too small");
  const int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  QDSP_REQUIRE(tiles < 0x7fffffffLL, QDSP_ERR_INVALID, "scan: too many elements");
  unsigned n_tiles = static_cast<unsigned>(tiles);
  QDSP_CUDA(cudaMemsetAsync(tmp, 0, need, stream));
  auto *counter = static_cast<unsigned *>(tmp);
  auto *desc = reinterpret_cast<unsigned long long *>(static_cast<char *>(tmp) + SCAN_HEADER_BYTES);
  // per-device set-up: opt-in shared-memory size, resident CTAs per SM, cooperative-launch support
  struct DevCfg { int dev = -1, occ = 0, coop = 0; };
  static thread_local DevCfg cfg;
  int dev = 0;
  QDSP_CUDA(cudaGetDevice(&dev));
  if (cfg.dev != dev) {
    QDSP_CUDA(cudaFuncSetAttribute(scan_inclusive_i32_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SCAN_SMEM_BYTES)));
    QDSP_CUDA(cudaFuncSetAttribute(scan_inclusive_i32_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SCAN_SMEM_BYTES)));
    QDSP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cfg.occ, scan_inclusive_i32_kernel<true>, SCAN_THREADS, SCAN_SMEM_BYTES));
    QDSP_CUDA(cudaDeviceGetAttribute(&cfg.coop, cudaDevAttrCooperativeLaunch, dev));
    const char *e = std::getenv("QDSP_SCAN_ROUNDS");  // "0": force the ticket / Merrill-Garland fallback (testing aid)
    if (e && e[0] == '0') cfg.coop = 0;
    cfg.dev = dev;
  }
  QDSP_REQUIRE(cfg.occ >= 1, QDSP_ERR_CUDA, "scan: kernel does not fit an SM");
  const int per_sm = std::min(cfg.occ, 2);
  const int grid = static_cast<int>(std::min<int64_t>(tiles, std::min(device_props().sm_count * per_sm, SCAN_MAX_GRID)));
  if (cfg.coop) {
    // every CTA waits for aggregates of CTAs of the same round, so all of them must be resident: cooperative launch
    // (fails instead of deadlocking if they cannot be)
    void *args[] = {&inout, &n, &head_skip, &n_tiles, &counter, &desc};
    QDSP_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void *>(scan_inclusive_i32_kernel<true>), dim3(static_cast<unsigned>(grid)),
                                          dim3(SCAN_THREADS), args, SCAN_SMEM_BYTES, stream));
  } else {
    scan_inclusive_i32_kernel<false><<<grid, SCAN_THREADS, SCAN_SMEM_BYTES, stream>>>(inout, n, head_skip, n_tiles, counter, desc);
  }
  QDSP_LAUNCH_CHECK();
}

// ---- the same scan for callers that cannot allocate device memory ----------------------------------------------------
// qdsp_inclusive_scan_i32_static: the tile descriptors and the ticket counter live in a fixed array that is part of the
// library image instead of in a buffer the caller brings.  Meant for hosts that own no allocator of their own to draw a
// tmp buffer from and must not call cudaMalloc either: an embedding whose device memory was handed, whole, to a
// pre-sized arena at start-up (the reference's runtime works like that: one pool, ndarrays carved out of it), or a
// fixed-rate stepping loop in which cudaMalloc / cudaFree -- which stall every stream of the device -- are banned.
// Cost: SCAN_STATIC_BYTES (about 1 MB) of every GPU that loads libqdsp.so, held until the process exits, and a length
// limit of SCAN_STATIC_MAX_N elements.
constexpr int64_t SCAN_STATIC_MAX_N = int64_t(1) << 26;
constexpr size_t SCAN_STATIC_BYTES =
    SCAN_HEADER_BYTES + (2 * static_cast<size_t>((SCAN_STATIC_MAX_N + 3 + SCAN_TILE - 1) / SCAN_TILE) + 2) *
                            (sizeof(unsigned long long) << SCAN_DESC_SHIFT);
static __device__ __align__(256) unsigned char g_scan_static_scratch[SCAN_STATIC_BYTES];

int64_t scan_static_max_n() { return SCAN_STATIC_MAX_N; }

void inclusive_scan_i32_static(int32_t *inout, int64_t n, cudaStream_t stream) {
  QDSP_REQUIRE(n >= 0, QDSP_ERR_INVALID, "scan (static scratch): negative length");
  QDSP_REQUIRE(n <= SCAN_STATIC_MAX_N && scan_tmp_bytes(n) <= SCAN_STATIC_BYTES, QDSP_ERR_INVALID,
               "scan (static scratch): " + std::to_string(n) + " elements exceed the built-in scratch array, which covers " +
                   std::to_string(SCAN_STATIC_MAX_N) + "; use qdsp_inclusive_scan_i32 with a caller-owned tmp buffer");
  if (n == 0) return;
  // In host code `g_scan_static_scratch` is only a handle for the runtime, not an address a kernel may be given.  The
  // real address differs per device (each GPU has its own copy of the array), so it is looked up once per device, with
  // that device current, on the first call made there.
  constexpr int MAX_DEVICES = 64;
  static std::mutex mu;
  static void *dev_ptr[MAX_DEVICES] = {};
  int dev = 0;
  QDSP_CUDA(cudaGetDevice(&dev));
  QDSP_REQUIRE(dev >= 0 && dev < MAX_DEVICES, QDSP_ERR_UNSUPPORTED, "scan (static scratch): device ordinal above 63");
  void *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev_ptr[dev] == nullptr) {
      void *p = nullptr;
      QDSP_CUDA(cudaGetSymbolAddress(&p, g_scan_static_scratch));
      QDSP_REQUIRE(p != nullptr, QDSP_ERR_CUDA, "scan (static scratch): the scratch array has no device address");
      dev_ptr[dev] = p;
    }
    scratch = dev_ptr[dev];
  }
  // From here on this is the ordinary scan: it zeroes the part of the scratch it is about to use with a stream-ordered
  // memset (the array still holds the previous call's descriptors) and launches the same kernel.  Nothing on this path
  // allocates or frees device memory.
  inclusive_scan_i32(inout, n, scratch, SCAN_STATIC_BYTES, stream);
}

}  // namespace qdsp
