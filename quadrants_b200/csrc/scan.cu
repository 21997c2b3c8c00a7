// scan.cu -- single-pass decoupled look-back inclusive prefix sum (i32, wrapping), the kernel behind
// qd.algorithms.PrefixSumExecutor.run (python/quadrants/algorithms/_algorithms.py:109-163, which runs 2 blits +
// 5 scan levels + 4 uniform-add levels of 64-thread blocks, python/quadrants/_kernels.py:298-352).
//
// HBM-bound: algorithmic traffic 2*N*4 bytes (read once, write once).  One CTA = one 16384-element tile:
// 128-bit coalesced loads (warp w owns 1024 consecutive elements, 8 x int4 per lane), warp-shuffle scans, one
// look-back per tile by warp 0, 128-bit stores.  In place.
#include "lookback.cuh"

namespace qdsp {

// Tile size: the look-back is a serial chain -- one L2 round trip (~0.4 us) lets a tile look at 32 predecessors, so the
// prefix front advances by at most ~32 tiles per round trip whatever the occupancy.  At 4096-element tiles that caps
// the kernel at ~2.8 TB/s (measured, profiles/r01d); 16384-element tiles move 4x the bytes per chain step.
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_VEC = 8;                                    // int4 chunks per lane
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_VEC * 4;         // 16384
constexpr int SCAN_WARP_SPAN = 32 * SCAN_VEC * 4;              // 1024 elements per warp
constexpr size_t SCAN_HEADER_BYTES = 128;                      // ticket counter, padded

__global__ void __launch_bounds__(SCAN_THREADS)
scan_inclusive_i32_kernel(int *__restrict__ data, long long n, unsigned *__restrict__ counter,
                          unsigned long long *__restrict__ desc) {
  __shared__ unsigned s_ticket;
  __shared__ unsigned s_warp_total[SCAN_THREADS / 32];
  __shared__ unsigned s_tile_prefix;
  const unsigned tile = take_ticket(counter, &s_ticket);
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  const long long warp_base = static_cast<long long>(tile) * SCAN_TILE + warp * SCAN_WARP_SPAN;
  const bool full = static_cast<long long>(tile + 1) * SCAN_TILE <= n;

  unsigned v[SCAN_VEC][4];
  if (full) {
#pragma unroll
    for (int j = 0; j < SCAN_VEC; ++j) {
      const int4 q = *reinterpret_cast<const int4 *>(data + warp_base + j * 128 + lane * 4);
      v[j][0] = q.x; v[j][1] = q.y; v[j][2] = q.z; v[j][3] = q.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < SCAN_VEC; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const long long i = warp_base + j * 128 + lane * 4 + e;
        v[j][e] = i < n ? static_cast<unsigned>(data[i]) : 0u;
      }
  }
  // warp-local inclusive scan of the warp's span in SCAN_VEC rounds of (4 per lane x 32 lanes)
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
  if (lane == 31) s_warp_total[warp] = running;
  __syncthreads();
  unsigned warp_prefix = 0, tile_total = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    const unsigned t = s_warp_total[w];
    if (w < (int)warp) warp_prefix += t;
    tile_total += t;
  }
  if (warp == 0) {
    const unsigned long long ex = lookback_exclusive(desc, tile, tile_total);
    if (lane == 0) s_tile_prefix = static_cast<unsigned>(ex);
  }
  __syncthreads();
  const unsigned add = s_tile_prefix + warp_prefix;
  if (full) {
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
        if (i < n) data[i] = static_cast<int>(v[j][e] + add);
      }
  }
}

size_t scan_tmp_bytes(int64_t n) {
  const int64_t tiles = n > 0 ? (n + SCAN_TILE - 1) / SCAN_TILE : 1;
  return SCAN_HEADER_BYTES + static_cast<size_t>(tiles) * sizeof(unsigned long long);
}

void inclusive_scan_i32(int32_t *inout, int64_t n, void *tmp, size_t tmp_bytes, cudaStream_t stream) {
  QDSP_REQUIRE(n >= 0, QDSP_ERR_INVALID, "scan: negative length");
  if (n == 0) return;
  QDSP_REQUIRE(inout != nullptr && tmp != nullptr, QDSP_ERR_INVALID, "scan: null pointer");
  QDSP_REQUIRE((reinterpret_cast<uintptr_t>(inout) & 15u) == 0, QDSP_ERR_INVALID, "scan: data must be 16-byte aligned");
  const size_t need = scan_tmp_bytes(n);
  QDSP_REQUIRE(tmp_bytes >= need, QDSP_ERR_INVALID, "scan: tmp buffer too small");
  const int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  QDSP_CUDA(cudaMemsetAsync(tmp, 0, need, stream));
  auto *counter = static_cast<unsigned *>(tmp);
  auto *desc = reinterpret_cast<unsigned long long *>(static_cast<char *>(tmp) + SCAN_HEADER_BYTES);
  scan_inclusive_i32_kernel<<<static_cast<unsigned>(tiles), SCAN_THREADS, 0, stream>>>(inout, n, counter, desc);
  QDSP_LAUNCH_CHECK();
}

}  // namespace qdsp
