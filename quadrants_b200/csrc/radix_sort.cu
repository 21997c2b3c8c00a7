// radix_sort.cu -- "onesweep" least-significant-digit radix sort (Adinets & Merrill): one histogram pass over the keys,
// then ONE read + ONE write of keys/values per 8-bit digit, the per-digit global offsets being resolved inside the
// sweep by decoupled look-back instead of a separate scan pass.  Stable.
//
// Replaces qd.algorithms.parallel_sort (python/quadrants/algorithms/_algorithms.py:19-56 +
// python/quadrants/_kernels.py:227-253: a Batcher odd-even merge network, 406 launches + sync() at N = 2^28) and the
// cusparseXcoosortByRow + cusparseGather of CuSparseMatrix::build_csr_from_coo
// (quadrants/program/sparse_matrix.cpp:358-368).
//
// Byte model (the roofline it is measured against): N*kb + P*2*N*(kb+vb) bytes for P digit passes.
// Tile = 256 threads x ITEMS keys, warp-striped so that (warp, item, lane) order == input order:
//   1. coalesced loads of keys (+values) into registers;
//   2. per-warp ranking: the lanes holding the same digit find each other with one ballot per digit bit and elect a
//      leader that bumps the warp's private digit counter in shared memory -> rank inside the warp, stable by
//      construction;
//   3. thread d (256 threads == 256 digits): scan of the 8 warp counters, publish the tile count of digit d;
//   4. scatter keys (+values) to shared memory in tile-sorted order;
//   5. thread d looks back over the preceding tiles' descriptors for digit d, publishes the inclusive prefix;
//   6. write out so that consecutive threads store runs of consecutive global addresses.
#include <cstdlib>

#include "common.cuh"

namespace qdsp {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_RADIX = 256;
constexpr size_t RS_HEADER_BYTES = 256;  // per-pass ticket counters (8 passes max)
constexpr unsigned RS_FLAG_AGG = 1u << 30, RS_FLAG_PREFIX = 2u << 30, RS_FLAG_MASK = 3u << 30, RS_VAL_MASK = ~RS_FLAG_MASK;

struct NoValue {};  // keys-only instantiation

template <typename KeyT>
__device__ __forceinline__ KeyT twiddle_in(KeyT k, int kind) {
  if constexpr (sizeof(KeyT) == 4) {
    if (kind == QDSP_KEY_I32) return k ^ 0x80000000u;
    if (kind == QDSP_KEY_F32) return (k & 0x80000000u) ? ~k : (k | 0x80000000u);
  } else {
    constexpr KeyT SIGN = KeyT(1) << 63;
    if (kind == QDSP_KEY_I64) return k ^ SIGN;
    if (kind == QDSP_KEY_F64) return (k & SIGN) ? ~k : (k | SIGN);
  }
  return k;
}
template <typename KeyT>
__device__ __forceinline__ KeyT twiddle_out(KeyT k, int kind) {
  if constexpr (sizeof(KeyT) == 4) {
    if (kind == QDSP_KEY_I32) return k ^ 0x80000000u;
    if (kind == QDSP_KEY_F32) return (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  } else {
    constexpr KeyT SIGN = KeyT(1) << 63;
    if (kind == QDSP_KEY_I64) return k ^ SIGN;
    if (kind == QDSP_KEY_F64) return (k & SIGN) ? (k & ~SIGN) : ~k;
  }
  return k;
}

// live count: min(n, *n_dev) when the caller keeps the count on the device (qd.algorithms.sort's 0-d `n`)
__device__ __forceinline__ long long rs_count(long long n, const int *n_dev) {
  if (n_dev == nullptr) return n;
  const long long m = *n_dev;
  return m < 0 ? 0 : (m < n ? m : n);
}

// ---- pass 0: digit histograms of every pass in one read of the keys -----------------------------------------------
template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS)
rs_histogram_kernel(const KeyT *__restrict__ keys, long long n, const int *__restrict__ n_dev, int kind, int begin_bit,
                    int end_bit, int passes, unsigned *__restrict__ hist /* passes x 256 */) {
  extern __shared__ unsigned s_hist[];  // passes x 256
  n = rs_count(n, n_dev);
  for (int i = threadIdx.x; i < passes * RS_RADIX; i += RS_THREADS) s_hist[i] = 0;
  __syncthreads();
  const long long stride = static_cast<long long>(gridDim.x) * RS_THREADS;
  for (long long i = static_cast<long long>(blockIdx.x) * RS_THREADS + threadIdx.x; i < n; i += stride) {
    const KeyT k = twiddle_in(keys[i], kind);
#pragma unroll 1
    for (int p = 0; p < passes; ++p) {
      const int shift = begin_bit + 8 * p;
      const int nb = min(8, end_bit - shift);
      const unsigned d = static_cast<unsigned>(k >> shift) & ((1u << nb) - 1u);
      atomicAdd(&s_hist[p * RS_RADIX + d], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * RS_RADIX; i += RS_THREADS)
    if (s_hist[i]) atomicAdd(&hist[i], s_hist[i]);
}

// exclusive scan of each pass's 256 bins (one CTA per pass)
__global__ void __launch_bounds__(RS_RADIX) rs_scan_bins_kernel(unsigned *__restrict__ hist) {
  __shared__ unsigned s_w[RS_RADIX / 32];
  unsigned *h = hist + blockIdx.x * RS_RADIX;
  const unsigned v = h[threadIdx.x];
  const unsigned incl = warp_inclusive_scan(v);
  if (lane_id() == 31) s_w[threadIdx.x >> 5] = incl;
  __syncthreads();
  unsigned base = 0;
  for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) base += s_w[w];
  h[threadIdx.x] = base + incl - v;
}

// ---- one digit pass -------------------------------------------------------------------------------------------------
template <typename KeyT, typename ValT, int ITEMS>
struct RsSmem {
  unsigned warp_hist[RS_WARPS][RS_RADIX];
  unsigned digit_start[RS_RADIX];  // first tile-sorted position of each digit
  unsigned gbase[RS_RADIX];        // global output index of tile-sorted position p of digit d is gbase[d] + p
  unsigned warp_tot[RS_WARPS];
  unsigned ticket;
  KeyT keys[RS_THREADS * ITEMS];
};

// the staged values follow the struct, 16-byte aligned (8-byte values after a 4-byte-aligned struct would fault)
template <typename KeyT, typename ValT, int ITEMS>
__host__ __device__ constexpr size_t rs_vals_offset() {
  return (sizeof(RsSmem<KeyT, ValT, ITEMS>) + 15) / 16 * 16;
}

// peers &= (lanes whose digit bit B equals mine): and+setp -> one predicate-setting LOP3, one VOTE, one SEL, one LOP3
// (the C++ form `peers &= bit ? m : ~m` compiles to six instructions per bit)
template <int B>
__device__ __forceinline__ void rs_peer_step(unsigned &peers, unsigned d) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b32 t, m, s;\n\t"
      "and.b32 t, %1, %2;\n\t"
      "setp.ne.u32 p, t, 0;\n\t"
      "vote.sync.ballot.b32 m, p, 0xffffffff;\n\t"
      "selp.b32 s, 0, 0xffffffff, p;\n\t"
      "xor.b32 m, m, s;\n\t"
      "and.b32 %0, %0, m;\n\t"
      "}"
      : "+r"(peers)
      : "r"(d), "n"(1u << B));
}

// ---- one digit pass ------------------------------------------------------------------------------------------------
// The sweep is issue- and latency-bound, not DRAM-bound (profiles/r01n_kernel_onesweep.txt: ~150 thread instructions per
// key at 50 % issue utilisation, DRAM at 21 %, a third of the stall samples in the look-back spin).  Hence:
//  * full tiles (all but the last) run a body without any per-item bounds test;
//  * the peer mask of a digit is built by rs_peer_step (3.5 SASS instructions per bit);
//  * the look-back runs after the shared-memory scatter, eight predecessors per round trip;
//  * the global output index of each tile-sorted position is computed once and reused for the value.
// Measured steps: profiles/r01_sort_variants.jsonl.

// Look-back of one digit column: `my` = this tile's descriptor (aggregate already published); returns the exclusive
// prefix and publishes the inclusive one.  The RS_LB_BATCH loads of a round are independent (one L2 latency for all);
// a descriptor that is not published yet ends the round and is polled again at the head of the next one.
constexpr int RS_LB_BATCH = 8;
__device__ __forceinline__ unsigned rs_lookback_batched(unsigned *my, unsigned tile, unsigned count) {
  if (tile == 0) return 0u;
  unsigned exclusive = 0;
  long long t = static_cast<long long>(tile) - 1;  // next predecessor to consume
  const unsigned *col = my - static_cast<size_t>(tile) * RS_RADIX;  // descriptor of tile 0 for this digit
  bool done = false;
  while (!done) {
    unsigned v[RS_LB_BATCH];
#pragma unroll
    for (int j = 0; j < RS_LB_BATCH; ++j)
      v[j] = (t - j >= 0) ? ld_relaxed_u32(col + static_cast<size_t>(t - j) * RS_RADIX) : RS_FLAG_PREFIX;
    int used = 0;
#pragma unroll
    for (int j = 0; j < RS_LB_BATCH; ++j) {
      if (!done && used == j) {
        const unsigned f = v[j] & RS_FLAG_MASK;
        if (f != 0u) {
          exclusive += v[j] & RS_VAL_MASK;
          used = j + 1;
          if (f == RS_FLAG_PREFIX) done = true;
        }
      }
    }
    t -= used;
  }
  st_relaxed_u32(my, RS_FLAG_PREFIX | ((exclusive + count) & RS_VAL_MASK));
  return exclusive;
}

// keys (+values) of one tile into registers, warp-striped, straight from global memory
template <typename KeyT, typename ValT, int ITEMS, bool FULL_TILE>
__device__ __forceinline__ void rs_load_tile(KeyT (&key)[ITEMS], ValT (&val)[ITEMS], const KeyT *__restrict__ kin,
                                             const ValT *__restrict__ vin, long long tile_base, int valid_n, int kind,
                                             int tw_in) {
  constexpr bool HAS_VAL = !std::is_same<ValT, NoValue>::value;
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  const int wbase = warp * 32 * ITEMS + lane;
  const KeyT *kp = kin + tile_base + wbase;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    if (FULL_TILE || wbase + i * 32 < valid_n) {
      const KeyT k = kp[i * 32];
      key[i] = tw_in ? twiddle_in(k, kind) : k;
    } else {
      key[i] = static_cast<KeyT>(~KeyT(0));
    }
  }
  if constexpr (HAS_VAL) {
    const ValT *vp = vin + tile_base + wbase;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i)
      if (FULL_TILE || wbase + i * 32 < valid_n) val[i] = vp[i * 32];
  }
}

// everything after the load: ranking, digit offsets, shared-memory scatter, look-back, coalesced write-out
template <typename KeyT, typename ValT, int ITEMS, bool FULL_TILE>
__device__ __forceinline__ void rs_sweep_tile(RsSmem<KeyT, ValT, ITEMS> &S, ValT *s_vals, KeyT (&key)[ITEMS],
                                              ValT (&val)[ITEMS], KeyT *__restrict__ kout, ValT *__restrict__ vout,
                                              int valid_n, int shift, unsigned dmask, int kind, int tw_out,
                                              unsigned tile, unsigned *__restrict__ lookback,
                                              const unsigned *__restrict__ bins) {
  constexpr bool HAS_VAL = !std::is_same<ValT, NoValue>::value;
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  const int wbase = warp * 32 * ITEMS + lane;
  unsigned *my_hist = &S.warp_hist[warp][0];
  unsigned rank[ITEMS];
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const bool ok = FULL_TILE || (wbase + i * 32 < valid_n);
    const unsigned d = static_cast<unsigned>(key[i] >> shift) & dmask;
    unsigned peers = FULL_TILE ? FULL : __ballot_sync(FULL, ok);
    rs_peer_step<0>(peers, d);
    rs_peer_step<1>(peers, d);
    rs_peer_step<2>(peers, d);
    rs_peer_step<3>(peers, d);
    rs_peer_step<4>(peers, d);
    rs_peer_step<5>(peers, d);
    rs_peer_step<6>(peers, d);
    rs_peer_step<7>(peers, d);
    const unsigned leader = __ffs(peers) - 1;
    unsigned old = 0;
    if (lane == leader && ok) {
      old = my_hist[d];
      my_hist[d] = old + __popc(peers);
    }
    old = __shfl_sync(FULL, old, leader);
    rank[i] = old + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  // thread d owns digit d: tile count, publication of the aggregate, tile-local digit offsets
  {
    unsigned count, start;
    unsigned *my = lookback + static_cast<size_t>(tile) * RS_RADIX + threadIdx.x;
    const unsigned d = threadIdx.x;
    unsigned sum = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
      const unsigned t = S.warp_hist[w][d];
      S.warp_hist[w][d] = sum;
      sum += t;
    }
    count = sum;
    st_relaxed_u32(my, (tile == 0 ? RS_FLAG_PREFIX : RS_FLAG_AGG) | count);
    const unsigned incl = warp_inclusive_scan(count);
    if (lane == 31) S.warp_tot[warp] = incl;
    __syncthreads();
    unsigned wpre = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w)
      if (w < (int)warp) wpre += S.warp_tot[w];
    start = wpre + incl - count;
    S.digit_start[d] = start;
    S.gbase[d] = count;  // parked in shared memory across the scatter (registers are at their cap there)
  }
  __syncthreads();
  unsigned pos[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const unsigned d = static_cast<unsigned>(key[i] >> shift) & dmask;
    pos[i] = S.digit_start[d] + my_hist[d] + rank[i];
    if (FULL_TILE || wbase + i * 32 < valid_n) S.keys[pos[i]] = key[i];
  }
  if constexpr (HAS_VAL) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i)
      if (FULL_TILE || wbase + i * 32 < valid_n) s_vals[pos[i]] = val[i];
  }
  {
    // the look-back runs only now: the predecessors had the whole shared-memory scatter to publish their counts, so
    // far fewer polls are spent spinning (the spin loop was 22 % of the stall samples and 20 % of the executed
    // instructions, profiles/r01n_kernel_onesweep.txt), and eight predecessors are fetched per round trip
    const unsigned cnt = S.gbase[threadIdx.x];
    const unsigned ex = rs_lookback_batched(lookback + static_cast<size_t>(tile) * RS_RADIX + threadIdx.x, tile, cnt);
    S.gbase[threadIdx.x] = bins[threadIdx.x] + ex - S.digit_start[threadIdx.x];
  }
  __syncthreads();
  unsigned out_idx[ITEMS];  // global index of tile-sorted position threadIdx.x + i * 256
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int p = threadIdx.x + i * RS_THREADS;
    if (FULL_TILE || p < valid_n) {
      const KeyT k = S.keys[p];
      const unsigned d = static_cast<unsigned>(k >> shift) & dmask;
      out_idx[i] = S.gbase[d] + p;
      kout[out_idx[i]] = tw_out ? twiddle_out(k, kind) : k;
    }
  }
  if constexpr (HAS_VAL) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int p = threadIdx.x + i * RS_THREADS;
      if (FULL_TILE || p < valid_n) vout[out_idx[i]] = s_vals[p];
    }
  }
}

template <typename KeyT, typename ValT, int ITEMS, int MINB>
__global__ void __launch_bounds__(RS_THREADS, MINB)
rs_onesweep_kernel(const KeyT *__restrict__ kin, KeyT *__restrict__ kout, const ValT *__restrict__ vin,
                      ValT *__restrict__ vout, long long n, const int *__restrict__ n_dev, int shift, int nbits,
                      int kind, int tw_in, int tw_out, unsigned *__restrict__ ticket,
                      unsigned *__restrict__ lookback /* tiles x 256 */, const unsigned *__restrict__ bins) {
  constexpr int TILE = RS_THREADS * ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  auto &S = *reinterpret_cast<RsSmem<KeyT, ValT, ITEMS> *>(s_raw);
  ValT *s_vals = reinterpret_cast<ValT *>(s_raw + rs_vals_offset<KeyT, ValT, ITEMS>());
  if (threadIdx.x == 0) S.ticket = atomicAdd(ticket, 1u);
  for (int i = threadIdx.x; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&S.warp_hist[0][0])[i] = 0;
  __syncthreads();
  const unsigned tile = S.ticket;
  const long long tile_base = static_cast<long long>(tile) * TILE;
  n = rs_count(n, n_dev);
  if (tile_base >= n) return;
  const int valid_n = static_cast<int>(min(static_cast<long long>(TILE), n - tile_base));
  const unsigned dmask = (1u << nbits) - 1u;
  KeyT key[ITEMS];
  ValT val[ITEMS];
  if (valid_n == TILE) {
    rs_load_tile<KeyT, ValT, ITEMS, true>(key, val, kin, vin, tile_base, valid_n, kind, tw_in);
    rs_sweep_tile<KeyT, ValT, ITEMS, true>(S, s_vals, key, val, kout, vout, valid_n, shift, dmask, kind, tw_out, tile,
                                           lookback, bins);
  } else {
    rs_load_tile<KeyT, ValT, ITEMS, false>(key, val, kin, vin, tile_base, valid_n, kind, tw_in);
    rs_sweep_tile<KeyT, ValT, ITEMS, false>(S, s_vals, key, val, kout, vout, valid_n, shift, dmask, kind, tw_out, tile,
                                            lookback, bins);
  }
}

// A persistent form of this kernel (each CTA loops over tiles; the next tile's keys and values arrive by TMA bulk copy
// into a landing buffer while the current one is ranked and scattered) was built and measured in round 2 and REMOVED:
// with a static b, b+G, ... tile assignment all G tiles of a round reach their look-back together and walk ~G
// predecessors (2^28 pairs 7.77 ms, keys only 7.07 ms against 7.20 / 5.72 ms for this kernel); with tiles handed out by
// ticket one ahead it was 7.59 / 8.01 ms.  The landing buffer costs a CTA per SM for pairs, and the extra shared-memory
// round trip + barrier per tile outweigh the hidden load latency, which the four resident CTAs already overlap.

// ---- host driver ---------------------------------------------------------------------------------------------------
// Keys per thread (tile = 256 x ITEMS) and resident CTAs per SM the kernel is compiled for.  Wider items -> fewer
// keys per thread so that key + value + rank registers stay under the cap.  QDSP_RS_ITEMS=8|12|16|20|24 overrides the choice
// for 4-byte keys with 0/4-byte values (measurement aid; read once).
static int rs_items(int kb, int vb) {
  static const int forced = [] {
    const char *e = std::getenv("QDSP_RS_ITEMS");
    const int v = e ? std::atoi(e) : 0;
    return (v == 8 || v == 12 || v == 16 || v == 20 || v == 24) ? v : 0;
  }();
  if (kb == 4) {
    if (vb == 8) return 12;
    return forced ? forced : 16;
  }
  return vb == 8 ? 8 : 12;
}
static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// aux = [tickets (8 passes) | histograms 8 x 256 | look-back words tiles x 256]
struct RsLayout {
  size_t hist_off, lookback_off, aux_bytes, keys_off, vals_off, total;
  int64_t tiles;
};
static RsLayout rs_layout(int64_t n, int kb, int vb) {
  RsLayout L{};
  const int tile = RS_THREADS * rs_items(kb, vb);
  L.tiles = n > 0 ? (n + tile - 1) / tile : 1;
  L.hist_off = RS_HEADER_BYTES;
  L.lookback_off = align_up(L.hist_off + 8 * RS_RADIX * sizeof(unsigned), 256);
  L.aux_bytes = align_up(L.lookback_off + static_cast<size_t>(L.tiles) * RS_RADIX * sizeof(unsigned), 256);
  L.keys_off = L.aux_bytes;
  L.vals_off = align_up(L.keys_off + static_cast<size_t>(n) * kb, 256);
  L.total = align_up(L.vals_off + static_cast<size_t>(n) * vb, 256);
  return L;
}

size_t sort_tmp_bytes(int64_t n, int key_bytes, int value_bytes) { return rs_layout(n, key_bytes, value_bytes).total; }
// aux only (the caller brings tmp_keys / tmp_values): sized for the smallest tile of this key width, any value width
// (8 keys per thread is the smallest tile ANY dispatch of rs_run can pick, QDSP_RS_ITEMS included, so the look-back area
// checked against the caller's scratch never undercounts the tiles a sweep will use)
size_t sort_modern_scratch_bytes(int64_t capacity, int key_bytes) {
  (void)key_bytes;
  const int64_t tile = RS_THREADS * 8;
  const int64_t tiles = capacity > 0 ? (capacity + tile - 1) / tile : 1;
  const size_t lookback_off = align_up(RS_HEADER_BYTES + 8 * RS_RADIX * sizeof(unsigned), 256);
  return align_up(lookback_off + static_cast<size_t>(tiles) * RS_RADIX * sizeof(unsigned), 256);
}

// Runs the digit passes of bits [begin_bit, end_bit) ping-ponging between (keys, vals) and (kalt, valt); `n` sizes the
// grids, the live count is min(n, *n_dev) when n_dev is given.  Returns where the sorted keys live.
template <typename KeyT, typename ValT, int ITEMS, int MINB>
static void *rs_run_items(KeyT *keys, ValT *vals, KeyT *kalt, ValT *valt, void *aux, int64_t n, const int *n_dev,
                          int kind, int begin_bit, int end_bit, cudaStream_t stream, void **vals_result) {
  constexpr int KB = sizeof(KeyT);
  constexpr int VB = std::is_same<ValT, NoValue>::value ? 0 : sizeof(ValT);
  constexpr int TILE = RS_THREADS * ITEMS;
  const RsLayout L = rs_layout(n, KB, VB);
  const int passes = (end_bit - begin_bit + 7) / 8;
  char *base = static_cast<char *>(aux);
  unsigned *tickets = reinterpret_cast<unsigned *>(base);
  unsigned *hist = reinterpret_cast<unsigned *>(base + L.hist_off);
  unsigned *lookback = reinterpret_cast<unsigned *>(base + L.lookback_off);
  const size_t lookback_bytes = static_cast<size_t>(L.tiles) * RS_RADIX * sizeof(unsigned);

  QDSP_CUDA(cudaMemsetAsync(base, 0, L.lookback_off, stream));  // tickets + histograms
  const int hist_grid = static_cast<int>(std::min<int64_t>((n + RS_THREADS - 1) / RS_THREADS,
                                                          static_cast<int64_t>(device_props().sm_count) * 8));
  rs_histogram_kernel<KeyT><<<hist_grid, RS_THREADS, passes * RS_RADIX * sizeof(unsigned), stream>>>(
      keys, n, n_dev, kind, begin_bit, end_bit, passes, hist);
  QDSP_LAUNCH_CHECK();
  rs_scan_bins_kernel<<<passes, RS_RADIX, 0, stream>>>(hist);
  QDSP_LAUNCH_CHECK();

  auto kernel = rs_onesweep_kernel<KeyT, ValT, ITEMS, MINB>;
  const size_t smem = rs_vals_offset<KeyT, ValT, ITEMS>() + static_cast<size_t>(TILE) * (VB ? VB : 0);
  static std::atomic<unsigned> attr_set{0};  // per instantiation; one bit per device (function attributes are per device)
  const unsigned dev_bit = 1u << (device_props().device & 31);
  if (!(attr_set.load(std::memory_order_acquire) & dev_bit)) {
    QDSP_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    attr_set.fetch_or(dev_bit, std::memory_order_release);
  }
  KeyT *kin = keys, *kout = kalt;
  ValT *vin = vals, *vout = valt;
  for (int p = 0; p < passes; ++p) {
    const int shift = begin_bit + 8 * p;
    const int nbits = std::min(8, end_bit - shift);
    QDSP_CUDA(cudaMemsetAsync(lookback, 0, lookback_bytes, stream));
    kernel<<<static_cast<unsigned>(L.tiles), RS_THREADS, smem, stream>>>(
        kin, kout, vin, vout, n, n_dev, shift, nbits, kind, p == 0, p == passes - 1, tickets + p, lookback,
        hist + p * RS_RADIX);
    QDSP_LAUNCH_CHECK();
    std::swap(kin, kout);
    std::swap(vin, vout);
  }
  if (vals_result) *vals_result = vin;
  return kin;  // the caller's buffer if `passes` is even
}

template <typename KeyT, typename ValT>
static void *rs_run(KeyT *keys, ValT *vals, KeyT *kalt, ValT *valt, void *aux, int64_t n, const int *n_dev, int kind,
                    int begin_bit, int end_bit, cudaStream_t stream, void **vals_result) {
  constexpr int KB = sizeof(KeyT);
  constexpr int VB = std::is_same<ValT, NoValue>::value ? 0 : sizeof(ValT);
#define QDSP_RS_GO(ITEMS, MINB)                                                                                   \
  return rs_run_items<KeyT, ValT, ITEMS, MINB>(keys, vals, kalt, valt, aux, n, n_dev, kind, begin_bit, end_bit, \
                                               stream, vals_result)
  const int items = rs_items(KB, VB);
  if constexpr (KB == 4 && VB <= 4) {
    if (items == 8) QDSP_RS_GO(8, 6);
    if (items == 12) QDSP_RS_GO(12, 5);
    if (items == 20) QDSP_RS_GO(20, 3);  // larger tiles at 3 CTAs per SM: not measured yet (time rose when tiles
    if (items == 24) QDSP_RS_GO(24, 3);  // shrank, profiles/r01_sort_variants.jsonl, so the other direction is open)
    QDSP_RS_GO(16, 4);
  } else if constexpr (KB == 8 && VB == 8) {
    QDSP_RS_GO(8, 4);
  } else {
    QDSP_RS_GO(12, 4);
  }
#undef QDSP_RS_GO
}

// old single-scratch form: the alternate buffers live behind the aux area of `tmp`
template <typename KeyT, typename ValT>
static void *rs_run_tmp(KeyT *keys, ValT *vals, int64_t n, int kind, int begin_bit, int end_bit, void *tmp,
                        cudaStream_t stream, void **vals_result) {
  constexpr int KB = sizeof(KeyT);
  constexpr int VB = std::is_same<ValT, NoValue>::value ? 0 : sizeof(ValT);
  const RsLayout L = rs_layout(n, KB, VB);
  char *base = static_cast<char *>(tmp);
  return rs_run<KeyT, ValT>(keys, vals, reinterpret_cast<KeyT *>(base + L.keys_off),
                            reinterpret_cast<ValT *>(base + L.vals_off), tmp, n, nullptr, kind, begin_bit, end_bit,
                            stream, vals_result);
}

// Sorts in place: if the pass count is odd the result is copied back from the scratch buffer.
template <typename KeyT, typename ValT>
static void rs_sort_inplace(KeyT *keys, ValT *vals, int64_t n, int kind, int begin_bit, int end_bit, void *tmp,
                            cudaStream_t stream) {
  void *vres = nullptr;
  void *kres = rs_run_tmp<KeyT, ValT>(keys, vals, n, kind, begin_bit, end_bit, tmp, stream, &vres);
  if (kres != keys) {
    QDSP_CUDA(cudaMemcpyAsync(keys, kres, static_cast<size_t>(n) * sizeof(KeyT), cudaMemcpyDeviceToDevice, stream));
    if constexpr (!std::is_same<ValT, NoValue>::value)
      QDSP_CUDA(cudaMemcpyAsync(vals, vres, static_cast<size_t>(n) * sizeof(ValT), cudaMemcpyDeviceToDevice, stream));
  }
}

static int key_bytes_of(int kind) { return kind >= QDSP_KEY_U64 ? 8 : 4; }

// Public entry: in-place ascending stable sort of `keys` (and `values`).
void sort_pairs(void *keys, void *values, int64_t n, int key_kind, int value_bytes, void *tmp, size_t tmp_bytes,
                cudaStream_t stream) {
  QDSP_REQUIRE(n >= 0, QDSP_ERR_INVALID, "sort: negative length");
  QDSP_REQUIRE(key_kind >= QDSP_KEY_U32 && key_kind <= QDSP_KEY_F64, QDSP_ERR_INVALID, "sort: bad key kind");
  QDSP_REQUIRE(value_bytes == 0 || value_bytes == 4 || value_bytes == 8, QDSP_ERR_INVALID,
               "sort: value_bytes must be 0, 4 or 8");
  QDSP_REQUIRE(n < (int64_t{1} << 30), QDSP_ERR_UNSUPPORTED, "sort: n must be below 2^30");
  if (values == nullptr) value_bytes = 0;
  if (n <= 1) return;
  QDSP_REQUIRE(keys != nullptr && tmp != nullptr, QDSP_ERR_INVALID, "sort: null pointer");
  const int kb = key_bytes_of(key_kind);
  QDSP_REQUIRE(tmp_bytes >= sort_tmp_bytes(n, kb, value_bytes), QDSP_ERR_INVALID, "sort: tmp buffer too small");
  const int end_bit = kb * 8;
  if (kb == 4) {
    auto *k = static_cast<uint32_t *>(keys);
    if (value_bytes == 0) rs_sort_inplace<uint32_t, NoValue>(k, nullptr, n, key_kind, 0, end_bit, tmp, stream);
    else if (value_bytes == 4) rs_sort_inplace<uint32_t, uint32_t>(k, static_cast<uint32_t *>(values), n, key_kind, 0, end_bit, tmp, stream);
    else rs_sort_inplace<uint32_t, uint64_t>(k, static_cast<uint64_t *>(values), n, key_kind, 0, end_bit, tmp, stream);
  } else {
    auto *k = static_cast<uint64_t *>(keys);
    if (value_bytes == 0) rs_sort_inplace<uint64_t, NoValue>(k, nullptr, n, key_kind, 0, end_bit, tmp, stream);
    else if (value_bytes == 4) rs_sort_inplace<uint64_t, uint32_t>(k, static_cast<uint32_t *>(values), n, key_kind, 0, end_bit, tmp, stream);
    else rs_sort_inplace<uint64_t, uint64_t>(k, static_cast<uint64_t *>(values), n, key_kind, 0, end_bit, tmp, stream);
  }
}

// qd.algorithms.sort (python/quadrants/algorithms/_radix_sort.py:277-344): caller-owned ping-pong buffers and scratch,
// device-resident count, low `end_bit` bits, even pass count so the result lands in keys / values.
void sort_modern(void *keys, void *tmp_keys, void *values, void *tmp_values, int64_t capacity, const int32_t *n_dev,
                 int key_elem, int value_bytes, int end_bit, void *scratch, size_t scratch_bytes, cudaStream_t stream) {
  static const int kind_of_elem[6] = {QDSP_KEY_I32, QDSP_KEY_U32, QDSP_KEY_F32, QDSP_KEY_I64, QDSP_KEY_U64, QDSP_KEY_F64};
  QDSP_REQUIRE(key_elem >= 0 && key_elem < 6, QDSP_ERR_UNSUPPORTED, "sort: key dtype not supported");
  QDSP_REQUIRE(capacity >= 0, QDSP_ERR_INVALID, "sort: negative capacity");
  QDSP_REQUIRE(capacity < (int64_t{1} << 30), QDSP_ERR_UNSUPPORTED, "sort: capacity must be below 2^30");
  QDSP_REQUIRE(value_bytes == 0 || value_bytes == 4 || value_bytes == 8, QDSP_ERR_INVALID,
               "sort: value_bytes must be 0, 4 or 8");
  const int kind = kind_of_elem[key_elem];
  const int kb = key_bytes_of(kind);
  QDSP_REQUIRE(end_bit > 0 && end_bit % 8 == 0 && end_bit <= kb * 8, QDSP_ERR_INVALID,
               "sort: end_bit must be a positive multiple of 8 no larger than the key width");
  QDSP_REQUIRE((end_bit / 8) % 2 == 0, QDSP_ERR_INVALID,
               "sort: end_bit must give an even number of digit passes so the result lands back in keys");
  if (capacity == 0) return;
  QDSP_REQUIRE(keys && tmp_keys && scratch, QDSP_ERR_INVALID, "sort: null pointer");
  QDSP_REQUIRE(keys != tmp_keys, QDSP_ERR_INVALID, "sort: keys and tmp_keys must be distinct buffers");
  if (value_bytes) {
    QDSP_REQUIRE(values && tmp_values, QDSP_ERR_INVALID, "sort: null values pointer");
    QDSP_REQUIRE(values != tmp_values, QDSP_ERR_INVALID, "sort: values and tmp_values must be distinct buffers");
  }
  QDSP_REQUIRE(scratch_bytes >= sort_modern_scratch_bytes(capacity, kb), QDSP_ERR_INVALID, "sort: scratch too small");
  QDSP_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 3u) == 0, QDSP_ERR_INVALID, "sort: scratch must be 4-byte aligned");
  // an IEEE / two's-complement twiddle flips high bits: partial-width sorts are only meaningful for unsigned keys,
  // exactly like the reference (it twiddles the whole key and sorts the low end_bit bits of the twiddled key)
  void *kres = nullptr;
  if (kb == 4) {
    auto *k = static_cast<uint32_t *>(keys), *ka = static_cast<uint32_t *>(tmp_keys);
    if (value_bytes == 0) kres = rs_run<uint32_t, NoValue>(k, nullptr, ka, nullptr, scratch, capacity, n_dev, kind, 0, end_bit, stream, nullptr);
    else if (value_bytes == 4) kres = rs_run<uint32_t, uint32_t>(k, static_cast<uint32_t *>(values), ka, static_cast<uint32_t *>(tmp_values), scratch, capacity, n_dev, kind, 0, end_bit, stream, nullptr);
    else kres = rs_run<uint32_t, uint64_t>(k, static_cast<uint64_t *>(values), ka, static_cast<uint64_t *>(tmp_values), scratch, capacity, n_dev, kind, 0, end_bit, stream, nullptr);
  } else {
    auto *k = static_cast<uint64_t *>(keys), *ka = static_cast<uint64_t *>(tmp_keys);
    if (value_bytes == 0) kres = rs_run<uint64_t, NoValue>(k, nullptr, ka, nullptr, scratch, capacity, n_dev, kind, 0, end_bit, stream, nullptr);
    else if (value_bytes == 4) kres = rs_run<uint64_t, uint32_t>(k, static_cast<uint32_t *>(values), ka, static_cast<uint32_t *>(tmp_values), scratch, capacity, n_dev, kind, 0, end_bit, stream, nullptr);
    else kres = rs_run<uint64_t, uint64_t>(k, static_cast<uint64_t *>(values), ka, static_cast<uint64_t *>(tmp_values), scratch, capacity, n_dev, kind, 0, end_bit, stream, nullptr);
  }
  QDSP_REQUIRE(kres == keys, QDSP_ERR_STATE, "sort: internal error, result not in keys");
}

// Internal entry for the CSR build: sorts only bits [0, key_bits) and returns where the result lives (no copy back).
void sort_build_keys(void *keys, void *values, int64_t n, int key_bytes, int value_bytes, int key_bits, void *tmp,
                     cudaStream_t stream, void **keys_out, void **values_out) {
  QDSP_REQUIRE(n < (int64_t{1} << 30), QDSP_ERR_UNSUPPORTED, "build: more than 2^30 triplets");
  if (key_bits < 1) key_bits = 1;
  void *kr = nullptr, *vr = nullptr;
  if (key_bytes == 4) {
    if (value_bytes == 4) kr = rs_run_tmp<uint32_t, uint32_t>(static_cast<uint32_t *>(keys), static_cast<uint32_t *>(values), n, QDSP_KEY_U32, 0, key_bits, tmp, stream, &vr);
    else kr = rs_run_tmp<uint32_t, uint64_t>(static_cast<uint32_t *>(keys), static_cast<uint64_t *>(values), n, QDSP_KEY_U32, 0, key_bits, tmp, stream, &vr);
  } else {
    if (value_bytes == 4) kr = rs_run_tmp<uint64_t, uint32_t>(static_cast<uint64_t *>(keys), static_cast<uint32_t *>(values), n, QDSP_KEY_U64, 0, key_bits, tmp, stream, &vr);
    else kr = rs_run_tmp<uint64_t, uint64_t>(static_cast<uint64_t *>(keys), static_cast<uint64_t *>(values), n, QDSP_KEY_U64, 0, key_bits, tmp, stream, &vr);
  }
  *keys_out = kr;
  *values_out = vr;
}

}  // namespace qdsp
