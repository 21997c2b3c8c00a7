// device_algos.cu -- device-wide reduce / exclusive scan / select / reduce-by-key, the kernels behind the reference's
// modern qd.algorithms API:
//   reduce_{add,min,max}          python/quadrants/algorithms/_reduce.py:448-507
//   exclusive_scan_{add,min,max}  python/quadrants/algorithms/_scan.py:496-541
//   select                        python/quadrants/algorithms/_select.py:124-169
//   reduce_by_key_add             python/quadrants/algorithms/_reduce_by_key.py:145-200
// The reference emits a staircase of 256-element phases (log256_max_n levels, ~N/256 scratch slots, one offloaded
// launch per phase, per-element index arrays in scratch for select / reduce_by_key).  Here every op is
//   (1) one streaming pass that reduces each 16 KB tile to one partial,
//   (2) one single-CTA pass over the partials (at most N/4096 of them) that also lands the total on the device,
//   (3) one streaming pass that re-reads the tile and emits the result,
// i.e. 3 N-sized HBM touches (2 reads + 1 write) and NO spin-waiting between CTAs: the live count `n` is read from
// device memory by every kernel, so the launch topology depends on the capacity only and a call can sit in a CUDA graph
// (the reference's design goal for these ops).  No per-element scratch: select / reduce_by_key recompute the in-tile
// positions from the flags / keys in pass 3 instead of storing N indices.
//
// HBM-bound.  Algorithmic bytes: reduce N*s; scan 3*N*s (the decoupled look-back scan of scan.cu does 2*N*s but its
// result order is timing dependent for floats; this one is fixed-order and hence bit-reproducible); select
// N*(4+4+s) + K*s; reduce_by_key N*(2*sk + sv) + R*(sk+sv).
// Tile layout (as in scan.cu): warp w owns DA_ROUNDS*32*VEC consecutive elements; in round j lane l holds the VEC
// consecutive elements starting at warp_base + (j*32 + l)*VEC, moved with one 128-bit access.
#include <type_traits>

#include "host.hpp"

namespace qdsp {

constexpr int DA_THREADS = 256;
constexpr int DA_WARPS = DA_THREADS / 32;
constexpr int DA_ROUNDS = 4;
constexpr int DA_SCAN_THREADS = 1024;  // the single-CTA pass over the partials

template <typename T>
struct DaTile {
  static constexpr int VEC = 16 / sizeof(T);                  // elements per 128-bit access: 4 or 2
  static constexpr int WARP_SPAN = DA_ROUNDS * 32 * VEC;      // 512 or 256
  static constexpr int TILE = DA_WARPS * WARP_SPAN;           // 4096 or 2048 elements = 16 KB
};

// ---- the three monoids ------------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ T da_highest();
template <> __device__ __forceinline__ int32_t da_highest<int32_t>() { return 0x7fffffff; }
template <> __device__ __forceinline__ uint32_t da_highest<uint32_t>() { return 0xffffffffu; }
template <> __device__ __forceinline__ float da_highest<float>() { return __int_as_float(0x7f800000); }
template <> __device__ __forceinline__ int64_t da_highest<int64_t>() { return 0x7fffffffffffffffll; }
template <> __device__ __forceinline__ uint64_t da_highest<uint64_t>() { return 0xffffffffffffffffull; }
template <> __device__ __forceinline__ double da_highest<double>() { return __longlong_as_double(0x7ff0000000000000ll); }
template <typename T> __device__ __forceinline__ T da_lowest();
template <> __device__ __forceinline__ int32_t da_lowest<int32_t>() { return static_cast<int32_t>(0x80000000u); }
template <> __device__ __forceinline__ uint32_t da_lowest<uint32_t>() { return 0u; }
template <> __device__ __forceinline__ float da_lowest<float>() { return __int_as_float(0xff800000); }
template <> __device__ __forceinline__ int64_t da_lowest<int64_t>() { return static_cast<int64_t>(0x8000000000000000ull); }
template <> __device__ __forceinline__ uint64_t da_lowest<uint64_t>() { return 0ull; }
template <> __device__ __forceinline__ double da_lowest<double>() { return __longlong_as_double(0xfff0000000000000ll); }

template <typename T, int OP>
struct DaOp {
  static __device__ __forceinline__ T identity() {
    if constexpr (OP == QDSP_OP_ADD) return T(0);
    else if constexpr (OP == QDSP_OP_MIN) return da_highest<T>();
    else return da_lowest<T>();
  }
  static __device__ __forceinline__ T apply(T a, T b) {
    if constexpr (OP == QDSP_OP_ADD) return a + b;
    else if constexpr (OP == QDSP_OP_MIN) return b < a ? b : a;
    else return a < b ? b : a;
  }
};

// live count: the device-resident count when given (clamped to the capacity), else the capacity
__device__ __forceinline__ long long da_count(long long capacity, const int *n_dev) {
  if (n_dev == nullptr) return capacity;
  const long long n = *n_dev;
  return n < 0 ? 0 : (n > capacity ? capacity : n);
}

template <typename T>
__device__ __forceinline__ T da_shfl_up(T v, int d) { return __shfl_up_sync(FULL, v, d); }
template <typename T>
__device__ __forceinline__ T da_shfl(T v, int src) { return __shfl_sync(FULL, v, src); }
template <typename T>
__device__ __forceinline__ T da_shfl_xor(T v, int m) { return __shfl_xor_sync(FULL, v, m); }

template <typename T, int OP>
__device__ __forceinline__ T da_warp_reduce(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = DaOp<T, OP>::apply(v, da_shfl_xor(v, o));
  return v;
}
// inclusive warp scan in lane order (left operand = lower lanes)
template <typename T, int OP>
__device__ __forceinline__ T da_warp_scan(T v) {
  const unsigned lane = lane_id();
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const T lo = da_shfl_up(v, o);
    if (lane >= (unsigned)o) v = DaOp<T, OP>::apply(lo, v);
  }
  return v;
}

// Loads the calling warp's span of a tile; positions at or beyond n hold `fill`.
template <typename T>
__device__ __forceinline__ void da_load_span(const T *__restrict__ in, long long warp_base, long long n, T fill,
                                             T (&v)[DA_ROUNDS][DaTile<T>::VEC]) {
  constexpr int VEC = DaTile<T>::VEC;
  const unsigned lane = lane_id();
  const bool vec_ok = (warp_base + DaTile<T>::WARP_SPAN <= n) && ((reinterpret_cast<uintptr_t>(in) & 15u) == 0);
  if (vec_ok) {
#pragma unroll
    for (int j = 0; j < DA_ROUNDS; ++j) {
      const int4 q = *reinterpret_cast<const int4 *>(in + warp_base + (j * 32 + lane) * VEC);
      if constexpr (VEC == 4) {
        const int w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) memcpy(&v[j][e], &w[e], 4);
      } else {
        const int2 a = make_int2(q.x, q.y), b = make_int2(q.z, q.w);
        memcpy(&v[j][0], &a, 8);
        memcpy(&v[j][1], &b, 8);
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < DA_ROUNDS; ++j)
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        const long long i = warp_base + (j * 32 + lane) * VEC + e;
        v[j][e] = i < n ? in[i] : fill;
      }
  }
}

template <typename T>
__device__ __forceinline__ void da_store_span(T *__restrict__ out, long long warp_base, long long n,
                                              const T (&v)[DA_ROUNDS][DaTile<T>::VEC]) {
  constexpr int VEC = DaTile<T>::VEC;
  const unsigned lane = lane_id();
  const bool vec_ok = (warp_base + DaTile<T>::WARP_SPAN <= n) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
  if (vec_ok) {
#pragma unroll
    for (int j = 0; j < DA_ROUNDS; ++j) {
      int4 q;
      memcpy(&q, &v[j][0], 16);
      *reinterpret_cast<int4 *>(out + warp_base + (j * 32 + lane) * VEC) = q;
    }
  } else {
#pragma unroll
    for (int j = 0; j < DA_ROUNDS; ++j)
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        const long long i = warp_base + (j * 32 + lane) * VEC + e;
        if (i < n) out[i] = v[j][e];
      }
  }
}

// Combines the 8 warp values of a CTA in warp order.  Every thread gets (exclusive prefix of its warp, CTA total).
template <typename T, int OP>
__device__ __forceinline__ void da_block_combine(T warp_value, T *s_warp /* [DA_WARPS] */, T &warp_prefix, T &total) {
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  if (lane == 0) s_warp[warp] = warp_value;
  __syncthreads();
  T acc = DaOp<T, OP>::identity();
  warp_prefix = acc;
#pragma unroll
  for (int w = 0; w < DA_WARPS; ++w) {
    if (w == (int)warp) warp_prefix = acc;
    acc = DaOp<T, OP>::apply(acc, s_warp[w]);
  }
  total = acc;
}

// ---- pass 1: one partial per tile -----------------------------------------------------------------------------------
template <typename T, int OP>
__global__ void __launch_bounds__(DA_THREADS)
da_tile_reduce_kernel(const T *__restrict__ in, long long capacity, const int *__restrict__ n_dev,
                      T *__restrict__ partials) {
  using Op = DaOp<T, OP>;
  constexpr int VEC = DaTile<T>::VEC;
  __shared__ T s_warp[DA_WARPS];
  const long long n = da_count(capacity, n_dev);
  const long long tile_base = static_cast<long long>(blockIdx.x) * DaTile<T>::TILE;
  if (tile_base >= n) return;
  const unsigned warp = threadIdx.x >> 5;
  T v[DA_ROUNDS][VEC];
  da_load_span(in, tile_base + warp * DaTile<T>::WARP_SPAN, n, Op::identity(), v);
  T acc = Op::identity();
#pragma unroll
  for (int j = 0; j < DA_ROUNDS; ++j)
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc = Op::apply(acc, v[j][e]);
  acc = da_warp_reduce<T, OP>(acc);
  T warp_prefix, total;
  da_block_combine<T, OP>(acc, s_warp, warp_prefix, total);
  if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// ---- pass 2: exclusive scan of the partials in place (one CTA), total -> *total_out ------------------------------------
template <typename T, int OP>
__global__ void __launch_bounds__(DA_SCAN_THREADS)
da_scan_partials_kernel(T *__restrict__ partials, long long capacity, const int *__restrict__ n_dev, int tile_elems,
                        T *__restrict__ total_out) {
  using Op = DaOp<T, OP>;
  __shared__ T s_warp[DA_SCAN_THREADS / 32];
  __shared__ T s_chunk_total;
  const long long n = da_count(capacity, n_dev);
  const long long count = (n + tile_elems - 1) / tile_elems;
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  T carry = Op::identity();
  for (long long base = 0; base < count; base += DA_SCAN_THREADS) {
    const long long i = base + threadIdx.x;
    const T v = i < count ? partials[i] : Op::identity();
    const T incl = da_warp_scan<T, OP>(v);
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const T w = s_warp[lane];
      const T wi = da_warp_scan<T, OP>(w);
      T we = da_shfl_up(wi, 1);
      if (lane == 0) we = Op::identity();
      s_warp[lane] = we;  // exclusive prefix of each warp
      if (lane == 31) s_chunk_total = wi;
    }
    __syncthreads();
    T lane_excl = da_shfl_up(incl, 1);
    if (lane == 0) lane_excl = Op::identity();
    const T excl = Op::apply(carry, Op::apply(s_warp[warp], lane_excl));
    if (i < count) partials[i] = excl;
    carry = Op::apply(carry, s_chunk_total);
    __syncthreads();  // s_warp / s_chunk_total are rewritten by the next chunk
  }
  if (threadIdx.x == 0 && total_out != nullptr) *total_out = carry;
}

// In-tile exclusive scan shared by pass 3 of scan / select / reduce_by_key.  On return ex[j][e] is the exclusive
// prefix (within the CTA's tile, in element order) of v[j][e]; returns the tile-exclusive prefix of the calling warp
// already folded in.  `s_warp` is CTA-shared scratch of DA_WARPS values.
template <typename T, int OP>
__device__ __forceinline__ void da_tile_exclusive(const T (&v)[DA_ROUNDS][DaTile<T>::VEC],
                                                  T (&ex)[DA_ROUNDS][DaTile<T>::VEC], T tile_prefix, T *s_warp) {
  using Op = DaOp<T, OP>;
  constexpr int VEC = DaTile<T>::VEC;
  const unsigned lane = lane_id();
  T running = Op::identity();  // everything of this warp before round j
  T base[DA_ROUNDS];           // exclusive prefix, inside the warp's span, of the thread's first element of round j
#pragma unroll
  for (int j = 0; j < DA_ROUNDS; ++j) {
    T t = v[j][0];
#pragma unroll
    for (int e = 1; e < VEC; ++e) t = Op::apply(t, v[j][e]);
    const T incl = da_warp_scan<T, OP>(t);
    T lane_excl = da_shfl_up(incl, 1);
    if (lane == 0) lane_excl = Op::identity();
    base[j] = Op::apply(running, lane_excl);
    running = Op::apply(running, da_shfl(incl, 31));
  }
  T warp_prefix, total;
  da_block_combine<T, OP>(running, s_warp, warp_prefix, total);
  const T pre = Op::apply(tile_prefix, warp_prefix);
#pragma unroll
  for (int j = 0; j < DA_ROUNDS; ++j) {
    T acc = Op::apply(pre, base[j]);
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      ex[j][e] = acc;
      acc = Op::apply(acc, v[j][e]);
    }
  }
}

// ---- pass 3 of exclusive_scan ---------------------------------------------------------------------------------------
template <typename T, int OP>
__global__ void __launch_bounds__(DA_THREADS)
da_tile_scan_kernel(const T *__restrict__ in, T *__restrict__ out, long long capacity, const int *__restrict__ n_dev,
                    const T *__restrict__ tile_prefix) {
  using Op = DaOp<T, OP>;
  constexpr int VEC = DaTile<T>::VEC;
  __shared__ T s_warp[DA_WARPS];
  const long long n = da_count(capacity, n_dev);
  const long long tile_base = static_cast<long long>(blockIdx.x) * DaTile<T>::TILE;
  if (tile_base >= n) return;
  const long long warp_base = tile_base + (threadIdx.x >> 5) * DaTile<T>::WARP_SPAN;
  T v[DA_ROUNDS][VEC], ex[DA_ROUNDS][VEC];
  da_load_span(in, warp_base, n, Op::identity(), v);
  da_tile_exclusive<T, OP>(v, ex, tile_prefix[blockIdx.x], s_warp);
  da_store_span(out, warp_base, n, ex);
}

// ---- pass 3 of select: out[position] = in[i] for flagged i; WORDS 32-bit words per item ------------------------------
template <int WORDS>
__global__ void __launch_bounds__(DA_THREADS)
da_select_scatter_kernel(const uint32_t *__restrict__ in, const uint32_t *__restrict__ flags,
                         uint32_t *__restrict__ out, long long capacity, const int *__restrict__ n_dev,
                         const uint32_t *__restrict__ tile_prefix, int words_rt) {
  using T = uint32_t;
  constexpr int VEC = DaTile<T>::VEC;
  __shared__ T s_warp[DA_WARPS];
  const long long n = da_count(capacity, n_dev);
  const long long tile_base = static_cast<long long>(blockIdx.x) * DaTile<T>::TILE;
  if (tile_base >= n) return;
  const unsigned lane = lane_id();
  const long long warp_base = tile_base + (threadIdx.x >> 5) * DaTile<T>::WARP_SPAN;
  T f[DA_ROUNDS][VEC], pos[DA_ROUNDS][VEC];
  da_load_span(flags, warp_base, n, 0u, f);
  da_tile_exclusive<T, QDSP_OP_ADD>(f, pos, tile_prefix[blockIdx.x], s_warp);
  const int words = WORDS > 0 ? WORDS : words_rt;
#pragma unroll
  for (int j = 0; j < DA_ROUNDS; ++j)
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      const long long i = warp_base + (j * 32 + lane) * VEC + e;
      if (i < n && f[j][e] != 0u) {
        const uint32_t *src = in + i * words;
        uint32_t *dst = out + static_cast<long long>(pos[j][e]) * words;
        if constexpr (WORDS == 1) {
          dst[0] = src[0];
        } else if constexpr (WORDS == 2) {
          *reinterpret_cast<uint2 *>(dst) = *reinterpret_cast<const uint2 *>(src);
        } else {
          for (int w = 0; w < words; ++w) dst[w] = src[w];
        }
      }
    }
}

// ---- reduce_by_key ----------------------------------------------------------------------------------------------------
template <bool FLOAT_KEYS>
__device__ __forceinline__ bool da_key_differs(uint32_t a, uint32_t b) {
  if constexpr (FLOAT_KEYS) return __uint_as_float(a) != __uint_as_float(b);  // NaN != NaN, +0 == -0
  else return a != b;
}

// head[j][e] = 1 when element (j, e) of the calling thread opens a run (differs from its predecessor, or is element 0)
template <bool FLOAT_KEYS>
__device__ __forceinline__ void da_head_flags(const uint32_t *__restrict__ keys, long long warp_base, long long n,
                                              const uint32_t (&k)[DA_ROUNDS][4], uint32_t (&head)[DA_ROUNDS][4]) {
  const unsigned lane = lane_id();
  uint32_t before_warp = 0;  // key just before the warp's span (lane 0 only)
  if (lane == 0 && warp_base > 0 && warp_base < n) before_warp = keys[warp_base - 1];
  uint32_t last_of_prev_round = before_warp;
#pragma unroll
  for (int j = 0; j < DA_ROUNDS; ++j) {
    uint32_t prev = __shfl_up_sync(FULL, k[j][3], 1);
    if (lane == 0) prev = last_of_prev_round;
    last_of_prev_round = __shfl_sync(FULL, k[j][3], 31);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const long long i = warp_base + (j * 32 + lane) * 4 + e;
      const uint32_t p = e == 0 ? prev : k[j][e - 1];
      head[j][e] = (i < n && (i == 0 || da_key_differs<FLOAT_KEYS>(k[j][e], p))) ? 1u : 0u;
    }
  }
}

template <bool FLOAT_KEYS>
__global__ void __launch_bounds__(DA_THREADS)
da_rbk_count_kernel(const uint32_t *__restrict__ keys, long long capacity, const int *__restrict__ n_dev,
                    uint32_t *__restrict__ partials) {
  using T = uint32_t;
  __shared__ T s_warp[DA_WARPS];
  const long long n = da_count(capacity, n_dev);
  const long long tile_base = static_cast<long long>(blockIdx.x) * DaTile<T>::TILE;
  if (tile_base >= n) return;
  const long long warp_base = tile_base + (threadIdx.x >> 5) * DaTile<T>::WARP_SPAN;
  T k[DA_ROUNDS][4], head[DA_ROUNDS][4];
  da_load_span(keys, warp_base, n, 0u, k);
  da_head_flags<FLOAT_KEYS>(keys, warp_base, n, k, head);
  T acc = 0;
#pragma unroll
  for (int j = 0; j < DA_ROUNDS; ++j)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc += head[j][e];
  acc = da_warp_reduce<T, QDSP_OP_ADD>(acc);
  T warp_prefix, total;
  da_block_combine<T, QDSP_OP_ADD>(acc, s_warp, warp_prefix, total);
  if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// values_out[0 : *count) = 0 (the reference's _rbk_zero_values_out_phase, restricted to the runs that exist)
__global__ void __launch_bounds__(DA_THREADS)
da_zero_prefix_kernel(uint32_t *__restrict__ out, const int *__restrict__ count_dev, long long capacity) {
  long long n = *count_dev;
  if (n > capacity) n = capacity;
  const long long stride = static_cast<long long>(gridDim.x) * DA_THREADS;
  for (long long i = static_cast<long long>(blockIdx.x) * DA_THREADS + threadIdx.x; i < n; i += stride) out[i] = 0u;
}

template <typename VT>
__device__ __forceinline__ void da_atomic_add(VT *p, VT v) { atomicAdd(p, v); }

// pass 3: run index = (number of heads up to and including the element) - 1; heads write the key, every thread adds
// the values of its (up to 4) consecutive elements run by run.  A warp round that lies inside ONE run is summed by
// shuffles first (one atomic per 128 elements: long runs do not serialise on one address).
template <bool FLOAT_KEYS, typename VT>
__global__ void __launch_bounds__(DA_THREADS)
da_rbk_scatter_kernel(const uint32_t *__restrict__ keys, const VT *__restrict__ vals, uint32_t *__restrict__ keys_out,
                      VT *__restrict__ vals_out, long long capacity, const int *__restrict__ n_dev,
                      const uint32_t *__restrict__ tile_prefix) {
  using T = uint32_t;
  __shared__ T s_warp[DA_WARPS];
  const long long n = da_count(capacity, n_dev);
  const long long tile_base = static_cast<long long>(blockIdx.x) * DaTile<T>::TILE;
  if (tile_base >= n) return;
  const unsigned lane = lane_id();
  const long long warp_base = tile_base + (threadIdx.x >> 5) * DaTile<T>::WARP_SPAN;
  T k[DA_ROUNDS][4], head[DA_ROUNDS][4], ex[DA_ROUNDS][4];
  VT v[DA_ROUNDS][4];
  da_load_span(keys, warp_base, n, 0u, k);
  da_load_span(vals, warp_base, n, VT(0), v);
  da_head_flags<FLOAT_KEYS>(keys, warp_base, n, k, head);
  da_tile_exclusive<T, QDSP_OP_ADD>(head, ex, tile_prefix[blockIdx.x], s_warp);
#pragma unroll
  for (int j = 0; j < DA_ROUNDS; ++j) {
    const long long i0 = warp_base + (j * 32 + lane) * 4;
    // run index of element e: heads before it, plus itself if it is a head, minus one
    T run[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) run[e] = ex[j][e] + head[j][e] - 1u;
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (head[j][e]) keys_out[run[e]] = k[j][e];
    const bool full = i0 + 3 < n;
    const T run0 = __shfl_sync(FULL, run[0], 0);
    const bool one_run = full && run[0] == run0 && run[3] == run0;  // run indices are non-decreasing
    if (__all_sync(FULL, one_run)) {
      VT s = (v[j][0] + v[j][1]) + (v[j][2] + v[j][3]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
      if (lane == 0) da_atomic_add(&vals_out[run0], s);
    } else {
      T cur = run[0];
      VT acc = VT(0);
      bool any = false;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (i0 + e < n) {
          if (any && run[e] != cur) {
            da_atomic_add(&vals_out[cur], acc);
            acc = VT(0);
          }
          cur = run[e];
          acc += v[j][e];
          any = true;
        }
      }
      if (any) da_atomic_add(&vals_out[cur], acc);
    }
  }
}

// ---- host side --------------------------------------------------------------------------------------------------------
namespace {
constexpr size_t DA_ALIGN = 256;
size_t da_align(size_t x) { return (x + DA_ALIGN - 1) / DA_ALIGN * DA_ALIGN; }
int64_t da_tiles(int64_t capacity, int tile) { return capacity > 0 ? (capacity + tile - 1) / tile : 1; }
int da_elem_bytes(int elem) { return elem >= QDSP_I64 ? 8 : 4; }

void da_check_common(const char *who, int64_t capacity, const void *scratch, size_t scratch_bytes, size_t need) {
  QDSP_REQUIRE(capacity >= 0, QDSP_ERR_INVALID, std::string(who) + ": negative capacity");
  QDSP_REQUIRE(capacity < (int64_t{1} << 31), QDSP_ERR_UNSUPPORTED, std::string(who) + ": capacity must be below 2^31");
  QDSP_REQUIRE(scratch != nullptr, QDSP_ERR_INVALID, std::string(who) + ": null scratch");
  QDSP_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 7u) == 0, QDSP_ERR_INVALID,
               std::string(who) + ": scratch must be 8-byte aligned");
  QDSP_REQUIRE(scratch_bytes >= need, QDSP_ERR_INVALID, std::string(who) + ": scratch buffer too small");
}

template <typename T, int OP>
void da_reduce_t(const void *in, void *out, int64_t capacity, const int *n_dev, void *scratch, cudaStream_t stream) {
  T *partials = static_cast<T *>(scratch);
  const int64_t tiles = da_tiles(capacity, DaTile<T>::TILE);
  if (capacity > 0) {
    da_tile_reduce_kernel<T, OP><<<static_cast<unsigned>(tiles), DA_THREADS, 0, stream>>>(
        static_cast<const T *>(in), capacity, n_dev, partials);
    QDSP_LAUNCH_CHECK();
  }
  da_scan_partials_kernel<T, OP><<<1, DA_SCAN_THREADS, 0, stream>>>(partials, capacity, n_dev, DaTile<T>::TILE,
                                                                    static_cast<T *>(out));
  QDSP_LAUNCH_CHECK();
}

template <typename T, int OP>
void da_scan_t(const void *in, void *out, int64_t capacity, const int *n_dev, void *scratch, cudaStream_t stream) {
  if (capacity == 0) return;
  T *partials = static_cast<T *>(scratch);
  const unsigned tiles = static_cast<unsigned>(da_tiles(capacity, DaTile<T>::TILE));
  da_tile_reduce_kernel<T, OP><<<tiles, DA_THREADS, 0, stream>>>(static_cast<const T *>(in), capacity, n_dev, partials);
  QDSP_LAUNCH_CHECK();
  da_scan_partials_kernel<T, OP><<<1, DA_SCAN_THREADS, 0, stream>>>(partials, capacity, n_dev, DaTile<T>::TILE, nullptr);
  QDSP_LAUNCH_CHECK();
  da_tile_scan_kernel<T, OP><<<tiles, DA_THREADS, 0, stream>>>(static_cast<const T *>(in), static_cast<T *>(out),
                                                               capacity, n_dev, partials);
  QDSP_LAUNCH_CHECK();
}

// add wraps for integers: signed add runs on the unsigned type of the same width (identical bits, no signed-overflow UB)
template <int OP, bool SCAN>
void da_run_op(int elem, const void *in, void *out, int64_t capacity, const int *n_dev, void *scratch,
               cudaStream_t stream) {
  auto run = [&](auto tag) {
    using T = decltype(tag);
    if constexpr (SCAN) da_scan_t<T, OP>(in, out, capacity, n_dev, scratch, stream);
    else da_reduce_t<T, OP>(in, out, capacity, n_dev, scratch, stream);
  };
  switch (elem) {
    case QDSP_I32:
      if constexpr (OP == QDSP_OP_ADD) run(uint32_t{}); else run(int32_t{});
      break;
    case QDSP_U32: run(uint32_t{}); break;
    case QDSP_ELEM_F32: run(float{}); break;
    case QDSP_I64:
      if constexpr (OP == QDSP_OP_ADD) run(uint64_t{}); else run(int64_t{});
      break;
    case QDSP_U64: run(uint64_t{}); break;
    case QDSP_ELEM_F64: run(double{}); break;
    default: throw Error(QDSP_ERR_INVALID, "bad element type");
  }
}

template <bool SCAN>
void da_run(int elem, int op, const void *in, void *out, int64_t capacity, const int *n_dev, void *scratch,
            cudaStream_t stream) {
  switch (op) {
    case QDSP_OP_ADD: da_run_op<QDSP_OP_ADD, SCAN>(elem, in, out, capacity, n_dev, scratch, stream); break;
    case QDSP_OP_MIN: da_run_op<QDSP_OP_MIN, SCAN>(elem, in, out, capacity, n_dev, scratch, stream); break;
    case QDSP_OP_MAX: da_run_op<QDSP_OP_MAX, SCAN>(elem, in, out, capacity, n_dev, scratch, stream); break;
    default: throw Error(QDSP_ERR_INVALID, "bad reduce op");
  }
}
}  // namespace

size_t sort_modern_scratch_bytes(int64_t capacity, int key_bytes);  // radix_sort.cu

size_t algo_scratch_bytes(int algo, int64_t capacity, int elem_bytes) {
  if (capacity < 0) return 0;
  switch (algo) {
    case QDSP_ALGO_REDUCE:
    case QDSP_ALGO_SCAN: {
      const int tile = elem_bytes == 8 ? DaTile<uint64_t>::TILE : DaTile<uint32_t>::TILE;
      return da_align(static_cast<size_t>(da_tiles(capacity, tile)) * (elem_bytes == 8 ? 8 : 4));
    }
    case QDSP_ALGO_SELECT:
    case QDSP_ALGO_REDUCE_BY_KEY:
      return da_align(static_cast<size_t>(da_tiles(capacity, DaTile<uint32_t>::TILE)) * 4);
    case QDSP_ALGO_SORT:
      return sort_modern_scratch_bytes(capacity, elem_bytes == 8 ? 8 : 4);
    default:
      return 0;
  }
}

void device_reduce(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op, void *scratch,
                   size_t scratch_bytes, cudaStream_t stream) {
  da_check_common("reduce", capacity, scratch, scratch_bytes,
                  algo_scratch_bytes(QDSP_ALGO_REDUCE, capacity, da_elem_bytes(elem)));
  QDSP_REQUIRE(out != nullptr && (in != nullptr || capacity == 0), QDSP_ERR_INVALID, "reduce: null pointer");
  da_run<false>(elem, op, in, out, capacity, n_dev, scratch, stream);
}

void device_exclusive_scan(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op,
                           void *scratch, size_t scratch_bytes, cudaStream_t stream) {
  da_check_common("exclusive_scan", capacity, scratch, scratch_bytes,
                  algo_scratch_bytes(QDSP_ALGO_SCAN, capacity, da_elem_bytes(elem)));
  if (capacity == 0) return;
  QDSP_REQUIRE(in != nullptr && out != nullptr, QDSP_ERR_INVALID, "exclusive_scan: null pointer");
  QDSP_REQUIRE(in != out, QDSP_ERR_INVALID, "exclusive_scan: out must be a distinct buffer from arr");
  da_run<true>(elem, op, in, out, capacity, n_dev, scratch, stream);
}

void device_select(const void *in, const int32_t *flags, void *out, int32_t *num_out, int64_t capacity,
                   const int32_t *n_dev, int item_bytes, void *scratch, size_t scratch_bytes, cudaStream_t stream) {
  da_check_common("select", capacity, scratch, scratch_bytes, algo_scratch_bytes(QDSP_ALGO_SELECT, capacity, 4));
  QDSP_REQUIRE(item_bytes > 0 && item_bytes % 4 == 0, QDSP_ERR_INVALID, "select: item_bytes must be a multiple of 4");
  QDSP_REQUIRE(num_out != nullptr, QDSP_ERR_INVALID, "select: null num_out");
  QDSP_REQUIRE(capacity == 0 || (in != nullptr && flags != nullptr && out != nullptr), QDSP_ERR_INVALID,
               "select: null pointer");
  QDSP_REQUIRE(in != out || capacity == 0, QDSP_ERR_INVALID, "select: out must be a distinct buffer from arr");
  using T = uint32_t;
  T *partials = static_cast<T *>(scratch);
  const unsigned tiles = static_cast<unsigned>(da_tiles(capacity, DaTile<T>::TILE));
  const T *f = reinterpret_cast<const T *>(flags);
  if (capacity > 0) {
    da_tile_reduce_kernel<T, QDSP_OP_ADD><<<tiles, DA_THREADS, 0, stream>>>(f, capacity, n_dev, partials);
    QDSP_LAUNCH_CHECK();
  }
  da_scan_partials_kernel<T, QDSP_OP_ADD><<<1, DA_SCAN_THREADS, 0, stream>>>(partials, capacity, n_dev, DaTile<T>::TILE,
                                                                             reinterpret_cast<T *>(num_out));
  QDSP_LAUNCH_CHECK();
  if (capacity == 0) return;
  const int words = item_bytes / 4;
  const T *src = static_cast<const T *>(in);
  T *dst = static_cast<T *>(out);
  const bool wide_ok = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 7u) == 0;
  if (words == 1)
    da_select_scatter_kernel<1><<<tiles, DA_THREADS, 0, stream>>>(src, f, dst, capacity, n_dev, partials, 1);
  else if (words == 2 && wide_ok)
    da_select_scatter_kernel<2><<<tiles, DA_THREADS, 0, stream>>>(src, f, dst, capacity, n_dev, partials, 2);
  else
    da_select_scatter_kernel<0><<<tiles, DA_THREADS, 0, stream>>>(src, f, dst, capacity, n_dev, partials, words);
  QDSP_LAUNCH_CHECK();
}

void device_reduce_by_key_add(const void *keys_in, const void *values_in, void *keys_out, void *values_out,
                              int32_t *num_runs, int64_t capacity, const int32_t *n_dev, int key_elem, int value_elem,
                              void *scratch, size_t scratch_bytes, cudaStream_t stream) {
  da_check_common("reduce_by_key_add", capacity, scratch, scratch_bytes,
                  algo_scratch_bytes(QDSP_ALGO_REDUCE_BY_KEY, capacity, 4));
  QDSP_REQUIRE(key_elem == QDSP_I32 || key_elem == QDSP_U32 || key_elem == QDSP_ELEM_F32, QDSP_ERR_UNSUPPORTED,
               "reduce_by_key_add: key dtype must be i32, u32 or f32");
  QDSP_REQUIRE(value_elem == QDSP_I32 || value_elem == QDSP_U32 || value_elem == QDSP_ELEM_F32, QDSP_ERR_UNSUPPORTED,
               "reduce_by_key_add: value dtype must be i32, u32 or f32");
  QDSP_REQUIRE(num_runs != nullptr, QDSP_ERR_INVALID, "reduce_by_key_add: null num_runs");
  QDSP_REQUIRE(capacity == 0 || (keys_in && values_in && keys_out && values_out), QDSP_ERR_INVALID,
               "reduce_by_key_add: null pointer");
  using T = uint32_t;
  T *partials = static_cast<T *>(scratch);
  const unsigned tiles = static_cast<unsigned>(da_tiles(capacity, DaTile<T>::TILE));
  const T *k = static_cast<const T *>(keys_in);
  const bool fk = key_elem == QDSP_ELEM_F32;
  if (capacity > 0) {
    if (fk) da_rbk_count_kernel<true><<<tiles, DA_THREADS, 0, stream>>>(k, capacity, n_dev, partials);
    else da_rbk_count_kernel<false><<<tiles, DA_THREADS, 0, stream>>>(k, capacity, n_dev, partials);
    QDSP_LAUNCH_CHECK();
  }
  da_scan_partials_kernel<T, QDSP_OP_ADD><<<1, DA_SCAN_THREADS, 0, stream>>>(partials, capacity, n_dev, DaTile<T>::TILE,
                                                                             reinterpret_cast<T *>(num_runs));
  QDSP_LAUNCH_CHECK();
  if (capacity == 0) return;
  const int zgrid = static_cast<int>(std::min<int64_t>((capacity + DA_THREADS - 1) / DA_THREADS,
                                                       static_cast<int64_t>(device_props().sm_count) * 8));
  da_zero_prefix_kernel<<<zgrid, DA_THREADS, 0, stream>>>(static_cast<T *>(values_out), num_runs, capacity);
  QDSP_LAUNCH_CHECK();
  T *ko = static_cast<T *>(keys_out);
  if (value_elem == QDSP_ELEM_F32) {
    auto *vi = static_cast<const float *>(values_in);
    auto *vo = static_cast<float *>(values_out);
    if (fk) da_rbk_scatter_kernel<true, float><<<tiles, DA_THREADS, 0, stream>>>(k, vi, ko, vo, capacity, n_dev, partials);
    else da_rbk_scatter_kernel<false, float><<<tiles, DA_THREADS, 0, stream>>>(k, vi, ko, vo, capacity, n_dev, partials);
  } else {
    auto *vi = static_cast<const uint32_t *>(values_in);
    auto *vo = static_cast<uint32_t *>(values_out);
    if (fk) da_rbk_scatter_kernel<true, uint32_t><<<tiles, DA_THREADS, 0, stream>>>(k, vi, ko, vo, capacity, n_dev, partials);
    else da_rbk_scatter_kernel<false, uint32_t><<<tiles, DA_THREADS, 0, stream>>>(k, vi, ko, vo, capacity, n_dev, partials);
  }
  QDSP_LAUNCH_CHECK();
}

}  // namespace qdsp
