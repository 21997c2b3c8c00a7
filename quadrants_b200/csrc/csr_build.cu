// csr_build.cu -- triplet buffer -> CSR, entirely on the device.
//
// Replaces SparseMatrixBuilder::build_cuda (quadrants/program/sparse_matrix.cpp:173-222: D2H of the whole buffer, a
// host unordered_map keyed row*cols+col -- int overflow past 2^31 --, three H2D copies) and
// CuSparseMatrix::build_csr_from_coo (:344-397: cusparseXcoosortByRow + Gather + Xcoo2csr).
// Contract (BASELINE.json): indptr / indices bit-exact with Eigen's row-major setFromTriplets
// (external/eigen/Eigen/src/SparseCore/SparseMatrix.h:1125-1173); values too, because duplicates are added in buffer
// order exactly like collapseDuplicates (:1475-1511) -- the sort is stable and the segment sum is sequential.
//
// Pipeline (all HBM-bound, byte model in DESIGN.md):
//   pack    : (row, col, bits) -> key = row << col_bits | col  (u32 when it fits, else u64), payload = value bits
//   sort    : onesweep LSD radix over the significant key bits only (radix_sort.cu)
//   collapse: ONE pass: head flags (key != previous key), ballot-based tile scan + decoupled look-back for the output
//             slot, sequential duplicate sum, indices, and indptr (each head that starts a new row fills the indptr
//             entries of the rows it skips over)
#include <algorithm>
#include <cstring>
#include <vector>

#include "host.hpp"
#include "lookback.cuh"

namespace qdsp {

constexpr int CB_THREADS = 256;

// ---- pack ------------------------------------------------------------------------------------------------------------
template <typename SlotT, typename KeyT, typename ValT>
__global__ void __launch_bounds__(CB_THREADS)
pack_triplets_kernel(const SlotT *__restrict__ trip, long long n, int rows, int cols, int col_bits,
                     KeyT *__restrict__ keys, ValT *__restrict__ vals, int *__restrict__ err) {
  const long long i = static_cast<long long>(blockIdx.x) * CB_THREADS + threadIdx.x;
  if (i >= n) return;
  const long long r = static_cast<long long>(trip[3 * i]);
  const long long c = static_cast<long long>(trip[3 * i + 1]);
  if (r < 0 || r >= rows || c < 0 || c >= cols) {
    atomicExch(err, 1);
    keys[i] = 0;
    vals[i] = 0;
    return;
  }
  keys[i] = (static_cast<KeyT>(r) << col_bits) | static_cast<KeyT>(c);
  vals[i] = static_cast<ValT>(trip[3 * i + 2]);  // raw value bits, same width as the slot
}

template <typename KeyT, typename ValT>
__global__ void __launch_bounds__(CB_THREADS)
pack_coo_kernel(const int *__restrict__ row, const int *__restrict__ col, const ValT *__restrict__ val, long long n,
                int rows, int cols, int col_bits, KeyT *__restrict__ keys, ValT *__restrict__ vals,
                int *__restrict__ err) {
  const long long i = static_cast<long long>(blockIdx.x) * CB_THREADS + threadIdx.x;
  if (i >= n) return;
  const int r = row[i], c = col[i];
  if (r < 0 || r >= rows || c < 0 || c >= cols) {
    atomicExch(err, 1);
    keys[i] = 0;
    vals[i] = 0;
    return;
  }
  keys[i] = (static_cast<KeyT>(r) << col_bits) | static_cast<KeyT>(c);
  vals[i] = val[i];
}

// ---- collapse ---------------------------------------------------------------------------------------------------------
constexpr int CL_ITEMS = 8;
constexpr int CL_SHORT_RUN = 6;                 // followers a head thread adds by itself
constexpr int CL_TILE = CB_THREADS * CL_ITEMS;   // 2048 sorted triplets per CTA
constexpr int CL_WARP_SPAN = 32 * CL_ITEMS;      // 256 per warp, item i of lane l <-> warp_base + i*32 + l

template <typename KeyT, typename T>
__global__ void __launch_bounds__(CB_THREADS)
collapse_kernel(const KeyT *__restrict__ keys, const T *__restrict__ vals, long long n, int rows, int col_bits,
                unsigned *__restrict__ counter, unsigned long long *__restrict__ desc, int *__restrict__ indptr,
                int *__restrict__ indices, T *__restrict__ values, long long *__restrict__ nnz_out) {
  __shared__ unsigned s_ticket;
  __shared__ unsigned s_warp_total[CB_THREADS / 32];
  __shared__ unsigned s_tile_prefix;
  const unsigned tile = take_ticket(counter, &s_ticket);
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  const long long warp_base = static_cast<long long>(tile) * CL_TILE + warp * CL_WARP_SPAN;
  const KeyT colmask = (KeyT(1) << col_bits) - 1;

  KeyT key[CL_ITEMS];
  bool head[CL_ITEMS];
  unsigned slot[CL_ITEMS];  // exclusive count of heads before this element inside the warp
  unsigned running = 0;
  KeyT carry = 0;           // key of the element just before this item's 32-element row (lane 31 of previous item)
#pragma unroll
  for (int i = 0; i < CL_ITEMS; ++i) {
    const long long p = warp_base + i * 32 + lane;
    const bool ok = p < n;
    key[i] = ok ? keys[p] : KeyT(0);
    KeyT prev = __shfl_up_sync(FULL, key[i], 1);
    if (lane == 0) prev = (i == 0) ? ((p > 0 && ok) ? keys[p - 1] : KeyT(0)) : carry;
    carry = __shfl_sync(FULL, key[i], 31);
    head[i] = ok && (p == 0 || key[i] != prev);
    const unsigned b = __ballot_sync(FULL, head[i]);
    slot[i] = running + __popc(b & lanemask_lt());
    running += __popc(b);
  }
  if (lane == 0) s_warp_total[warp] = running;
  __syncthreads();
  unsigned warp_prefix = 0, tile_total = 0;
#pragma unroll
  for (int w = 0; w < CB_THREADS / 32; ++w) {
    const unsigned t = s_warp_total[w];
    if (w < (int)warp) warp_prefix += t;
    tile_total += t;
  }
  if (warp == 0) {
    const unsigned long long ex = lookback_exclusive(desc, tile, tile_total);
    if (lane == 0) s_tile_prefix = static_cast<unsigned>(ex);
  }
  __syncthreads();
  const unsigned base = s_tile_prefix + warp_prefix;

#pragma unroll
  for (int i = 0; i < CL_ITEMS; ++i) {
    const long long p = warp_base + i * 32 + lane;
    // duplicates follow in buffer order (stable sort): v = ((v1 + v2) + v3) ... like Eigen's collapseDuplicates.
    // A head adds its first CL_SHORT_RUN followers itself; a longer run (FEM assembly: a hub vertex's diagonal can
    // collect 10^5 contributions) is finished by the whole warp -- 32 followers per coalesced load, still added
    // strictly left to right, so the value stays bit-identical to Eigen's.
    T v = T(0);
    long long q = p + 1;
    if (head[i]) {
      v = vals[p];
      for (int j = 0; j < CL_SHORT_RUN && q < n && keys[q] == key[i]; ++j, ++q) v = v + vals[q];
    }
    unsigned todo = __ballot_sync(FULL, head[i] && q < n && keys[q] == key[i]);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const KeyT k0 = __shfl_sync(FULL, key[i], src);
      long long q0 = __shfl_sync(FULL, q, src);
      T acc = __shfl_sync(FULL, v, src);
      while (true) {
        const long long qq = q0 + lane;
        const bool same = qq < n && keys[qq] == k0;
        const T x = same ? vals[qq] : T(0);
        const unsigned m = __ballot_sync(FULL, same);
        const int cnt = (m == FULL) ? 32 : __ffs(~m) - 1;  // leading run of matches
        for (int j = 0; j < cnt; ++j) acc = acc + __shfl_sync(FULL, x, j);
        if (cnt < 32) break;
        q0 += 32;
      }
      if (lane == src) v = acc;
    }
    if (head[i]) {
      const unsigned u = base + slot[i];
      indices[u] = static_cast<int>(key[i] & colmask);
      values[u] = v;
      const int r = static_cast<int>(key[i] >> col_bits);
      const int pr = (p == 0) ? -1 : static_cast<int>(keys[p - 1] >> col_bits);
      for (int rr = pr + 1; rr <= r; ++rr) indptr[rr] = static_cast<int>(u);
    }
    if (p == n - 1) {  // the last sorted triplet closes the matrix
      const unsigned total = base + slot[i] + (head[i] ? 1u : 0u);
      const int r = static_cast<int>(key[i] >> col_bits);
      for (int rr = r + 1; rr <= rows; ++rr) indptr[rr] = static_cast<int>(total);
      *nnz_out = total;
    }
  }
}

// ---- host side ----------------------------------------------------------------------------------------------------------
void CsrMatrix::allocate(int rows_, int cols_, int64_t nnz_, int dtype_) {
  rows = rows_; cols = cols_; nnz = nnz_; dtype = dtype_;
  indptr.alloc(sizeof(int32_t) * (static_cast<size_t>(rows) + 1) + 64);
  indices.alloc(sizeof(int32_t) * static_cast<size_t>(nnz) + 64);
  values.alloc(value_bytes() * static_cast<size_t>(nnz) + 64);
}

static size_t collapse_tmp_bytes(int64_t n) {
  const int64_t tiles = (n + CL_TILE - 1) / CL_TILE;
  return 128 + static_cast<size_t>(tiles) * sizeof(unsigned long long) + 16;
}

std::unique_ptr<CsrMatrix> csr_from_packed(int rows, int cols, int64_t n, int dtype, int key_bytes, int col_bits,
                                           DeviceBuffer &keys, DeviceBuffer &vals, cudaStream_t stream) {
  const int vb = dtype == QDSP_F64 ? 8 : 4;
  const int key_bits = bits_for(rows) + col_bits;
  DeviceBuffer sort_tmp = scratch_buffer(sort_tmp_bytes(n, key_bytes, vb));
  void *ks = nullptr, *vs = nullptr;
  sort_build_keys(keys.get(), vals.get(), n, key_bytes, vb, key_bits, sort_tmp.get(), stream, &ks, &vs);
  // the buffer pair that does not hold the sorted data is free: collapse writes indices/values there first, because
  // the exact nnz is only known after the pass
  const bool in_caller = (ks == keys.get());
  // when the result is in the scratch pair the caller pair is free; otherwise take two spare buffers
  DeviceBuffer spare_idx, spare_val;
  int *idx_tmp;
  void *val_tmp;
  if (!in_caller) {
    idx_tmp = keys.get<int>();
    val_tmp = vals.get();
  } else {
    spare_idx.alloc_scratch(sizeof(int) * static_cast<size_t>(n));
    spare_val.alloc_scratch(static_cast<size_t>(vb) * n);
    idx_tmp = spare_idx.get<int>();
    val_tmp = spare_val.get();
  }
  auto out = std::make_unique<CsrMatrix>();
  out->rows = rows; out->cols = cols; out->dtype = dtype;
  out->indptr.alloc(sizeof(int32_t) * (static_cast<size_t>(rows) + 1) + 64);
  DeviceBuffer ctl(collapse_tmp_bytes(n));
  QDSP_CUDA(cudaMemsetAsync(ctl.get(), 0, ctl.bytes(), stream));
  unsigned *counter = ctl.get<unsigned>();
  auto *desc = reinterpret_cast<unsigned long long *>(ctl.get<char>() + 128);
  const int64_t tiles = (n + CL_TILE - 1) / CL_TILE;
  long long *nnz_dev = reinterpret_cast<long long *>(ctl.get<char>() + 128 + tiles * sizeof(unsigned long long));
  // 8-byte alignment of nnz_dev holds: 128 + 8*tiles
  const unsigned grid = static_cast<unsigned>(tiles);
#define QDSP_COLLAPSE(KeyT, T)                                                                                      \
  collapse_kernel<KeyT, T><<<grid, CB_THREADS, 0, stream>>>(static_cast<const KeyT *>(ks), static_cast<const T *>(vs), \
                                                            n, rows, col_bits, counter, desc,                         \
                                                            out->indptr.get<int>(), idx_tmp,                          \
                                                            static_cast<T *>(val_tmp), nnz_dev)
  if (key_bytes == 4 && vb == 4) QDSP_COLLAPSE(uint32_t, float);
  else if (key_bytes == 4) QDSP_COLLAPSE(uint32_t, double);
  else if (vb == 4) QDSP_COLLAPSE(uint64_t, float);
  else QDSP_COLLAPSE(uint64_t, double);
#undef QDSP_COLLAPSE
  QDSP_LAUNCH_CHECK();
  long long nnz = 0;
  QDSP_CUDA(cudaMemcpyAsync(&nnz, nnz_dev, sizeof(nnz), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  out->nnz = nnz;
  out->indices.alloc(sizeof(int32_t) * static_cast<size_t>(nnz) + 64);
  out->values.alloc(static_cast<size_t>(vb) * nnz + 64);
  QDSP_CUDA(cudaMemsetAsync(out->indices.get<char>() + sizeof(int32_t) * nnz, 0, 64, stream));
  QDSP_CUDA(cudaMemsetAsync(out->values.get<char>() + static_cast<size_t>(vb) * nnz, 0, 64, stream));
  QDSP_CUDA(cudaMemcpyAsync(out->indices.get(), idx_tmp, sizeof(int32_t) * nnz, cudaMemcpyDeviceToDevice, stream));
  QDSP_CUDA(cudaMemcpyAsync(out->values.get(), val_tmp, static_cast<size_t>(vb) * nnz, cudaMemcpyDeviceToDevice, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));  // scratch buffers die with this scope
  return out;
}

static std::unique_ptr<CsrMatrix> empty_matrix(int rows, int cols, int dtype, cudaStream_t stream) {
  auto out = std::make_unique<CsrMatrix>();
  out->allocate(rows, cols, 0, dtype);
  QDSP_CUDA(cudaMemsetAsync(out->indptr.get(), 0, out->indptr.bytes(), stream));
  QDSP_CUDA(cudaMemsetAsync(out->indices.get(), 0, out->indices.bytes(), stream));
  QDSP_CUDA(cudaMemsetAsync(out->values.get(), 0, out->values.bytes(), stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  return out;
}

static void check_pack_error(DeviceBuffer &err, cudaStream_t stream) {
  int h = 0;
  QDSP_CUDA(cudaMemcpyAsync(&h, err.get(), sizeof(int), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  QDSP_REQUIRE(h == 0, QDSP_ERR_INVALID, "triplet index out of range for the matrix shape");
}

Builder::Builder(int rows, int cols, int64_t max_triplets, int dtype, const std::string &storage_format)
    : rows_(rows), cols_(cols), max_triplets_(max_triplets), dtype_(dtype) {
  QDSP_REQUIRE(rows > 0 && cols > 0, QDSP_ERR_INVALID, "SparseMatrixBuilder: rows and cols must be positive");
  QDSP_REQUIRE(max_triplets >= 0, QDSP_ERR_INVALID, "SparseMatrixBuilder: negative max_num_triplets");
  QDSP_REQUIRE(dtype == QDSP_F32 || dtype == QDSP_F64, QDSP_ERR_INVALID, "Unsupported sparse matrix data type!");
  QDSP_REQUIRE(storage_format == "col_major" || storage_format == "row_major", QDSP_ERR_INVALID,
               "Unsupported sparse matrix storage format: " + storage_format);
  // the reference's ndarray shape is an int: 3*max+1 must fit (sparse_matrix.cpp:83)
  QDSP_REQUIRE(3 * max_triplets + 1 <= 2147483647LL, QDSP_ERR_INVALID,
               "SparseMatrixBuilder: 3*max_num_triplets+1 exceeds the int32 ndarray shape of the reference");
  row_major_ = storage_format == "row_major";
  buf_.alloc(slot_bytes() * (3 * static_cast<size_t>(max_triplets) + 1));
  QDSP_CUDA(cudaMemset(buf_.get(), 0, slot_bytes()));  // slot 0 = count
}

int64_t Builder::count(cudaStream_t stream) const {
  int64_t c = 0;
  if (dtype_ == QDSP_F64) {
    QDSP_CUDA(cudaMemcpyAsync(&c, buf_.get(), 8, cudaMemcpyDeviceToHost, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
  } else {
    int32_t c32 = 0;
    QDSP_CUDA(cudaMemcpyAsync(&c32, buf_.get(), 4, cudaMemcpyDeviceToHost, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
    c = c32;
  }
  return c;
}

void Builder::fill_from_host(const void *triplets, int64_t count, cudaStream_t stream) {
  QDSP_REQUIRE(count >= 0 && count <= max_triplets_, QDSP_ERR_INVALID, "fill_from_host: count exceeds max_num_triplets");
  const size_t sb = slot_bytes();
  if (count) QDSP_CUDA(cudaMemcpyAsync(buf_.get<char>() + sb, triplets, sb * 3 * count, cudaMemcpyHostToDevice, stream));
  int64_t c64 = count;
  int32_t c32 = static_cast<int32_t>(count);
  QDSP_CUDA(cudaMemcpyAsync(buf_.get(), sb == 8 ? static_cast<void *>(&c64) : static_cast<void *>(&c32), sb,
                            cudaMemcpyHostToDevice, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
}

// ---- bulk triplet emission ---------------------------------------------------------------------------
// The reference's fill kernels call insert_triplet_f32/f64 once per triplet: one atomic_add on slot 0 each
// (internal_functions.h:12-39), i.e. 2e8 serialised atomics on one address at C4.  A bulk append needs none: slot 0 is
// only READ while the triplets are written at slot0 + k, and a one-thread kernel bumps it afterwards (stream order
// separates the two).  Array order is kept, so a build from appended triplets is deterministic.
template <typename SlotT, typename BitsT>
__global__ void __launch_bounds__(CB_THREADS)
append_coo_kernel(SlotT *__restrict__ buf, const int32_t *__restrict__ row, const int32_t *__restrict__ col,
                  const BitsT *__restrict__ val, long long count, long long max_triplets, int negate) {
  const long long start = static_cast<long long>(buf[0]);
  if (start < 0 || start + count > max_triplets) return;  // overflow: nothing written, the bump makes build() fail
  constexpr BitsT SIGN = BitsT(1) << (8 * sizeof(BitsT) - 1);
  const long long stride = static_cast<long long>(gridDim.x) * CB_THREADS;
  for (long long k = static_cast<long long>(blockIdx.x) * CB_THREADS + threadIdx.x; k < count; k += stride) {
    SlotT *t = buf + 1 + 3 * (start + k);
    BitsT bits = val[k];
    if (negate) bits ^= SIGN;  // A[i, j] -= v inserts -v
    t[0] = static_cast<SlotT>(row[k]);
    t[1] = static_cast<SlotT>(col[k]);
    t[2] = static_cast<SlotT>(bits);
  }
}

// element e of k nodes contributes (nodes[e*k+a], nodes[e*k+b], mats[(e*k+a)*k+b]), a-major: one thread per triplet
template <typename SlotT, typename BitsT>
__global__ void __launch_bounds__(CB_THREADS)
append_elements_kernel(SlotT *__restrict__ buf, const int32_t *__restrict__ nodes, const BitsT *__restrict__ mats,
                       long long n_elems, int k, long long max_triplets) {
  const long long kk = static_cast<long long>(k) * k;
  const long long count = n_elems * kk;
  const long long start = static_cast<long long>(buf[0]);
  if (start < 0 || start + count > max_triplets) return;
  const long long stride = static_cast<long long>(gridDim.x) * CB_THREADS;
  for (long long t = static_cast<long long>(blockIdx.x) * CB_THREADS + threadIdx.x; t < count; t += stride) {
    const long long e = t / kk;
    const int ab = static_cast<int>(t - e * kk);
    const int a = ab / k, b = ab - a * k;
    SlotT *o = buf + 1 + 3 * (start + t);
    o[0] = static_cast<SlotT>(nodes[e * k + a]);
    o[1] = static_cast<SlotT>(nodes[e * k + b]);
    o[2] = static_cast<SlotT>(mats[t]);
  }
}

template <typename SlotT>
__global__ void append_bump_kernel(SlotT *buf, long long count) {
  buf[0] = static_cast<SlotT>(static_cast<long long>(buf[0]) + count);
}

static unsigned append_grid(int64_t count) {
  return static_cast<unsigned>(std::min<int64_t>((count + CB_THREADS - 1) / CB_THREADS,
                                                 static_cast<int64_t>(device_props().sm_count) * 16));
}

void Builder::append(const int32_t *row, const int32_t *col, const void *val, int64_t count, bool negate,
                     cudaStream_t stream) {
  QDSP_REQUIRE(count >= 0, QDSP_ERR_INVALID, "append: negative count");
  if (count == 0) return;
  QDSP_REQUIRE(row && col && val, QDSP_ERR_INVALID, "append: null pointer");
  QDSP_REQUIRE(count <= max_triplets_, QDSP_ERR_INVALID,
               "append: " + std::to_string(count) + " triplets exceed max_num_triplets " + std::to_string(max_triplets_));
  if (dtype_ == QDSP_F64) {
    append_coo_kernel<int64_t, uint64_t><<<append_grid(count), CB_THREADS, 0, stream>>>(
        buf_.get<int64_t>(), row, col, static_cast<const uint64_t *>(val), count, max_triplets_, negate ? 1 : 0);
    QDSP_LAUNCH_CHECK();
    append_bump_kernel<int64_t><<<1, 1, 0, stream>>>(buf_.get<int64_t>(), count);
  } else {
    append_coo_kernel<int32_t, uint32_t><<<append_grid(count), CB_THREADS, 0, stream>>>(
        buf_.get<int32_t>(), row, col, static_cast<const uint32_t *>(val), count, max_triplets_, negate ? 1 : 0);
    QDSP_LAUNCH_CHECK();
    append_bump_kernel<int32_t><<<1, 1, 0, stream>>>(buf_.get<int32_t>(), count);
  }
  QDSP_LAUNCH_CHECK();
}

void Builder::append_elements(const int32_t *nodes, const void *mats, int64_t n_elems, int k, cudaStream_t stream) {
  QDSP_REQUIRE(n_elems >= 0 && k > 0 && k <= 64, QDSP_ERR_INVALID, "append_elements: bad element count or nodes per element");
  if (n_elems == 0) return;
  QDSP_REQUIRE(nodes && mats, QDSP_ERR_INVALID, "append_elements: null pointer");
  const int64_t count = n_elems * k * k;
  QDSP_REQUIRE(count <= max_triplets_, QDSP_ERR_INVALID,
               "append_elements: " + std::to_string(count) + " triplets exceed max_num_triplets " +
                   std::to_string(max_triplets_));
  if (dtype_ == QDSP_F64) {
    append_elements_kernel<int64_t, uint64_t><<<append_grid(count), CB_THREADS, 0, stream>>>(
        buf_.get<int64_t>(), nodes, static_cast<const uint64_t *>(mats), n_elems, k, max_triplets_);
    QDSP_LAUNCH_CHECK();
    append_bump_kernel<int64_t><<<1, 1, 0, stream>>>(buf_.get<int64_t>(), count);
  } else {
    append_elements_kernel<int32_t, uint32_t><<<append_grid(count), CB_THREADS, 0, stream>>>(
        buf_.get<int32_t>(), nodes, static_cast<const uint32_t *>(mats), n_elems, k, max_triplets_);
    QDSP_LAUNCH_CHECK();
    append_bump_kernel<int32_t><<<1, 1, 0, stream>>>(buf_.get<int32_t>(), count);
  }
  QDSP_LAUNCH_CHECK();
}

void Builder::clear(cudaStream_t stream) {
  built_ = false;
  QDSP_CUDA(cudaMemsetAsync(buf_.get(), 0, slot_bytes(), stream));
}

std::unique_ptr<CsrMatrix> Builder::build(cudaStream_t stream) {
  QDSP_REQUIRE(!built_, QDSP_ERR_STATE, "SparseMatrixBuilder::build called twice (QD_ASSERT(built_ == false))");
  const int64_t n = count(stream);
  QDSP_REQUIRE(n >= 0 && n <= max_triplets_, QDSP_ERR_INVALID,
               "triplet count " + std::to_string(n) + " exceeds max_num_triplets " + std::to_string(max_triplets_));
  built_ = true;
  std::unique_ptr<CsrMatrix> out;
  if (n == 0) {
    out = empty_matrix(rows_, cols_, dtype_, stream);
  } else {
    const int col_bits = bits_for(cols_), row_bits = bits_for(rows_);
    const int key_bytes = (col_bits + row_bits <= 32) ? 4 : 8;
    const int vb = static_cast<int>(slot_bytes());
    DeviceBuffer keys = scratch_buffer(static_cast<size_t>(key_bytes) * n), vals = scratch_buffer(static_cast<size_t>(vb) * n),
                 err(sizeof(int));
    QDSP_CUDA(cudaMemsetAsync(err.get(), 0, sizeof(int), stream));
    const unsigned grid = static_cast<unsigned>((n + CB_THREADS - 1) / CB_THREADS);
    const char *trip = buf_.get<char>() + slot_bytes();
    if (vb == 4 && key_bytes == 4)
      pack_triplets_kernel<int32_t, uint32_t, uint32_t><<<grid, CB_THREADS, 0, stream>>>(
          reinterpret_cast<const int32_t *>(trip), n, rows_, cols_, col_bits, keys.get<uint32_t>(), vals.get<uint32_t>(), err.get<int>());
    else if (vb == 4)
      pack_triplets_kernel<int32_t, uint64_t, uint32_t><<<grid, CB_THREADS, 0, stream>>>(
          reinterpret_cast<const int32_t *>(trip), n, rows_, cols_, col_bits, keys.get<uint64_t>(), vals.get<uint32_t>(), err.get<int>());
    else if (key_bytes == 4)
      pack_triplets_kernel<int64_t, uint32_t, uint64_t><<<grid, CB_THREADS, 0, stream>>>(
          reinterpret_cast<const int64_t *>(trip), n, rows_, cols_, col_bits, keys.get<uint32_t>(), vals.get<uint64_t>(), err.get<int>());
    else
      pack_triplets_kernel<int64_t, uint64_t, uint64_t><<<grid, CB_THREADS, 0, stream>>>(
          reinterpret_cast<const int64_t *>(trip), n, rows_, cols_, col_bits, keys.get<uint64_t>(), vals.get<uint64_t>(), err.get<int>());
    QDSP_LAUNCH_CHECK();
    check_pack_error(err, stream);
    out = csr_from_packed(rows_, cols_, n, dtype_, key_bytes, col_bits, keys, vals, stream);
  }
  out->row_major_order = row_major_;
  clear(stream);  // like build_template / build_cuda: the builder can be refilled
  return out;
}

void Builder::print_triplets() const {
  const int64_t n = count(nullptr);
  std::printf("n=%d, m=%d, num_triplets=%lld (max=%lld)\n", rows_, cols_, (long long)n, (long long)max_triplets_);
  if (n <= 0) return;
  const size_t sb = slot_bytes();
  std::vector<char> h(sb * 3 * n);
  QDSP_CUDA(cudaMemcpy(h.data(), buf_.get<char>() + sb, h.size(), cudaMemcpyDeviceToHost));
  for (int64_t i = 0; i < n; ++i) {
    if (sb == 4) {
      const int32_t *t = reinterpret_cast<const int32_t *>(h.data()) + 3 * i;
      float v; std::memcpy(&v, &t[2], 4);
      std::printf("[%d, %d] = %g\n", t[0], t[1], v);
    } else {
      const int64_t *t = reinterpret_cast<const int64_t *>(h.data()) + 3 * i;
      double v; std::memcpy(&v, &t[2], 8);
      std::printf("[%lld, %lld] = %g\n", (long long)t[0], (long long)t[1], v);
    }
  }
}

std::unique_ptr<CsrMatrix> CsrMatrix::from_device_coo(int rows, int cols, int64_t n, const int32_t *row,
                                                      const int32_t *col, const void *val, int dtype,
                                                      cudaStream_t stream) {
  QDSP_REQUIRE(rows > 0 && cols > 0 && n >= 0, QDSP_ERR_INVALID, "build_csr_from_coo: bad shape");
  if (n == 0) return empty_matrix(rows, cols, dtype, stream);
  const int col_bits = bits_for(cols), row_bits = bits_for(rows);
  const int key_bytes = (col_bits + row_bits <= 32) ? 4 : 8;
  const int vb = dtype == QDSP_F64 ? 8 : 4;
  DeviceBuffer keys = scratch_buffer(static_cast<size_t>(key_bytes) * n), vals = scratch_buffer(static_cast<size_t>(vb) * n),
               err(sizeof(int));
  QDSP_CUDA(cudaMemsetAsync(err.get(), 0, sizeof(int), stream));
  const unsigned grid = static_cast<unsigned>((n + CB_THREADS - 1) / CB_THREADS);
  if (vb == 4 && key_bytes == 4)
    pack_coo_kernel<uint32_t, uint32_t><<<grid, CB_THREADS, 0, stream>>>(row, col, static_cast<const uint32_t *>(val), n, rows, cols, col_bits, keys.get<uint32_t>(), vals.get<uint32_t>(), err.get<int>());
  else if (vb == 4)
    pack_coo_kernel<uint64_t, uint32_t><<<grid, CB_THREADS, 0, stream>>>(row, col, static_cast<const uint32_t *>(val), n, rows, cols, col_bits, keys.get<uint64_t>(), vals.get<uint32_t>(), err.get<int>());
  else if (key_bytes == 4)
    pack_coo_kernel<uint32_t, uint64_t><<<grid, CB_THREADS, 0, stream>>>(row, col, static_cast<const uint64_t *>(val), n, rows, cols, col_bits, keys.get<uint32_t>(), vals.get<uint64_t>(), err.get<int>());
  else
    pack_coo_kernel<uint64_t, uint64_t><<<grid, CB_THREADS, 0, stream>>>(row, col, static_cast<const uint64_t *>(val), n, rows, cols, col_bits, keys.get<uint64_t>(), vals.get<uint64_t>(), err.get<int>());
  QDSP_LAUNCH_CHECK();
  check_pack_error(err, stream);
  return csr_from_packed(rows, cols, n, dtype, key_bytes, col_bits, keys, vals, stream);
}

std::unique_ptr<CsrMatrix> CsrMatrix::from_device_csr(int rows, int cols, int64_t nnz, const int32_t *indptr,
                                                      const int32_t *indices, const void *values, int dtype,
                                                      bool row_major_order, cudaStream_t stream) {
  auto out = std::make_unique<CsrMatrix>();
  out->allocate(rows, cols, nnz, dtype);
  out->row_major_order = row_major_order;
  QDSP_CUDA(cudaMemsetAsync(out->indices.get(), 0, out->indices.bytes(), stream));
  QDSP_CUDA(cudaMemsetAsync(out->values.get(), 0, out->values.bytes(), stream));
  QDSP_CUDA(cudaMemcpyAsync(out->indptr.get(), indptr, sizeof(int32_t) * (static_cast<size_t>(rows) + 1), cudaMemcpyDefault, stream));
  if (nnz) {
    QDSP_CUDA(cudaMemcpyAsync(out->indices.get(), indices, sizeof(int32_t) * nnz, cudaMemcpyDefault, stream));
    QDSP_CUDA(cudaMemcpyAsync(out->values.get(), values, out->value_bytes() * nnz, cudaMemcpyDefault, stream));
  }
  QDSP_CUDA(cudaStreamSynchronize(stream));
  return out;
}

}  // namespace qdsp
