// csr_ops.cu -- the remaining CuSparseMatrix algebra (quadrants/program/sparse_matrix.cpp:413-652): A+B / A-B
// (cusparseScsrgeam2), alpha*A, A^T (cusparseCsr2cscEx2), A*B (cusparseSpGEMM).  Not a performance target of any
// BASELINE config: each one is expressed as "emit COO triplets on the device, then run the same
// sort + collapse pipeline as the builder", so they inherit its kernels and its determinism (duplicates summed in a
// fixed order).
#include <cstdlib>
#include <vector>

#include "host.hpp"

namespace qdsp {

constexpr int OP_THREADS = 256;

// row index of every stored entry (CSR -> COO expansion), warp per row
__global__ void __launch_bounds__(OP_THREADS)
expand_rows_kernel(int rows, const int *__restrict__ indptr, int *__restrict__ coo_row, long long offset) {
  const int warp = (blockIdx.x * OP_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  for (int k = indptr[warp] + lane; k < indptr[warp + 1]; k += 32) coo_row[offset + k] = warp;
}

template <typename T>
__global__ void __launch_bounds__(OP_THREADS)
scale_copy_kernel(long long n, const T *__restrict__ in, T alpha, T *__restrict__ out) {
  const long long i = static_cast<long long>(blockIdx.x) * OP_THREADS + threadIdx.x;
  if (i < n) out[i] = alpha * in[i];
}

__global__ void __launch_bounds__(OP_THREADS)
copy_int_kernel(long long n, const int *__restrict__ in, int *__restrict__ out) {
  const long long i = static_cast<long long>(blockIdx.x) * OP_THREADS + threadIdx.x;
  if (i < n) out[i] = in[i];
}

// SpGEMM by expansion: entry k of A = (i, j, a) contributes a * B[j, :]; cnt[k] = len(B row j)
__global__ void __launch_bounds__(OP_THREADS)
spgemm_count_kernel(long long nnz_a, const int *__restrict__ a_idx, const int *__restrict__ b_ptr,
                    int *__restrict__ cnt) {
  const long long k = static_cast<long long>(blockIdx.x) * OP_THREADS + threadIdx.x;
  if (k < nnz_a) cnt[k] = b_ptr[a_idx[k] + 1] - b_ptr[a_idx[k]];
}

template <typename T>
__global__ void __launch_bounds__(OP_THREADS)
spgemm_expand_kernel(long long nnz_a, const int *__restrict__ a_row, const int *__restrict__ a_idx,
                     const T *__restrict__ a_val, const int *__restrict__ b_ptr, const int *__restrict__ b_idx,
                     const T *__restrict__ b_val, const int *__restrict__ incl /* inclusive scan of cnt */,
                     int *__restrict__ o_row, int *__restrict__ o_col, T *__restrict__ o_val) {
  const long long k = static_cast<long long>(blockIdx.x) * OP_THREADS + threadIdx.x;
  if (k >= nnz_a) return;
  const int j = a_idx[k];
  const int len = b_ptr[j + 1] - b_ptr[j];
  int o = incl[k] - len;
  for (int q = b_ptr[j]; q < b_ptr[j + 1]; ++q, ++o) {
    o_row[o] = a_row[k];
    o_col[o] = b_idx[q];
    o_val[o] = a_val[k] * b_val[q];
  }
}

__device__ __forceinline__ float add_unfused(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_unfused(double a, double b) { return __dadd_rn(a, b); }

static unsigned grid_for(long long n) { return static_cast<unsigned>(std::max<long long>(1, (n + OP_THREADS - 1) / OP_THREADS)); }

static void expand_rows(const CsrMatrix &m, int *coo_row, long long offset, cudaStream_t s) {
  if (m.rows == 0) return;
  expand_rows_kernel<<<grid_for(static_cast<long long>(m.rows) * 32), OP_THREADS, 0, s>>>(m.rows, m.indptr.get<int>(), coo_row, offset);
  QDSP_LAUNCH_CHECK();
}

template <typename T>
static void scale_copy(long long n, const T *in, double alpha, T *out, cudaStream_t s) {
  if (n == 0) return;
  scale_copy_kernel<T><<<grid_for(n), OP_THREADS, 0, s>>>(n, in, static_cast<T>(alpha), out);
  QDSP_LAUNCH_CHECK();
}

std::unique_ptr<CsrMatrix> CsrMatrix::scale(double alpha, cudaStream_t stream) const {
  auto out = from_device_csr(rows, cols, nnz, indptr.get<int>(), indices.get<int>(), values.get(), dtype, row_major_order, stream);
  if (dtype == QDSP_F32) scale_copy<float>(nnz, values.get<float>(), alpha, out->values.get<float>(), stream);
  else scale_copy<double>(nnz, values.get<double>(), alpha, out->values.get<double>(), stream);
  QDSP_CUDA(cudaStreamSynchronize(stream));
  return out;
}

static std::unique_ptr<CsrMatrix> add_by_sort(const CsrMatrix &a, const CsrMatrix &b, double alpha, double beta, cudaStream_t stream) {
    const long long nnz = a.nnz;
  const int rows = a.rows, cols = a.cols, dtype = a.dtype;
  const long long n = nnz + b.nnz;
  const size_t vb = a.value_bytes();
  DeviceBuffer r = scratch_buffer(sizeof(int) * n), c = scratch_buffer(sizeof(int) * n), v = scratch_buffer(vb * n);
  expand_rows(a, r.get<int>(), 0, stream);
  expand_rows(b, r.get<int>(), nnz, stream);
  if (nnz) { copy_int_kernel<<<grid_for(nnz), OP_THREADS, 0, stream>>>(nnz, a.indices.get<int>(), c.get<int>()); QDSP_LAUNCH_CHECK(); }
  if (b.nnz) { copy_int_kernel<<<grid_for(b.nnz), OP_THREADS, 0, stream>>>(b.nnz, b.indices.get<int>(), c.get<int>() + nnz); QDSP_LAUNCH_CHECK(); }
  if (dtype == QDSP_F32) {
    scale_copy<float>(nnz, a.values.get<float>(), alpha, v.get<float>(), stream);
    scale_copy<float>(b.nnz, b.values.get<float>(), beta, v.get<float>() + nnz, stream);
  } else {
    scale_copy<double>(nnz, a.values.get<double>(), alpha, v.get<double>(), stream);
    scale_copy<double>(b.nnz, b.values.get<double>(), beta, v.get<double>() + nnz, stream);
  }
  auto out = CsrMatrix::from_device_coo(rows, cols, n, r.get<int>(), c.get<int>(), v.get(), dtype, stream);
  out->row_major_order = a.row_major_order;
  return out;
}


// ---- alpha A + beta B by row-wise merge (cusparseScsrgeam2, sparse_matrix.cpp:413-509) ------------------------------------------
// Rows of canonical CSR are sorted and duplicate-free, so the union of two rows needs no sort: a two-pointer merge per row,
// once to count and once to fill -- two passes over the operands and one write of the result instead of the builder's
// sort of |A| + |B| triplets.  An entry present in both is alpha*a + beta*b with the products rounded first (the order the
// sort path sums in, and scipy's); structural zeros from cancellation are kept, like csrgeam2.  One thread per row (rows
// are short where this matters; neighbouring threads read neighbouring rows).  `unsorted` is raised when a row is not
// strictly ascending -- matrices adopted through qdsp_csr_from_device_csr need not be canonical -- and the caller then
// takes the sort path above.
__global__ void __launch_bounds__(OP_THREADS)
geam_count_kernel(int rows, const int *__restrict__ a_ptr, const int *__restrict__ a_idx, const int *__restrict__ b_ptr,
                  const int *__restrict__ b_idx, int *__restrict__ out_cnt, int *__restrict__ unsorted) {
  const int r = blockIdx.x * OP_THREADS + threadIdx.x;
  if (r >= rows) return;
  int i = a_ptr[r], j = b_ptr[r];
  const int ie = a_ptr[r + 1], je = b_ptr[r + 1];
  int n = 0, last = -1;
  bool bad = false;
  while (i < ie && j < je) {
    const int ca = a_idx[i], cb = b_idx[j];
    const int c = min(ca, cb);
    bad |= c <= last;
    last = c;
    i += ca <= cb;
    j += cb <= ca;
    ++n;
  }
  for (; i < ie; ++i, ++n) { bad |= a_idx[i] <= last; last = a_idx[i]; }
  for (; j < je; ++j, ++n) { bad |= b_idx[j] <= last; last = b_idx[j]; }
  out_cnt[r] = n;
  if (bad) *unsorted = 1;
}

template <typename T>
__global__ void __launch_bounds__(OP_THREADS)
geam_fill_kernel(int rows, const int *__restrict__ a_ptr, const int *__restrict__ a_idx, const T *__restrict__ a_val,
                 const int *__restrict__ b_ptr, const int *__restrict__ b_idx, const T *__restrict__ b_val, T alpha, T beta,
                 const int *__restrict__ o_ptr, int *__restrict__ o_idx, T *__restrict__ o_val) {
  const int r = blockIdx.x * OP_THREADS + threadIdx.x;
  if (r >= rows) return;
  int i = a_ptr[r], j = b_ptr[r], o = o_ptr[r];
  const int ie = a_ptr[r + 1], je = b_ptr[r + 1];
  while (i < ie && j < je) {
    const int ca = a_idx[i], cb = b_idx[j];
    T v;
    if (ca < cb) { v = alpha * a_val[i]; ++i; }
    else if (cb < ca) { v = beta * b_val[j]; ++j; }
    else {
      const T pa = alpha * a_val[i], pb = beta * b_val[j];
      v = add_unfused(pa, pb);
      ++i; ++j;
    }
    o_idx[o] = min(ca, cb);
    o_val[o] = v;
    ++o;
  }
  for (; i < ie; ++i, ++o) { o_idx[o] = a_idx[i]; o_val[o] = alpha * a_val[i]; }
  for (; j < je; ++j, ++o) { o_idx[o] = b_idx[j]; o_val[o] = beta * b_val[j]; }
}

std::unique_ptr<CsrMatrix> CsrMatrix::add(const CsrMatrix &b, double alpha, double beta, cudaStream_t stream) const {
  QDSP_REQUIRE(rows == b.rows && cols == b.cols, QDSP_ERR_INVALID, "sparse matrix addition: dimension mismatch");
  QDSP_REQUIRE(dtype == b.dtype, QDSP_ERR_INVALID, "sparse matrix addition: dtype mismatch");
  auto out = std::make_unique<CsrMatrix>();
  out->rows = rows; out->cols = cols; out->dtype = dtype;
  out->indptr.alloc(sizeof(int32_t) * (static_cast<size_t>(rows) + 1) + 64);
  DeviceBuffer flag(sizeof(int)), scan_tmp(scan_tmp_bytes(rows));
  QDSP_CUDA(cudaMemsetAsync(flag.get(), 0, sizeof(int), stream));
  QDSP_CUDA(cudaMemsetAsync(out->indptr.get(), 0, sizeof(int32_t), stream));
  int *optr = out->indptr.get<int>();
  geam_count_kernel<<<grid_for(rows), OP_THREADS, 0, stream>>>(rows, indptr.get<int>(), indices.get<int>(), b.indptr.get<int>(),
                                                              b.indices.get<int>(), optr + 1, flag.get<int>());
  QDSP_LAUNCH_CHECK();
  inclusive_scan_i32(optr + 1, rows, scan_tmp.get(), scan_tmp.bytes(), stream);
  int h[2] = {0, 0};
  QDSP_CUDA(cudaMemcpyAsync(&h[0], flag.get(), sizeof(int), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaMemcpyAsync(&h[1], optr + rows, sizeof(int), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  if (h[0] != 0 || h[1] < 0) return add_by_sort(*this, b, alpha, beta, stream);  // not canonical, or > 2^31 entries
  out->nnz = h[1];
  out->indices.alloc(sizeof(int32_t) * static_cast<size_t>(out->nnz) + 64);
  out->values.alloc(value_bytes() * static_cast<size_t>(out->nnz) + 64);
  QDSP_CUDA(cudaMemsetAsync(out->indices.get<char>() + sizeof(int32_t) * out->nnz, 0, 64, stream));
  QDSP_CUDA(cudaMemsetAsync(out->values.get<char>() + value_bytes() * out->nnz, 0, 64, stream));
  if (dtype == QDSP_F32)
    geam_fill_kernel<float><<<grid_for(rows), OP_THREADS, 0, stream>>>(rows, indptr.get<int>(), indices.get<int>(), values.get<float>(), b.indptr.get<int>(), b.indices.get<int>(), b.values.get<float>(), static_cast<float>(alpha), static_cast<float>(beta), optr, out->indices.get<int>(), out->values.get<float>());
  else
    geam_fill_kernel<double><<<grid_for(rows), OP_THREADS, 0, stream>>>(rows, indptr.get<int>(), indices.get<int>(), values.get<double>(), b.indptr.get<int>(), b.indices.get<int>(), b.values.get<double>(), alpha, beta, optr, out->indices.get<int>(), out->values.get<double>());
  QDSP_LAUNCH_CHECK();
  QDSP_CUDA(cudaStreamSynchronize(stream));
  out->row_major_order = row_major_order;
  return out;
}

std::unique_ptr<CsrMatrix> CsrMatrix::transpose(cudaStream_t stream) const {
  DeviceBuffer r = scratch_buffer(sizeof(int) * std::max<int64_t>(nnz, 1));
  expand_rows(*this, r.get<int>(), 0, stream);
  // (row, col, v) -> (col, row, v); the stable sort keeps each output row ordered by the old row index
  auto out = from_device_coo(cols, rows, nnz, indices.get<int>(), r.get<int>(), values.get(), dtype, stream);
  out->row_major_order = row_major_order;
  return out;
}

// partial products of every row of A*B in 64 bits (the i32 scans below work on one bounded chunk at a time)
__global__ void __launch_bounds__(OP_THREADS)
spgemm_row_products_kernel(int rows, const int *__restrict__ a_ptr, const int *__restrict__ a_idx,
                           const int *__restrict__ b_ptr, long long *__restrict__ out) {
  const int warp = (blockIdx.x * OP_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  long long t = 0;
  for (int k = a_ptr[warp] + lane; k < a_ptr[warp + 1]; k += 32) t += b_ptr[a_idx[k] + 1] - b_ptr[a_idx[k]];
  t = warp_sum(t);
  if (lane == 0) out[warp] = t;
}

// dst[i] = src[i] + add  (row pointers of one chunk into the final indptr)
__global__ void __launch_bounds__(OP_THREADS)
offset_copy_kernel(long long n, const int *__restrict__ src, int add, int *__restrict__ dst) {
  const long long i = static_cast<long long>(blockIdx.x) * OP_THREADS + threadIdx.x;
  if (i < n) dst[i] = src[i] + add;
}

// Partial products one chunk of rows may expand to.  Bounds the intermediate of A*B (16-20 bytes per product plus the
// sort's ping-pong buffers): 2^26 products ~ 2.5 GB.  QDSP_SPGEMM_CHUNK overrides (tests force several chunks with it).
static long long spgemm_chunk_products() {
  if (const char *e = std::getenv("QDSP_SPGEMM_CHUNK")) {
    const long long v = std::atoll(e);
    if (v > 0) return std::min<long long>(v, 1ll << 29);
  }
  return 1ll << 26;
}

// A*B (cusparseSpGEMM, sparse_matrix.cpp:522-598) by expand - sort - collapse, in row chunks: the rows of A are cut so
// that no chunk expands to more than spgemm_chunk_products() partial products (a single heavier row forms its own chunk,
// up to the sort's 2^30 limit); each chunk becomes the CSR of its rows through the same deterministic pipeline as the
// builder, and the chunks are concatenated.  The result does not depend on the chunking: every output entry is the sum of
// its partial products in (A entry, B entry) order.
std::unique_ptr<CsrMatrix> CsrMatrix::matmul(const CsrMatrix &b, cudaStream_t stream) const {
  QDSP_REQUIRE(cols == b.rows, QDSP_ERR_INVALID, "sparse matrix product: dimension mismatch");
  QDSP_REQUIRE(dtype == b.dtype, QDSP_ERR_INVALID, "sparse matrix product: dtype mismatch");
  if (nnz == 0 || b.nnz == 0) {
    DeviceBuffer dummy(16);
    return from_device_coo(rows, b.cols, 0, dummy.get<int>(), dummy.get<int>(), dummy.get(), dtype, stream);
  }
  // symbolic pass: products per row, then the chunk boundaries on the host
  DeviceBuffer d_rowprod(sizeof(long long) * rows);
  spgemm_row_products_kernel<<<grid_for(static_cast<long long>(rows) * 32), OP_THREADS, 0, stream>>>(
      rows, indptr.get<int>(), indices.get<int>(), b.indptr.get<int>(), d_rowprod.get<long long>());
  QDSP_LAUNCH_CHECK();
  std::vector<long long> rowprod(rows);
  std::vector<int> a_ptr(rows + 1);
  QDSP_CUDA(cudaMemcpyAsync(rowprod.data(), d_rowprod.get(), sizeof(long long) * rows, cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaMemcpyAsync(a_ptr.data(), indptr.get(), sizeof(int) * (rows + 1), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  const long long bound = spgemm_chunk_products();
  std::vector<int> cuts{0};
  std::vector<long long> chunk_total;
  long long acc = 0;
  for (int r = 0; r < rows; ++r) {
    // the reference (cusparseSpGEMM) errors out as well when its work buffers do not fit
    QDSP_REQUIRE(rowprod[r] < (1ll << 30), QDSP_ERR_UNSUPPORTED,
                 "sparse matrix product: one row expands to more than 2^30 partial products");
    if (acc > 0 && acc + rowprod[r] > bound) {
      cuts.push_back(r);
      chunk_total.push_back(acc);
      acc = 0;
    }
    acc += rowprod[r];
  }
  cuts.push_back(rows);
  chunk_total.push_back(acc);
  const size_t vb = value_bytes();
  std::vector<std::unique_ptr<CsrMatrix>> parts;
  for (size_t ch = 0; ch + 1 < cuts.size(); ++ch) {
    const int r0 = cuts[ch], r1 = cuts[ch + 1];
    const long long k0 = a_ptr[r0], nk = a_ptr[r1] - k0;
    const int total = static_cast<int>(chunk_total[ch]);
    if (nk == 0 || total == 0) {
      DeviceBuffer dummy(16);
      parts.push_back(from_device_coo(r1 - r0, b.cols, 0, dummy.get<int>(), dummy.get<int>(), dummy.get(), dtype, stream));
      continue;
    }
    DeviceBuffer a_row = scratch_buffer(sizeof(int) * nk), cnt = scratch_buffer(sizeof(int) * nk + 64), scan_tmp(scan_tmp_bytes(nk));
    expand_rows_kernel<<<grid_for(static_cast<long long>(r1 - r0) * 32), OP_THREADS, 0, stream>>>(
        r1 - r0, indptr.get<int>() + r0, a_row.get<int>(), -k0);
    QDSP_LAUNCH_CHECK();
    spgemm_count_kernel<<<grid_for(nk), OP_THREADS, 0, stream>>>(nk, indices.get<int>() + k0, b.indptr.get<int>(), cnt.get<int>());
    QDSP_LAUNCH_CHECK();
    inclusive_scan_i32(cnt.get<int>(), nk, scan_tmp.get(), scan_tmp.bytes(), stream);
    DeviceBuffer o_row = scratch_buffer(sizeof(int) * total), o_col = scratch_buffer(sizeof(int) * total), o_val = scratch_buffer(vb * total);
    if (dtype == QDSP_F32)
      spgemm_expand_kernel<float><<<grid_for(nk), OP_THREADS, 0, stream>>>(nk, a_row.get<int>(), indices.get<int>() + k0, values.get<float>() + k0, b.indptr.get<int>(), b.indices.get<int>(), b.values.get<float>(), cnt.get<int>(), o_row.get<int>(), o_col.get<int>(), o_val.get<float>());
    else
      spgemm_expand_kernel<double><<<grid_for(nk), OP_THREADS, 0, stream>>>(nk, a_row.get<int>(), indices.get<int>() + k0, values.get<double>() + k0, b.indptr.get<int>(), b.indices.get<int>(), b.values.get<double>(), cnt.get<int>(), o_row.get<int>(), o_col.get<int>(), o_val.get<double>());
    QDSP_LAUNCH_CHECK();
    parts.push_back(from_device_coo(r1 - r0, b.cols, total, o_row.get<int>(), o_col.get<int>(), o_val.get(), dtype, stream));
  }
  if (parts.size() == 1) {
    parts[0]->row_major_order = row_major_order;
    return std::move(parts[0]);
  }
  long long out_nnz = 0;
  for (auto &p : parts) out_nnz += p->nnz;
  QDSP_REQUIRE(out_nnz < (1ll << 31), QDSP_ERR_UNSUPPORTED, "sparse matrix product: more than 2^31 non-zeros (INDEX_32I)");
  auto out = std::make_unique<CsrMatrix>();
  out->allocate(rows, b.cols, out_nnz, dtype);
  long long at = 0;
  for (size_t ch = 0; ch < parts.size(); ++ch) {
    const CsrMatrix &p = *parts[ch];
    const int r0 = cuts[ch];
    offset_copy_kernel<<<grid_for(p.rows + 1), OP_THREADS, 0, stream>>>(p.rows + 1, p.indptr.get<int>(), static_cast<int>(at),
                                                                      out->indptr.get<int>() + r0);
    QDSP_LAUNCH_CHECK();
    if (p.nnz) {
      QDSP_CUDA(cudaMemcpyAsync(out->indices.get<int>() + at, p.indices.get(), sizeof(int) * p.nnz, cudaMemcpyDeviceToDevice, stream));
      QDSP_CUDA(cudaMemcpyAsync(static_cast<char *>(out->values.get()) + vb * at, p.values.get(), vb * p.nnz, cudaMemcpyDeviceToDevice, stream));
    }
    at += p.nnz;
  }
  QDSP_CUDA(cudaStreamSynchronize(stream));  // the chunk matrices are released on return
  out->row_major_order = row_major_order;
  return out;
}

}  // namespace qdsp
