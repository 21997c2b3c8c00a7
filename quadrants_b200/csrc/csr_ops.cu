// csr_ops.cu -- the remaining CuSparseMatrix algebra (quadrants/program/sparse_matrix.cpp:413-652): A+B / A-B
// (cusparseScsrgeam2), alpha*A, A^T (cusparseCsr2cscEx2), A*B (cusparseSpGEMM).  Not a performance target of any
// BASELINE config: each one is expressed as "emit COO triplets on the device, then run the same
// sort + collapse pipeline as the builder", so they inherit its kernels and its determinism (duplicates summed in a
// fixed order).
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

std::unique_ptr<CsrMatrix> CsrMatrix::add(const CsrMatrix &b, double alpha, double beta, cudaStream_t stream) const {
  QDSP_REQUIRE(rows == b.rows && cols == b.cols, QDSP_ERR_INVALID, "sparse matrix addition: dimension mismatch");
  QDSP_REQUIRE(dtype == b.dtype, QDSP_ERR_INVALID, "sparse matrix addition: dtype mismatch");
  const long long n = nnz + b.nnz;
  const size_t vb = value_bytes();
  DeviceBuffer r(sizeof(int) * n), c(sizeof(int) * n), v(vb * n);
  expand_rows(*this, r.get<int>(), 0, stream);
  expand_rows(b, r.get<int>(), nnz, stream);
  if (nnz) { copy_int_kernel<<<grid_for(nnz), OP_THREADS, 0, stream>>>(nnz, indices.get<int>(), c.get<int>()); QDSP_LAUNCH_CHECK(); }
  if (b.nnz) { copy_int_kernel<<<grid_for(b.nnz), OP_THREADS, 0, stream>>>(b.nnz, b.indices.get<int>(), c.get<int>() + nnz); QDSP_LAUNCH_CHECK(); }
  if (dtype == QDSP_F32) {
    scale_copy<float>(nnz, values.get<float>(), alpha, v.get<float>(), stream);
    scale_copy<float>(b.nnz, b.values.get<float>(), beta, v.get<float>() + nnz, stream);
  } else {
    scale_copy<double>(nnz, values.get<double>(), alpha, v.get<double>(), stream);
    scale_copy<double>(b.nnz, b.values.get<double>(), beta, v.get<double>() + nnz, stream);
  }
  auto out = from_device_coo(rows, cols, n, r.get<int>(), c.get<int>(), v.get(), dtype, stream);
  out->row_major_order = row_major_order;
  return out;
}

std::unique_ptr<CsrMatrix> CsrMatrix::transpose(cudaStream_t stream) const {
  DeviceBuffer r(sizeof(int) * std::max<int64_t>(nnz, 1));
  expand_rows(*this, r.get<int>(), 0, stream);
  // (row, col, v) -> (col, row, v); the stable sort keeps each output row ordered by the old row index
  auto out = from_device_coo(cols, rows, nnz, indices.get<int>(), r.get<int>(), values.get(), dtype, stream);
  out->row_major_order = row_major_order;
  return out;
}

// 64-bit total of the per-entry product counts (the i32 scan below wraps silently past 2^31)
__global__ void sum_counts_kernel(int64_t n, const int *__restrict__ cnt, unsigned long long *__restrict__ total) {
  unsigned long long t = 0;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    t += static_cast<unsigned long long>(cnt[i]);
  t = warp_sum(t);
  if (lane_id() == 0 && t) atomicAdd(total, t);
}

std::unique_ptr<CsrMatrix> CsrMatrix::matmul(const CsrMatrix &b, cudaStream_t stream) const {
  QDSP_REQUIRE(cols == b.rows, QDSP_ERR_INVALID, "sparse matrix product: dimension mismatch");
  QDSP_REQUIRE(dtype == b.dtype, QDSP_ERR_INVALID, "sparse matrix product: dtype mismatch");
  if (nnz == 0 || b.nnz == 0) {
    DeviceBuffer dummy(16);
    return from_device_coo(rows, b.cols, 0, dummy.get<int>(), dummy.get<int>(), dummy.get(), dtype, stream);
  }
  DeviceBuffer a_row(sizeof(int) * nnz), cnt(sizeof(int) * nnz + 64), scan_tmp(scan_tmp_bytes(nnz));
  expand_rows(*this, a_row.get<int>(), 0, stream);
  spgemm_count_kernel<<<grid_for(nnz), OP_THREADS, 0, stream>>>(nnz, indices.get<int>(), b.indptr.get<int>(), cnt.get<int>());
  QDSP_LAUNCH_CHECK();
  DeviceBuffer total64(sizeof(unsigned long long));
  QDSP_CUDA(cudaMemsetAsync(total64.get(), 0, sizeof(unsigned long long), stream));
  sum_counts_kernel<<<std::min<int64_t>(grid_for(nnz), 1184), OP_THREADS, 0, stream>>>(nnz, cnt.get<int>(), total64.get<unsigned long long>());
  QDSP_LAUNCH_CHECK();
  unsigned long long total_ull = 0;
  QDSP_CUDA(cudaMemcpyAsync(&total_ull, total64.get(), sizeof(total_ull), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  // the expansion is sorted by the onesweep sort (< 2^30 items); the reference (cusparseSpGEMM) errors out as well when
  // its work buffers do not fit
  QDSP_REQUIRE(total_ull < (1ull << 30), QDSP_ERR_UNSUPPORTED,
               "sparse matrix product: more than 2^30 partial products (intermediate does not fit one sort)");
  const int total = static_cast<int>(total_ull);
  inclusive_scan_i32(cnt.get<int>(), nnz, scan_tmp.get(), scan_tmp.bytes(), stream);
  const size_t vb = value_bytes();
  DeviceBuffer o_row(sizeof(int) * std::max(total, 1)), o_col(sizeof(int) * std::max(total, 1)), o_val(vb * std::max(total, 1));
  if (dtype == QDSP_F32)
    spgemm_expand_kernel<float><<<grid_for(nnz), OP_THREADS, 0, stream>>>(nnz, a_row.get<int>(), indices.get<int>(), values.get<float>(), b.indptr.get<int>(), b.indices.get<int>(), b.values.get<float>(), cnt.get<int>(), o_row.get<int>(), o_col.get<int>(), o_val.get<float>());
  else
    spgemm_expand_kernel<double><<<grid_for(nnz), OP_THREADS, 0, stream>>>(nnz, a_row.get<int>(), indices.get<int>(), values.get<double>(), b.indptr.get<int>(), b.indices.get<int>(), b.values.get<double>(), cnt.get<int>(), o_row.get<int>(), o_col.get<int>(), o_val.get<double>());
  QDSP_LAUNCH_CHECK();
  auto out = from_device_coo(rows, b.cols, total, o_row.get<int>(), o_col.get<int>(), o_val.get(), dtype, stream);
  out->row_major_order = row_major_order;
  return out;
}

}  // namespace qdsp
