// spmv.cu -- CsrMatrix: SpMV launcher and the small host-visible accessors of CuSparseMatrix
// (quadrants/program/sparse_matrix.cpp:654-780).
#include <cstring>
#include <fstream>
#include <vector>

#include "host.hpp"
#include "spmv_kernels.cuh"

namespace qdsp {

std::atomic<int64_t> g_launch_count{0};

const DeviceProps &device_props() {
  static thread_local DeviceProps cached;
  int dev = 0;
  QDSP_CUDA(cudaGetDevice(&dev));
  if (cached.device != dev) {
    cudaDeviceProp p;
    QDSP_CUDA(cudaGetDeviceProperties(&p, dev));
    cached.device = dev;
    cached.sm_count = p.multiProcessorCount;
    cached.cc_major = p.major;
    cached.cc_minor = p.minor;
    cached.l2_bytes = static_cast<size_t>(p.l2CacheSize);
    cached.hbm_bytes = p.totalGlobalMem;
  }
  return cached;
}

template <typename T, bool ROW_MAJOR_ORDER>
__global__ void __launch_bounds__(SPMV_ROWS)
spmv_kernel(int rows, const int *__restrict__ indptr, const int *__restrict__ indices, const T *__restrict__ values,
            const T *__restrict__ x, T *__restrict__ y) {
  __shared__ SpmvSmem<T> S;
  spmv_cta<T, ROW_MAJOR_ORDER>(S, rows, indptr, indices, values, x, y);
}

void CsrMatrix::spmv(const void *x, void *y, cudaStream_t stream) const {
  QDSP_REQUIRE(x != nullptr && y != nullptr, QDSP_ERR_INVALID, "spmv: null vector");
  QDSP_REQUIRE(x != y, QDSP_ERR_INVALID, "spmv: x and y must not alias");
  if (rows == 0) return;
  const unsigned grid = static_cast<unsigned>((rows + SPMV_ROWS - 1) / SPMV_ROWS);
  const int *ip = indptr.get<int>(), *ix = indices.get<int>();
  if (dtype == QDSP_F32) {
    if (row_major_order)
      spmv_kernel<float, true><<<grid, SPMV_ROWS, 0, stream>>>(rows, ip, ix, values.get<float>(), static_cast<const float *>(x), static_cast<float *>(y));
    else
      spmv_kernel<float, false><<<grid, SPMV_ROWS, 0, stream>>>(rows, ip, ix, values.get<float>(), static_cast<const float *>(x), static_cast<float *>(y));
  } else {
    if (row_major_order)
      spmv_kernel<double, true><<<grid, SPMV_ROWS, 0, stream>>>(rows, ip, ix, values.get<double>(), static_cast<const double *>(x), static_cast<double *>(y));
    else
      spmv_kernel<double, false><<<grid, SPMV_ROWS, 0, stream>>>(rows, ip, ix, values.get<double>(), static_cast<const double *>(x), static_cast<double *>(y));
  }
  QDSP_LAUNCH_CHECK();
}

void CsrMatrix::to_host(std::vector<int32_t> &ip, std::vector<int32_t> &ix, std::vector<char> &vals) const {
  ip.resize(static_cast<size_t>(rows) + 1);
  ix.resize(static_cast<size_t>(nnz));
  vals.resize(static_cast<size_t>(nnz) * value_bytes());
  QDSP_CUDA(cudaMemcpy(ip.data(), indptr.get(), ip.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (nnz) {
    QDSP_CUDA(cudaMemcpy(ix.data(), indices.get(), ix.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    QDSP_CUDA(cudaMemcpy(vals.data(), values.get(), vals.size(), cudaMemcpyDeviceToHost));
  }
}

// CuSparseMatrix::get_element copies the whole CSR to the host on every call (sparse_matrix.cpp:688-712); here only
// the two indptr entries and the row's slice travel.
double CsrMatrix::get_element(int r, int c) const {
  QDSP_REQUIRE(r >= 0 && r < rows && c >= 0 && c < cols, QDSP_ERR_INVALID, "get_element: index out of range");
  int32_t be[2];
  QDSP_CUDA(cudaMemcpy(be, indptr.get<int32_t>() + r, sizeof(be), cudaMemcpyDeviceToHost));
  const int len = be[1] - be[0];
  if (len <= 0) return 0.0;
  std::vector<int32_t> ix(len);
  QDSP_CUDA(cudaMemcpy(ix.data(), indices.get<int32_t>() + be[0], sizeof(int32_t) * len, cudaMemcpyDeviceToHost));
  for (int k = 0; k < len; ++k) {
    if (ix[k] == c) {
      if (dtype == QDSP_F32) {
        float v;
        QDSP_CUDA(cudaMemcpy(&v, values.get<float>() + be[0] + k, sizeof(v), cudaMemcpyDeviceToHost));
        return v;
      }
      double v;
      QDSP_CUDA(cudaMemcpy(&v, values.get<double>() + be[0] + k, sizeof(v), cudaMemcpyDeviceToHost));
      return v;
    }
  }
  return 0.0;
}

// "%%MatrixMarket matrix coordinate real general", 1-based, row by row (sparse_matrix.cpp:748-780)
void CsrMatrix::mmwrite(const std::string &path) const {
  std::vector<int32_t> ip, ix;
  std::vector<char> vals;
  to_host(ip, ix, vals);
  std::ofstream f(path);
  QDSP_REQUIRE(f.good(), QDSP_ERR_INVALID, "mmwrite: cannot open " + path);
  f << "%%MatrixMarket matrix coordinate real general\n%" << std::endl;
  f << rows << " " << cols << " " << nnz << std::endl;
  for (int r = 0; r < rows; ++r)
    for (int k = ip[r]; k < ip[r + 1]; ++k) {
      f << r + 1 << " " << ix[k] + 1 << " ";
      if (dtype == QDSP_F32) f << reinterpret_cast<const float *>(vals.data())[k];
      else f << reinterpret_cast<const double *>(vals.data())[k];
      f << std::endl;
    }
}

}  // namespace qdsp
