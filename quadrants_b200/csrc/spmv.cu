// spmv.cu -- CsrMatrix: SpMV launcher and the small host-visible accessors of CuSparseMatrix
// (quadrants/program/sparse_matrix.cpp:654-780).
#include <cstring>
#include <fstream>
#include <algorithm>
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

template <typename T, int LPR, bool ROW_MAJOR_ORDER>
__global__ void __launch_bounds__(SPMV_THREADS)
spmv_kernel(int rows, const int *__restrict__ indptr, const int *__restrict__ indices, const T *__restrict__ values,
            const T *__restrict__ x, T *__restrict__ y, DistDev d) {
  extern __shared__ unsigned char smem_raw[];
  spmv_persistent<T, LPR, ROW_MAJOR_ORDER>(smem_raw, rows, indptr, indices, values, x, y, nullptr, d);
}

namespace {

// persistent grid = SM count x resident CTAs of this instantiation (queried once per instantiation)
template <typename T, int LPR>
int spmv_ctas_per_sm() {
  static int occ = 0;
  if (occ == 0) {
    const size_t smem = spmv_smem_bytes<T>();
    QDSP_CUDA(cudaFuncSetAttribute(spmv_kernel<T, LPR, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    QDSP_CUDA(cudaFuncSetAttribute(spmv_kernel<T, LPR, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int o = 0;
    QDSP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, spmv_kernel<T, LPR, false>, SPMV_THREADS, smem));
    occ = std::max(1, o);
  }
  return occ;
}

template <typename T, int LPR>
void launch_spmv_t(const CsrMatrix &A, const T *x, T *y, const DistDev &d, cudaStream_t stream) {
  constexpr int R = SPMV_CONSUMERS / LPR;
  const int grid = std::max(1, std::min(ceil_div(A.rows, R), device_props().sm_count * spmv_ctas_per_sm<T, LPR>()));
  const size_t smem = spmv_smem_bytes<T>();
  const int *ip = A.indptr.get<int>(), *ix = A.indices.get<int>();
  if (A.row_major_order && LPR == 1)
    spmv_kernel<T, LPR, true><<<grid, SPMV_THREADS, smem, stream>>>(A.rows, ip, ix, A.values.get<T>(), x, y, d);
  else
    spmv_kernel<T, LPR, false><<<grid, SPMV_THREADS, smem, stream>>>(A.rows, ip, ix, A.values.get<T>(), x, y, d);
  QDSP_LAUNCH_CHECK();
}

template <typename T>
void launch_spmv_lpr(const CsrMatrix &A, const void *x, void *y, const DistDev &d, cudaStream_t stream) {
  const T *xv = static_cast<const T *>(x);
  T *yv = static_cast<T *>(y);
  switch (A.spmv_lpr) {
    case 1: launch_spmv_t<T, 1>(A, xv, yv, d, stream); break;
    case 2: launch_spmv_t<T, 2>(A, xv, yv, d, stream); break;
    case 4: launch_spmv_t<T, 4>(A, xv, yv, d, stream); break;
    case 8: launch_spmv_t<T, 8>(A, xv, yv, d, stream); break;
    case 16: launch_spmv_t<T, 16>(A, xv, yv, d, stream); break;
    default: launch_spmv_t<T, 32>(A, xv, yv, d, stream); break;
  }
}

}  // namespace

void launch_spmv(const CsrMatrix &A, const void *x, void *y, const DistDev &d, cudaStream_t stream) {
  if (A.rows == 0) return;
  if (A.spmv_lpr == 0) A.spmv_lpr = spmv_lanes_per_row(A.nnz, A.rows);
  if (A.dtype == QDSP_F32) launch_spmv_lpr<float>(A, x, y, d, stream);
  else launch_spmv_lpr<double>(A, x, y, d, stream);
}

void CsrMatrix::spmv(const void *x, void *y, cudaStream_t stream) const {
  QDSP_REQUIRE(x != nullptr && y != nullptr, QDSP_ERR_INVALID, "spmv: null vector");
  QDSP_REQUIRE(x != y, QDSP_ERR_INVALID, "spmv: x and y must not alias");
  launch_spmv(*this, x, y, DistDev{}, stream);
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
