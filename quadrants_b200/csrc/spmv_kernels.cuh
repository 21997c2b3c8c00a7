// spmv_kernels.cuh -- row-length-binned CSR SpMV, y = A x  (fp32 and fp64), optionally fused with the p.Ap dot of CG.
//
// Replaces CuSparseMatrix::spmv (quadrants/program/sparse_matrix.cpp:654-686: cusparseCreate + SpMV_bufferSize +
// cuMemAlloc + cusparseSpMV(CSR_ALG1, f32) + destroy + cuMemFree on EVERY call) and is held to the arithmetic of
// EigenSparseMatrix::spmv (:265-279).
//
// HBM-bound.  Algorithmic bytes per SpMV: nnz*(sv+4) + (n+1)*4 + m*sv + n*sv  (SURVEY 8(d)); the matrix stream
// (indices + values) is ~82 % of that for the 7-point Poisson matrix, so the kernel is organised around streaming it
// with perfectly coalesced 128-bit loads, independent of how the non-zeros fall into rows:
//
//   one CTA = SPMV_ROWS (256) consecutive rows, i.e. ONE contiguous slice [indptr[r0], indptr[r0+256]) of the matrix.
//   Bin by the slice's mean row length (decided per CTA from two indptr reads -- no pre-pass, no plan to store):
//     * short rows (mean <= SPMV_VECTOR_MEAN): "stream" path.  The slice is staged through shared memory in chunks:
//       every thread issues 128-bit L1-bypassing loads of 4 indices + 4 values, gathers the 4 x entries through the
//       read-only path (L1/L2 resident: the gather footprint of a Poisson slice is 3 short runs of x), and parks
//       (value, x) pairs in shared memory.  Then thread t walks row r0+t in shared memory and accumulates with fma in
//       Eigen's order, which makes the result BIT-IDENTICAL to the reference's Eigen product
//       (col_major: one accumulator left to right; row_major: two alternating accumulators).
//     * long rows: "vector" path, one warp per row, lanes stride the row with coalesced loads, warp-shuffle
//       reduction (order differs from Eigen's: held to the 1e-5 / 1e-12 per-row tolerance instead).
//   Tensor cores are deliberately unused: 2 flop per 8-12 bytes.
#pragma once
#include "common.cuh"

namespace qdsp {

constexpr int SPMV_ROWS = 256;          // rows per CTA == threads per CTA
constexpr int SPMV_CHUNK = 2048;        // staged non-zeros per round (fp32: 16 KB, fp64: 32 KB of shared memory)
constexpr int SPMV_VECTOR_MEAN = 24;    // mean row length above which a slice takes the warp-per-row path

template <typename T> struct Vec4;
template <> struct Vec4<float> {
  __device__ static __forceinline__ void load(const float *p, float (&v)[4]) {
    const int4 q = ld_stream_v4(p);
    v[0] = __int_as_float(q.x); v[1] = __int_as_float(q.y); v[2] = __int_as_float(q.z); v[3] = __int_as_float(q.w);
  }
};
template <> struct Vec4<double> {
  __device__ static __forceinline__ void load(const double *p, double (&v)[4]) {
    const int4 a = ld_stream_v4(p), b = ld_stream_v4(p + 2);
    v[0] = __hiloint2double(a.y, a.x); v[1] = __hiloint2double(a.w, a.z);
    v[2] = __hiloint2double(b.y, b.x); v[3] = __hiloint2double(b.w, b.z);
  }
};

__device__ __forceinline__ float fma_t(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fma_t(double a, double b, double c) { return fma(a, b, c); }
// unfused multiply-add (the compiler may not contract these intrinsics)
__device__ __forceinline__ float mad_unfused(float a, float b, float c) { return __fadd_rn(__fmul_rn(a, b), c); }
__device__ __forceinline__ double mad_unfused(double a, double b, double c) { return __dadd_rn(__dmul_rn(a, b), c); }

template <typename T>
struct SpmvSmem {
  int ptr[SPMV_ROWS + 1];
  T val[SPMV_CHUNK];
  T xg[SPMV_CHUNK];
  double red[SPMV_ROWS / 32];
};

// Computes y for the CTA's rows; returns (to the caller, per thread) y of row r0 + threadIdx.x or 0.
// `indices` and `values` must be readable up to 3 elements past nnz rounded up to a multiple of 4 (the matrices of
// this library carry 64 bytes of padding).
template <typename T, bool ROW_MAJOR_ORDER>
__device__ __forceinline__ T spmv_cta(SpmvSmem<T> &S, int rows, const int *__restrict__ indptr,
                                      const int *__restrict__ indices, const T *__restrict__ values,
                                      const T *__restrict__ x, T *__restrict__ y) {
  const int r0 = blockIdx.x * SPMV_ROWS;
  const int nr = min(SPMV_ROWS, rows - r0);
  const int t = threadIdx.x;
  if (t <= nr) S.ptr[t] = indptr[r0 + t];
  if (t == 0 && nr == SPMV_ROWS) S.ptr[SPMV_ROWS] = indptr[r0 + SPMV_ROWS];
  __syncthreads();
  const int s = S.ptr[0], e = S.ptr[nr];
  T result = T(0);

  if (e - s <= SPMV_VECTOR_MEAN * nr) {
    // ---- stream path ----
    const int a = t < nr ? S.ptr[t] : 0, b = t < nr ? S.ptr[t + 1] : 0;
    T acc_a = T(0), acc_b = T(0);
    for (int c0 = s & ~3; c0 < e; c0 += SPMV_CHUNK) {
      const int c1 = min(c0 + SPMV_CHUNK, e);
      for (int k = c0 + t * 4; k < c1; k += SPMV_ROWS * 4) {
        const int4 ci = ld_stream_v4(indices + k);
        T v[4];
        Vec4<T>::load(values + k, v);
        const int cc[4] = {ci.x, ci.y, ci.z, ci.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool in = (k + j >= s) && (k + j < e);
          S.val[k - c0 + j] = v[j];
          S.xg[k - c0 + j] = in ? __ldg(x + cc[j]) : T(0);
        }
      }
      __syncthreads();
      const int lo = max(a, c0), hi = min(b, c1);
      if (ROW_MAJOR_ORDER) {
        // Eigen row-major processRow: entries 0,2,4.. of the row -> accumulator a, entries 1,3,5.. -> accumulator b.
        // The reference's shipped build (GCC 13, -O3 -march=x86-64-v3) leaves a as mul+add and fuses b; mirrored so
        // that the result equals oracle/_ref bit for bit (tests/test_oracle_vs_eigen.py pins the same pattern).
        for (int k = lo; k < hi; ++k) {
          if (((k - a) & 1) == 0) acc_a = mad_unfused(S.val[k - c0], S.xg[k - c0], acc_a);
          else acc_b = fma_t(S.val[k - c0], S.xg[k - c0], acc_b);
        }
      } else {
        // Eigen col-major scatter: y_i += a_ij * x_j for j ascending, one fused multiply-add each
        for (int k = lo; k < hi; ++k) acc_a = fma_t(S.val[k - c0], S.xg[k - c0], acc_a);
      }
      __syncthreads();
    }
    result = ROW_MAJOR_ORDER ? (T(0) + (acc_a + acc_b)) : acc_a;
    if (t < nr) y[r0 + t] = result;
  } else {
    // ---- vector path: warp per row ----
    const int lane = t & 31, warp = t >> 5;
    for (int r = warp; r < nr; r += SPMV_ROWS / 32) {
      const int a = S.ptr[r], b = S.ptr[r + 1];
      T acc = T(0);
      for (int k = a + lane; k < b; k += 32) acc = fma_t(values[k], __ldg(x + indices[k]), acc);
      acc = warp_sum(acc);
      if (lane == 0) {
        y[r0 + r] = acc;
        S.val[r] = acc;  // hand the row result to thread r for the fused dot
      }
    }
    __syncthreads();
    result = t < nr ? S.val[t] : T(0);
  }
  return result;
}

}  // namespace qdsp
