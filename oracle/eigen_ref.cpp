// eigen_ref.cpp -- the reference's own CPU arithmetic, compiled from the
// reference's vendored Eigen (/root/reference/external/eigen, 3.4.90).
//
// TEST INFRASTRUCTURE ONLY (see oracle/qd_oracle.c header).  This TU is ours;
// it contains no reference source.  It performs, call for call, what the
// reference's C++ "program" layer does on arch=x64, so that the C restatement
// in qd_oracle.c and the CUDA path can both be pinned against the reference's
// real numerics:
//   build : SparseMatrixBuilder::build_template<T,G>  (quadrants/program/sparse_matrix.cpp:139-152)
//           -> EigenSparseMatrix::build_triplets -> matrix_.setFromTriplets (:12-17,254-263)
//   spmv  : EigenSparseMatrix::spmv  (:265-279)   y = matrix_.cast<T>() * x
//   cg    : CG<EigenT,DT>::solve     (quadrants/program/conjugate_gradient.h:36-49)
//           Eigen::ConjugateGradient<SparseMatrix<DT>, Lower|Upper>, setMaxIterations, setTolerance,
//           compute, solve, iterations(), error(), info()
//   direct: EigenSparseSolver<...>::analyze_pattern / factorize / solve / info  (quadrants/program/sparse_solver.cpp:74-123)
//           with the twelve solver types of sparse_solver.h:9-20,60-71: Simplicial{LLT,LDLT}<SparseMatrix<dt>, Lower,
//           {AMD,COLAMD}Ordering<int>> and SparseLU<SparseMatrix<dt>, {AMD,COLAMD}Ordering<int>>
// The full Quadrants package cannot be built here (LLVM 22, nanobind, spdlog,
// dlpack, SPIRV-Cross are absent), which is why the calls are restated rather
// than linked.  Built by oracle/Makefile into oracle/_ref/ (git-ignored) with
// the reference's flags: -O3 -DNDEBUG -std=c++17 -march=x86-64-v3
// (CMakeLists.txt:20,78-83; cmake/QuadrantsCXXFlags.cmake:123-124); the _omp
// variant adds -fopenmp (Eigen then threads the row-major product only,
// SparseDenseProduct.h:47-58) and is a best-effort baseline, not "as shipped".
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "Eigen/Sparse"
#include "Eigen/IterativeLinearSolvers"
#include "Eigen/SparseCholesky"
#include "Eigen/SparseLU"
#include "Eigen/OrderingMethods"

#define REF_API extern "C" __attribute__((visibility("default")))

namespace {

struct RefMatrix {
  int rows = 0, cols = 0;
  bool f64 = false, row_major = false;
  Eigen::SparseMatrix<float, Eigen::ColMajor> fc;
  Eigen::SparseMatrix<float, Eigen::RowMajor> fr;
  Eigen::SparseMatrix<double, Eigen::ColMajor> dc;
  Eigen::SparseMatrix<double, Eigen::RowMajor> dr;
};

template <typename T, typename G, typename M>
void build_into(M &m, const void *buf) {
  using V = Eigen::Triplet<T>;
  std::vector<V> triplets;
  const G *data = reinterpret_cast<const G *>(buf);
  const uint64_t n = static_cast<uint64_t>(data[0]);
  data += 1;
  for (uint64_t i = 0; i < n; i++) {
    T v;
    std::memcpy(&v, &data[i * 3 + 2], sizeof(T));
    triplets.push_back(V(static_cast<int>(data[i * 3]), static_cast<int>(data[i * 3 + 1]), v));
  }
  m.setFromTriplets(triplets.begin(), triplets.end());
}

template <typename M, typename T>
void export_arrays(const M &m, int32_t *outer, int32_t *inner, T *vals) {
  std::memcpy(outer, m.outerIndexPtr(), sizeof(int32_t) * (m.outerSize() + 1));
  std::memcpy(inner, m.innerIndexPtr(), sizeof(int32_t) * m.nonZeros());
  std::memcpy(vals, m.valuePtr(), sizeof(T) * m.nonZeros());
}

}  // namespace

REF_API void *ref_matrix_build(const void *buf, int rows, int cols, int is_f64, int row_major) {
  auto *m = new RefMatrix();
  m->rows = rows;
  m->cols = cols;
  m->f64 = is_f64 != 0;
  m->row_major = row_major != 0;
  if (!m->f64 && !m->row_major) { m->fc.resize(rows, cols); build_into<float, int32_t>(m->fc, buf); }
  if (!m->f64 && m->row_major)  { m->fr.resize(rows, cols); build_into<float, int32_t>(m->fr, buf); }
  if (m->f64 && !m->row_major)  { m->dc.resize(rows, cols); build_into<double, int64_t>(m->dc, buf); }
  if (m->f64 && m->row_major)   { m->dr.resize(rows, cols); build_into<double, int64_t>(m->dr, buf); }
  return m;
}

// Adopts compressed arrays (outer has outer_size+1 entries) instead of going through setFromTriplets: used by
// bench.py to hand the 117 M non-zero C2 matrix to Eigen without a second 1.4 GB triplet copy.  The arithmetic timed
// afterwards (product, ConjugateGradient) is the same Eigen code on the same compressed storage.
REF_API void *ref_matrix_from_compressed(int rows, int cols, int64_t nnz, const int32_t *outer, const int32_t *inner,
                                         const void *vals, int is_f64, int row_major) {
  auto *m = new RefMatrix();
  m->rows = rows;
  m->cols = cols;
  m->f64 = is_f64 != 0;
  m->row_major = row_major != 0;
  if (!m->f64) {
    const float *v = static_cast<const float *>(vals);
    if (m->row_major) m->fr = Eigen::Map<const Eigen::SparseMatrix<float, Eigen::RowMajor>>(rows, cols, nnz, outer, inner, v);
    else m->fc = Eigen::Map<const Eigen::SparseMatrix<float, Eigen::ColMajor>>(rows, cols, nnz, outer, inner, v);
  } else {
    const double *v = static_cast<const double *>(vals);
    if (m->row_major) m->dr = Eigen::Map<const Eigen::SparseMatrix<double, Eigen::RowMajor>>(rows, cols, nnz, outer, inner, v);
    else m->dc = Eigen::Map<const Eigen::SparseMatrix<double, Eigen::ColMajor>>(rows, cols, nnz, outer, inner, v);
  }
  return m;
}

REF_API void ref_matrix_free(void *h) { delete static_cast<RefMatrix *>(h); }

REF_API int64_t ref_matrix_nnz(void *h) {
  auto *m = static_cast<RefMatrix *>(h);
  if (!m->f64) return m->row_major ? m->fr.nonZeros() : m->fc.nonZeros();
  return m->row_major ? m->dr.nonZeros() : m->dc.nonZeros();
}

REF_API int ref_matrix_is_compressed(void *h) {
  auto *m = static_cast<RefMatrix *>(h);
  if (!m->f64) return m->row_major ? m->fr.isCompressed() : m->fc.isCompressed();
  return m->row_major ? m->dr.isCompressed() : m->dc.isCompressed();
}

REF_API void ref_matrix_export(void *h, int32_t *outer, int32_t *inner, void *vals) {
  auto *m = static_cast<RefMatrix *>(h);
  if (!m->f64) {
    if (m->row_major) export_arrays(m->fr, outer, inner, static_cast<float *>(vals));
    else export_arrays(m->fc, outer, inner, static_cast<float *>(vals));
  } else {
    if (m->row_major) export_arrays(m->dr, outer, inner, static_cast<double *>(vals));
    else export_arrays(m->dc, outer, inner, static_cast<double *>(vals));
  }
}

// EigenSparseMatrix::spmv: y is the (zero-filled) result ndarray, assigned, not accumulated.
REF_API void ref_matrix_spmv(void *h, const void *x, void *y) {
  auto *m = static_cast<RefMatrix *>(h);
  if (!m->f64) {
    Eigen::Map<const Eigen::VectorXf> xv(static_cast<const float *>(x), m->cols);
    Eigen::Map<Eigen::VectorXf> yv(static_cast<float *>(y), m->rows);
    if (m->row_major) yv = m->fr.cast<float>() * xv;
    else yv = m->fc.cast<float>() * xv;
  } else {
    Eigen::Map<const Eigen::VectorXd> xv(static_cast<const double *>(x), m->cols);
    Eigen::Map<Eigen::VectorXd> yv(static_cast<double *>(y), m->rows);
    if (m->row_major) yv = m->dr.cast<double>() * xv;
    else yv = m->dc.cast<double>() * xv;
  }
}

// CG<EigenT,DT>::solve.  The reference only accepts col-major matrices here
// (static_cast at conjugate_gradient.h:40); `allow_row_major` != 0 lets the
// best-effort OpenMP baseline run the same solver on a row-major matrix so
// that Eigen threads the product.  Returns Eigen's info() (0 = Success).
REF_API int ref_cg_solve(void *h, const void *b, void *x, int max_iters, double tol, int allow_row_major,
                         int *iters, double *err) {
  auto *m = static_cast<RefMatrix *>(h);
  if (m->row_major && !allow_row_major) return -1;
  int info = -1;
  if (!m->f64) {
    Eigen::Map<const Eigen::VectorXf> bv(static_cast<const float *>(b), m->rows);
    Eigen::VectorXf b_ = bv, x_;
    const float tol_ = static_cast<float>(tol);
    if (!m->row_major) {
      Eigen::ConjugateGradient<Eigen::SparseMatrix<float>, Eigen::Lower | Eigen::Upper> cg;
      cg.setMaxIterations(max_iters); cg.setTolerance(tol_); cg.compute(m->fc);
      x_ = cg.solve(b_); *iters = static_cast<int>(cg.iterations()); *err = cg.error(); info = cg.info();
    } else {
      Eigen::ConjugateGradient<Eigen::SparseMatrix<float, Eigen::RowMajor>, Eigen::Lower | Eigen::Upper> cg;
      cg.setMaxIterations(max_iters); cg.setTolerance(tol_); cg.compute(m->fr);
      x_ = cg.solve(b_); *iters = static_cast<int>(cg.iterations()); *err = cg.error(); info = cg.info();
    }
    std::memcpy(x, x_.data(), sizeof(float) * m->cols);
  } else {
    Eigen::Map<const Eigen::VectorXd> bv(static_cast<const double *>(b), m->rows);
    Eigen::VectorXd b_ = bv, x_;
    // tol travels through a float ctor argument (conjugate_gradient.h:12,64; export_lang.cpp:1287-1292)
    const double tol_ = static_cast<double>(static_cast<float>(tol));
    if (!m->row_major) {
      Eigen::ConjugateGradient<Eigen::SparseMatrix<double>, Eigen::Lower | Eigen::Upper> cg;
      cg.setMaxIterations(max_iters); cg.setTolerance(tol_); cg.compute(m->dc);
      x_ = cg.solve(b_); *iters = static_cast<int>(cg.iterations()); *err = cg.error(); info = cg.info();
    } else {
      Eigen::ConjugateGradient<Eigen::SparseMatrix<double, Eigen::RowMajor>, Eigen::Lower | Eigen::Upper> cg;
      cg.setMaxIterations(max_iters); cg.setTolerance(tol_); cg.compute(m->dr);
      x_ = cg.solve(b_); *iters = static_cast<int>(cg.iterations()); *err = cg.error(); info = cg.info();
    }
    std::memcpy(x, x_.data(), sizeof(double) * m->cols);
  }
  return info;
}

namespace {
// analyzePattern + factorize + solve + info, exactly the member calls of EigenSparseSolver (sparse_solver.cpp:86-123)
template <typename Solver, typename M, typename V>
int direct_solve(const M &mat, const V &b, V &x, double *factor_ms) {
  Solver solver;
  solver.analyzePattern(mat);
  solver.factorize(mat);
  if (solver.info() != Eigen::Success) return static_cast<int>(solver.info());
  x = solver.solve(b);
  (void)factor_ms;
  return static_cast<int>(solver.info());
}
template <typename T>
int direct_dispatch(const Eigen::SparseMatrix<T> &mat, int type, int ordering, const T *b, T *x, int n) {
  using M = Eigen::SparseMatrix<T>;
  using V = Eigen::Matrix<T, Eigen::Dynamic, 1>;
  V bv = Eigen::Map<const V>(b, n), xv;
  int info = -1;
  const int key = type * 2 + ordering;
  switch (key) {
    case 0: info = direct_solve<Eigen::SimplicialLLT<M, Eigen::Lower, Eigen::AMDOrdering<int>>>(mat, bv, xv, nullptr); break;
    case 1: info = direct_solve<Eigen::SimplicialLLT<M, Eigen::Lower, Eigen::COLAMDOrdering<int>>>(mat, bv, xv, nullptr); break;
    case 2: info = direct_solve<Eigen::SimplicialLDLT<M, Eigen::Lower, Eigen::AMDOrdering<int>>>(mat, bv, xv, nullptr); break;
    case 3: info = direct_solve<Eigen::SimplicialLDLT<M, Eigen::Lower, Eigen::COLAMDOrdering<int>>>(mat, bv, xv, nullptr); break;
    case 4: info = direct_solve<Eigen::SparseLU<M, Eigen::AMDOrdering<int>>>(mat, bv, xv, nullptr); break;
    case 5: info = direct_solve<Eigen::SparseLU<M, Eigen::COLAMDOrdering<int>>>(mat, bv, xv, nullptr); break;
    default: return -2;
  }
  if (info == 0) std::memcpy(x, xv.data(), sizeof(T) * n);
  return info;
}
}  // namespace

// type: 0 LLT, 1 LDLT, 2 LU; ordering: 0 AMD, 1 COLAMD.  The matrix must be col-major (the solvers are declared on
// Eigen::SparseMatrix<dt>, sparse_solver.h:9-20).  Returns Eigen's info() (0 = Success), -1 for a row-major handle.
REF_API int ref_direct_solve(void *h, int type, int ordering, const void *b, void *x) {
  auto *m = static_cast<RefMatrix *>(h);
  if (m->row_major) return -1;
  if (!m->f64) return direct_dispatch<float>(m->fc, type, ordering, static_cast<const float *>(b), static_cast<float *>(x), m->rows);
  return direct_dispatch<double>(m->dc, type, ordering, static_cast<const double *>(b), static_cast<double *>(x), m->rows);
}

REF_API int ref_num_threads(void) { return Eigen::nbThreads(); }
REF_API void ref_set_num_threads(int n) { Eigen::setNbThreads(n); }
REF_API const char *ref_eigen_version(void) {
  static char v[32];
  std::snprintf(v, sizeof v, "%d.%d.%d", EIGEN_WORLD_VERSION, EIGEN_MAJOR_VERSION, EIGEN_MINOR_VERSION);
  return v;
}
