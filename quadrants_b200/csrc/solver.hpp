// solver.hpp -- sparse direct solver (LLT / LDLT / LU) behind qd.linalg.SparseSolver.
// Replaces CuSparseSolver (quadrants/program/sparse_solver.h:81-136, sparse_solver.cpp:149-476: cuSOLVER low-level
// Cholesky on the device, csrlu*Host on the HOST) and is held to the results of the reference's Eigen solvers
// (sparse_solver.h:9-20: SimplicialLLT / SimplicialLDLT / SparseLU).  SURVEY 8(a) A16, 8(f) N4.
#pragma once
#include <cstdint>
#include <vector>

#include "host.hpp"

namespace qdsp {

enum SolverType { SOLVER_LLT = 0, SOLVER_LDLT = 1, SOLVER_LU = 2 };

// Host-side analysis (the reference also analyses on the host: csrsymamdHost + csrperm, sparse_solver.cpp:169-231).
// The factor pattern is that of the Cholesky factor of P (A + A^T) P^T:
//   CSC:  column j = rowidx[colptr[j] .. colptr[j+1]), ascending, the diagonal first;
//   rows: row j   = the columns k < j with L(j,k) != 0: rowcol[rowptr[j] .. rowptr[j+1]) ascending, and rowpos = the CSC
//         slot of entry (j,k) -- so a row can be walked without searching.
struct SymbolicFactor {
  int n = 0;
  std::vector<int32_t> perm, pinv;      // perm[new] = old, pinv[old] = new
  std::vector<int32_t> parent;          // elimination tree (-1 = root)
  std::vector<int32_t> colptr, rowidx;  // n + 1, |L|
  std::vector<int32_t> rowptr, rowcol, rowpos;  // n + 1, |L| - n, |L| - n
  int64_t factor_nnz = 0;               // |L| including the diagonal
  double flops = 0;                     // sum over columns of (column length)^2
  int height = 0;                       // elimination-tree height = length of the dependency chain
};

// Approximate-minimum-degree ordering of the pattern of A + A^T (quotient graph, approximate external degrees and
// aggressive element absorption after Amestoy, Davis & Duff 1996).  perm[new] = old.
void solver_order_host(int n, const int32_t *indptr, const int32_t *indices, std::vector<int32_t> &perm);
void solver_symbolic_host(int n, const int32_t *indptr, const int32_t *indices, const std::vector<int32_t> &perm,
                          SymbolicFactor &out);

// LU only: the row permutation of the static pivoting (solver.cu header).  w[q] = |value| of stored entry q;
// rowof[i] = the row of A that becomes row i.  False when no zero-free diagonal exists (rowof is still a permutation).
bool solver_lu_row_matching_host(int n, const int32_t *indptr, const int32_t *indices, const std::vector<double> &w,
                                 std::vector<int32_t> &rowof);

class DirectSolver {
 public:
  DirectSolver(int dtype, int solver_type, int ordering);
  ~DirectSolver();
  void analyze_pattern(const CsrMatrix &A, cudaStream_t stream);
  void factorize(const CsrMatrix &A, cudaStream_t stream);  // sets info()
  void solve(const void *b, void *x, cudaStream_t stream);  // device vectors of n elements
  void solve_host(const void *b_host, void *x_host, cudaStream_t stream);
  bool info() const { return factorized_ && zero_pivot_ < 0; }
  int zero_pivot() const { return zero_pivot_; }  // first failing pivot in the PERMUTED numbering, -1 = none
  int n() const { return sym_.n; }
  int dtype() const { return dtype_; }
  const SymbolicFactor &symbolic() const { return sym_; }

 private:
  int dtype_, type_, ordering_;
  bool analyzed_ = false, factorized_ = false;
  int zero_pivot_ = -1;
  int64_t nnz_a_ = 0;
  SymbolicFactor sym_;
  DeviceBuffer d_colptr_, d_rowidx_, d_rowptr_, d_rowcol_, d_rowpos_, d_perm_, d_bperm_, d_amap_;
  DeviceBuffer d_lx_, d_ux_, d_flags_, d_work_;
};

}  // namespace qdsp
