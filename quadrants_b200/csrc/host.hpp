// host.hpp -- C++ host orchestration layer above the kernels and below the C ABI (include/qdsp.h).
// One class per reference class of quadrants/program/{sparse_matrix,conjugate_gradient}.h.
#pragma once
#include <atomic>
#include <memory>
#include <string>
#include <vector>

#include "common.cuh"
#include "dist_dev.cuh"

namespace qdsp {

// ---- algorithms (scan.cu, radix_sort.cu) -----------------------------------------------------------------------------
size_t scan_tmp_bytes(int64_t n);
void inclusive_scan_i32(int32_t *inout, int64_t n, void *tmp, size_t tmp_bytes, cudaStream_t stream);
int64_t scan_static_max_n();
void inclusive_scan_i32_static(int32_t *inout, int64_t n, cudaStream_t stream);  // scratch = a static __device__ array
size_t sort_tmp_bytes(int64_t n, int key_bytes, int value_bytes);
void sort_pairs(void *keys, void *values, int64_t n, int key_kind, int value_bytes, void *tmp, size_t tmp_bytes,
                cudaStream_t stream);
void sort_build_keys(void *keys, void *values, int64_t n, int key_bytes, int value_bytes, int key_bits, void *tmp,
                     cudaStream_t stream, void **keys_out, void **values_out);
void sort_modern(void *keys, void *tmp_keys, void *values, void *tmp_values, int64_t capacity, const int32_t *n_dev,
                 int key_elem, int value_bytes, int end_bit, void *scratch, size_t scratch_bytes, cudaStream_t stream);

// ---- device-wide reduce / scan / select / reduce-by-key (device_algos.cu) ---------------------------------------------
size_t algo_scratch_bytes(int algo, int64_t capacity, int elem_bytes);
void device_reduce(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op, void *scratch,
                   size_t scratch_bytes, cudaStream_t stream);
void device_exclusive_scan(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op,
                           void *scratch, size_t scratch_bytes, cudaStream_t stream);
void device_select(const void *in, const int32_t *flags, void *out, int32_t *num_out, int64_t capacity,
                   const int32_t *n_dev, int item_bytes, void *scratch, size_t scratch_bytes, cudaStream_t stream);
void device_reduce_by_key_add(const void *keys_in, const void *values_in, void *keys_out, void *values_out,
                              int32_t *num_runs, int64_t capacity, const int32_t *n_dev, int key_elem, int value_elem,
                              void *scratch, size_t scratch_bytes, cudaStream_t stream);

// ---- CuSparseMatrix (quadrants/program/sparse_matrix.h:206-299) -------------------------------------------------------
// Owns csr_row_ptr_/csr_col_ind_/csr_val_ on the device.  Unlike the reference's shallow copy-ctor
// (sparse_matrix.h:233-234, a double-free hazard) it is move-only.
class CsrMatrix {
 public:
  int rows = 0, cols = 0;
  int64_t nnz = 0;
  int dtype = QDSP_F32;
  // which Eigen product order the short-row SpMV path mirrors bit-for-bit: false = col_major (the reference's
  // default storage_format: one accumulator, entries of a row added left to right with fma), true = row_major
  // (two alternating accumulators, SparseDenseProduct.h:66-80)
  bool row_major_order = false;
  DeviceBuffer indptr, indices, values;  // each carries 64 bytes of padding: TMA bulk copies read whole 16-byte units
  // lanes per row of the SpMV kernels (1, 2, 4 ... 32), chosen from the mean row length on first use; 0 = not yet
  mutable int spmv_lpr = 0;
  // SpMV plan, made on first use (spmv_prepare): 1 = uniform row blocks (regular matrices; bit-exact Eigen order when
  // spmv_lpr == 1), 2 = stream path over nnz-balanced plan blocks (irregular matrices), 0 = not planned yet
  mutable int spmv_mode = 0;
  mutable int spmv_max_row_len = 0;
  mutable int spmv_max_block_nnz = 0;  // largest stage footprint of a uniform row block (spmv_prepare)
  mutable DeviceBuffer spmv_plan_blocks, spmv_plan_long, spmv_plan_partials, spmv_plan_ticket;
  mutable int spmv_plan_nblk = 0, spmv_plan_nlong = 0;
  // uniform path, diagonal-coded indices (spmv_kernels.cuh SpmvSched): one byte per non-zero + the table of offsets
  mutable DeviceBuffer spmv_codes, spmv_code_tab;
  mutable int spmv_ntab = 0;        // 0 = canonical indices are streamed
  mutable int spmv_ntab_built = 0;  // entries of the table once made (kept when the codes are switched off); -1 = not codable
  mutable int spmv_compress = -1;   // -1: QDSP_SPMV_COMPRESS decides (default on), 0: off, 1: on (qdsp_csr_spmv_config)
  // stream path, hot columns (spmv_kernels.cuh): the most referenced columns, and `indices` with them stored as ~slot
  mutable DeviceBuffer spmv_plan_hot_cols, spmv_plan_hot_idx;
  mutable int spmv_plan_hot_n = 0;
  mutable double spmv_plan_hot_coverage = 0.0;  // share of the non-zeros whose column is hot
  // work counters of the dynamically scheduled tail of the uniform path (spmv_kernels.cuh SpmvSched): 16 slots used
  // round-robin (atomic: host threads launching on the same matrix draw distinct slots), so up to 16 products of one
  // matrix may be in flight at once.  A 17th concurrent product would share a counter with the 1st and both would skip
  // row blocks: callers that keep more than 16 products of ONE matrix in flight on different streams must order them.
  // The plan itself (first product of a matrix, spmv_prepare) is not thread-safe: make the first call from one thread.
  mutable DeviceBuffer spmv_sched_buf;
  mutable std::atomic<unsigned> spmv_sched_next{0};

  size_t value_bytes() const { return dtype == QDSP_F64 ? 8 : 4; }
  void allocate(int rows_, int cols_, int64_t nnz_, int dtype_);
  void spmv(const void *x, void *y, cudaStream_t stream) const;
  double get_element(int r, int c) const;
  void mmwrite(const std::string &path) const;
  void to_host(std::vector<int32_t> &ip, std::vector<int32_t> &ix, std::vector<char> &vals) const;

  static std::unique_ptr<CsrMatrix> from_device_coo(int rows, int cols, int64_t n, const int32_t *row,
                                                    const int32_t *col, const void *val, int dtype,
                                                    cudaStream_t stream);
  static std::unique_ptr<CsrMatrix> from_device_csr(int rows, int cols, int64_t nnz, const int32_t *indptr,
                                                    const int32_t *indices, const void *values, int dtype,
                                                    bool row_major_order, cudaStream_t stream);
  std::unique_ptr<CsrMatrix> add(const CsrMatrix &b, double alpha, double beta, cudaStream_t stream) const;
  std::unique_ptr<CsrMatrix> scale(double alpha, cudaStream_t stream) const;
  std::unique_ptr<CsrMatrix> transpose(cudaStream_t stream) const;
  std::unique_ptr<CsrMatrix> matmul(const CsrMatrix &b, cudaStream_t stream) const;
};

// Shared by Builder::build and CsrMatrix::from_device_coo: keys/vals already packed on the device.
std::unique_ptr<CsrMatrix> csr_from_packed(int rows, int cols, int64_t n, int dtype, int key_bytes, int col_bits,
                                           DeviceBuffer &keys, DeviceBuffer &vals, cudaStream_t stream);

// ---- SparseMatrixBuilder (quadrants/program/sparse_matrix.h:16-53) ----------------------------------------------------
class Builder {
 public:
  Builder(int rows, int cols, int64_t max_triplets, int dtype, const std::string &storage_format);
  void *buffer_ptr() const { return buf_.get(); }
  int64_t count(cudaStream_t stream) const;
  void fill_from_host(const void *triplets, int64_t count, cudaStream_t stream);
  void clear(cudaStream_t stream);
  // bulk, atomic-free replacements of per-triplet insert_triplet_f32/f64; array order is kept
  void append(const int32_t *row, const int32_t *col, const void *val, int64_t count, bool negate, cudaStream_t stream);
  void append_elements(const int32_t *nodes, const void *mats, int64_t n_elems, int k, cudaStream_t stream);
  std::unique_ptr<CsrMatrix> build(cudaStream_t stream);
  void print_triplets() const;
  int rows() const { return rows_; }
  int cols() const { return cols_; }
  int dtype() const { return dtype_; }

 private:
  int rows_, cols_;
  int64_t max_triplets_;
  int dtype_;
  bool row_major_;
  bool built_ = false;
  DeviceBuffer buf_;
  size_t slot_bytes() const { return dtype_ == QDSP_F64 ? 8 : 4; }
};

// Chooses the SpMV path of A (mode, lanes per row, plan blocks); idempotent.  Synchronises `stream` the first time.
void spmv_prepare(const CsrMatrix &A, cudaStream_t stream);
struct SpmvPlanDev;
struct SpmvSched;
// scheduling of a uniform-path launch of `grid` CTAs over `nblk` row blocks; counters = 2 zeroed unsigned words (or nullptr)
SpmvSched spmv_sched_for(int nblk, int grid, unsigned *counters);
// attaches the diagonal codes of A (if the plan made any) to a launch
void spmv_attach_codes(const CsrMatrix &A, SpmvSched &sched);
void spmv_set_compress(const CsrMatrix &A, int compress, cudaStream_t stream);
// pure host: the stream-path work partition (blocks {row0, nrows | -(segment+1), nz0, nz1}; long_rows {row, first segment})
void spmv_plan_host(const int32_t *indptr, int rows, std::vector<int4> &blocks, std::vector<int2> &long_rows,
                    int &n_segments);
SpmvPlanDev spmv_plan_dev(const CsrMatrix &A, bool allow_hot = true);
// y = A x with the persistent TMA-fed kernel; `d` (world > 1) makes every CTA wait for the halo of x first (spmv.cu)
void launch_spmv(const CsrMatrix &A, const void *x, void *y, const DistDev &d, cudaStream_t stream);

// ---- CG<EigenT,DT> and CUCG (quadrants/program/conjugate_gradient.h) ---------------------------------------------------
struct CgScalars;    // device-resident scalar block, cg.cu
class DistContext;   // dist.hpp

class CgSolver {
 public:
  CgSolver(const CsrMatrix &A, int max_iters, double tol, int mode, bool verbose, DistContext *dist = nullptr);
  ~CgSolver();
  void solve(void *x, const void *b, cudaStream_t stream, int *iterations, double *error, int *success);
  void solve_host(void *x_host, const void *b_host, cudaStream_t stream, int *iterations, double *error, int *success);
  void run_fixed(void *x, const void *b, int iters, cudaStream_t stream);
  struct Impl;

 private:
  std::unique_ptr<Impl> impl_;
};

// ---- dense vector kernels of the matrix-free Krylov solvers (krylov.cu; python/quadrants/linalg/matrixfree_cg.py) ------
void vec_dot2(const void *a, const void *b, const void *c, const void *d, int64_t n, int dtype, cudaStream_t stream,
              double *out);
void vec_lincomb(void *out, int64_t n, int dtype, int k, const void *const *xs, const double *coefs, cudaStream_t stream);
void vec_cg_step(void *x, void *r, const void *p, const void *ap, double alpha, int64_t n, int dtype, cudaStream_t stream,
                 double *rr);

}  // namespace qdsp
