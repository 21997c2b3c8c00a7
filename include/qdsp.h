/*
 * qdsp.h -- C ABI of the B200-native sparse linear-algebra path that sits behind Quadrants'
 *           qd.linalg.{SparseMatrixBuilder,SparseMatrix,SparseCG} and qd.algorithms.{parallel_sort,PrefixSumExecutor}.
 *
 * This is the drop-in boundary.  The reference has no C API; its boundary is the nanobind surface of
 * quadrants/python/export_lang.cpp:518-531,1166-1295 over the C++ classes of quadrants/program/sparse_matrix.{h,cpp}
 * and conjugate_gradient.{h,cpp}.  Every entry point below names the reference member it replaces; INTEGRATION.md
 * shows the ~100-line patch that makes those members call this ABI.
 *
 * Conventions: plain C, opaque handles, raw device pointers (void*), `void *stream` is a cudaStream_t (NULL = legacy
 * default stream, which is what the reference's library handles use).  Every function returns 0 (QDSP_OK) or a
 * qdsp_status; the message of the last failure on the calling thread is qdsp_last_error().  No exception crosses the
 * ABI.  There is no CPU fallback: without a CUDA device every compute entry point returns QDSP_ERR_CUDA.
 * All sizes are in elements unless the name says bytes.  Indices are int32 end to end like the reference
 * (cuSPARSE INDEX_32I, quadrants/program/sparse_matrix.cpp:375-377).
 */
#ifndef QDSP_H_
#define QDSP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden; only this ABI is exported */
#endif

#define QDSP_ABI_VERSION 1

typedef enum {
  QDSP_OK = 0,
  QDSP_ERR_INVALID = 1, /* bad argument / out-of-range triplet index (reference: QD_ASSERT / eigen_assert)           */
  QDSP_ERR_CUDA = 2,    /* CUDA runtime failure (reference: QD_ERROR_IF in CUDADriverFunction, cuda_driver.h:92-95)  */
  QDSP_ERR_STATE = 3,   /* call order misuse, e.g. build() twice (reference: QD_ASSERT(built_ == false))             */
  QDSP_ERR_UNSUPPORTED = 4, /* reference: QD_NOT_IMPLEMENTED                                                         */
  QDSP_ERR_COMM = 5     /* multi-GPU peer mapping / NCCL failure                                                     */
} qdsp_status;

typedef enum { QDSP_F32 = 0, QDSP_F64 = 1 } qdsp_dtype;

/* key kinds of qdsp_sort_pairs: order is numeric for every kind (sign / IEEE twiddles applied on the fly) */
typedef enum { QDSP_KEY_U32 = 0, QDSP_KEY_I32 = 1, QDSP_KEY_F32 = 2, QDSP_KEY_U64 = 3, QDSP_KEY_I64 = 4,
               QDSP_KEY_F64 = 5 } qdsp_key_kind;

/* CG semantics.  EIGEN_JACOBI = the reference's arch=x64 path, Eigen::ConjugateGradient with the default diagonal
 * preconditioner, zero start, relative tolerance (quadrants/program/conjugate_gradient.h:36-49).  CUCG = the
 * reference's arch=cuda path, plain CG on the x passed in, absolute r.r <= tol^2
 * (quadrants/program/conjugate_gradient.cpp:19-83). */
typedef enum { QDSP_CG_EIGEN_JACOBI = 0, QDSP_CG_CUCG = 1 } qdsp_cg_mode;

typedef struct qdsp_builder_s *qdsp_builder_t;
typedef struct qdsp_csr_s *qdsp_csr_t;
typedef struct qdsp_cg_s *qdsp_cg_t;
typedef struct qdsp_dist_s *qdsp_dist_t;

/* ---- library ------------------------------------------------------------------------------------------------- */
const char *qdsp_last_error(void);
int qdsp_abi_version(void);
/* Binds the calling thread to `device` (the reference hard-codes device 0, quadrants/rhi/cuda/cuda_context.cpp:16-26)
 * and reports what the launch heuristics use. */
int qdsp_init(int device);
int qdsp_device_info(int *device, int *sm_count, int *cc_major, int *cc_minor, size_t *l2_bytes, size_t *hbm_bytes);
/* Number of kernels of this library launched by the calling process so far (bench.py's gpu_launches). */
int64_t qdsp_launch_count(void);
/* The multi-GB temporaries of a build / sort are cached between calls instead of cudaMalloc'd and freed every time (cap:
 * QDSP_SCRATCH_POOL_MB, default 16384; 0 = no cache -- and never more than a quarter of the device memory that would be
 * free without the cache, re-checked whenever a block comes back, so the cache shrinks as other allocators fill the
 * device).  This returns the cached blocks to the driver.  (The reference cuMemAllocs and frees its work buffers on every
 * call, e.g. sparse_matrix.cpp:654-686.) */
int qdsp_release_scratch(void);
/* Bytes cached right now and the cap that applies right now on the current device (diagnostics; no reference counterpart). */
int qdsp_scratch_info(size_t *cached_bytes, size_t *cap_bytes);

/* ---- SparseMatrixBuilder (quadrants/program/sparse_matrix.h:16-53) -------------------------------------------- */
/* ctor + create_ndarray (sparse_matrix.cpp:70-85): allocates the device triplet buffer of 3*max_triplets+1 slots of
 * 4 (f32) or 8 (f64) bytes, layout [count | (row, col, value-bits)...] written by insert_triplet_f32/f64
 * (quadrants/runtime/llvm/runtime_module/internal_functions.h:12-39).  Slot 0 is zeroed (the reference relies on the
 * pool for that).  storage_format: "col_major" | "row_major"; like make_cu_sparse_matrix the device matrix is CSR
 * either way, the format only selects which Eigen summation order SpMV mirrors bit-for-bit. */
int qdsp_builder_create(int rows, int cols, int64_t max_triplets, int dtype, const char *storage_format,
                        qdsp_builder_t *out);
/* get_ndarray_data_ptr (sparse_matrix.cpp:135-137): the address JIT fill kernels receive as their u64 argument. */
int qdsp_builder_buffer_ptr(qdsp_builder_t b, void **device_ptr);
/* num_triplets_ as build_cuda reads it (sparse_matrix.cpp:178): D2H copy of slot 0 (synchronises `stream`). */
int qdsp_builder_count(qdsp_builder_t b, void *stream, int64_t *count);
/* Host-side fill for callers without a JIT kernel: copies `count` triplets already in the slot layout (without the
 * leading count slot) from host memory and sets slot 0.  (No reference counterpart; used by the e2e path.) */
int qdsp_builder_fill_from_host(qdsp_builder_t b, const void *host_triplets, int64_t count, void *stream);
/* clear (sparse_matrix.cpp:224-228) */
int qdsp_builder_clear(qdsp_builder_t b, void *stream);
/* build_cuda (sparse_matrix.cpp:173-222) + CuSparseMatrix::build_csr_from_coo (:344-397), entirely on the device:
 * pack -> onesweep radix sort by (row, col) -> head flags + decoupled look-back scan -> duplicate sum in buffer order
 * -> indptr.  Result equals Eigen's row-major setFromTriplets bit-for-bit (indptr, indices, values). */
int qdsp_builder_build(qdsp_builder_t b, void *stream, qdsp_csr_t *out);
/* print_triplets_cuda (sparse_matrix.cpp:119-133) */
int qdsp_builder_print_triplets(qdsp_builder_t b);
/* delete_ndarray + dtor */
int qdsp_builder_destroy(qdsp_builder_t b);

/* ---- CuSparseMatrix (quadrants/program/sparse_matrix.h:206-299) ------------------------------------------------ */
/* build_csr_from_coo (sparse_matrix.cpp:344-397): device COO (unsorted, duplicates allowed and summed in array
 * order) -> CSR.  The inputs are only read (the reference takes ownership of them). */
int qdsp_csr_from_device_coo(int rows, int cols, int64_t nnz, const int32_t *row, const int32_t *col, const void *val,
                             int dtype, void *stream, qdsp_csr_t *out);
/* Adopts device CSR arrays by copy (make_cu_sparse_matrix(mat, rows, cols, ...), sparse_matrix.cpp:304-316). */
int qdsp_csr_from_device_csr(int rows, int cols, int64_t nnz, const int32_t *indptr, const int32_t *indices,
                             const void *values, int dtype, int order_row_major, void *stream, qdsp_csr_t *out);
int qdsp_csr_info(qdsp_csr_t m, int *rows, int *cols, int64_t *nnz, int *dtype);
/* Borrowed device pointers to csr_row_ptr_/csr_col_ind_/csr_val_ (valid until qdsp_csr_destroy). */
int qdsp_csr_export(qdsp_csr_t m, const int32_t **indptr, const int32_t **indices, const void **values);
/* spmv / nd_spmv (sparse_matrix.cpp:654-686): y = A x, x has cols elements, y rows elements, both dense device.
 * Concurrency: the first product of a matrix plans it and must not race with another call on the same matrix; after
 * that, host threads may launch products of one matrix concurrently, with at most 16 of them in flight at a time
 * (on different streams) -- beyond that the caller has to order them. */
int qdsp_csr_spmv(qdsp_csr_t m, const void *x, void *y, void *stream);
/* SpMV storage option of one matrix.  compress: 1 = stream one-byte diagonal codes instead of the 4-byte column indices
 * when the matrix has at most 256 distinct (column - row) values (stencils, banded matrices; the plan keeps an extra
 * nnz-byte array, the canonical arrays stay exportable and the results are bit-identical), 0 = always stream the
 * canonical indices, -1 = the process default (QDSP_SPMV_COMPRESS, on unless "0").  Applies to products and solvers
 * started afterwards; a solver created earlier keeps the path it was created with (the code arrays are never moved or
 * freed while the matrix lives).  No reference counterpart (cuSPARSE has no such format for CSR_ALG1). */
int qdsp_csr_spmv_config(qdsp_csr_t m, int compress);
/* What the plan of `m` streams per product (plans the matrix if needed): *index_bytes_per_nnz = 1 (diagonal codes) or 4,
 * *path = 1 uniform row blocks / 2 irregular (stream) path, *hot_columns = columns cached in shared memory (stream path). */
int qdsp_csr_spmv_info(qdsp_csr_t m, void *stream, int *index_bytes_per_nnz, int *path, int *hot_columns);
/* Pure host helper (no CUDA needed): how the SpMV of an irregular matrix (row lengths far above the mean) cuts the rows
 * into work blocks of at most 2044 non-zeros and 256 rows; a longer row becomes several single-row segments.
 * blocks_out: 4 int32 per block {row0, nrows (> 0) or -(segment id + 1), nz_begin, nz_end}.  No reference counterpart
 * (cuSPARSE CSR_ALG1 hides its own partitioning); exposed so that the plan is testable without a GPU. */
int qdsp_spmv_plan_host(const int32_t *indptr, int rows, int32_t *blocks_out, int64_t block_capacity, int64_t *n_blocks,
                        int *n_long_rows, int *n_segments);
/* Same through host buffers (H2D x, SpMV, D2H y): the reference-facing call bench.py's e2e leg times. */
int qdsp_csr_spmv_host(qdsp_csr_t m, const void *x_host, void *y_host, void *stream);
/* get_element (sparse_matrix.cpp:688-712): 0 when absent */
int qdsp_csr_get_element(qdsp_csr_t m, int row, int col, double *out);
/* addition / subtraction / scalar multiply / transpose / matmul (sparse_matrix.cpp:413-652) */
int qdsp_csr_add(qdsp_csr_t a, qdsp_csr_t b, double alpha, double beta, void *stream, qdsp_csr_t *out);
int qdsp_csr_scale(qdsp_csr_t a, double alpha, void *stream, qdsp_csr_t *out);
int qdsp_csr_transpose(qdsp_csr_t a, void *stream, qdsp_csr_t *out);
int qdsp_csr_matmul(qdsp_csr_t a, qdsp_csr_t b, void *stream, qdsp_csr_t *out);
/* mmwrite (sparse_matrix.cpp:748-780): Matrix Market coordinate real general, 1-based */
int qdsp_csr_mmwrite(qdsp_csr_t m, const char *path);
int qdsp_csr_destroy(qdsp_csr_t m);

/* ---- CG<EigenT,DT> / CUCG (quadrants/program/conjugate_gradient.h) -------------------------------------------- */
/* make_{float,double}_cg_solver / make_cucg_solver (conjugate_gradient.h:67-70, conjugate_gradient.cpp:85-87).
 * The solver keeps a reference to `m` like the reference does; destroy the solver first. */
int qdsp_cg_create(qdsp_csr_t m, int max_iters, double tol, int mode, int verbose, qdsp_cg_t *out);
/* solve (conjugate_gradient.h:36-49 / conjugate_gradient.cpp:19-83).  x, b: dense device vectors.  EIGEN_JACOBI
 * overwrites x from a zero start; CUCG starts from the x passed in.  iterations/error/success follow Eigen's
 * iterations()/error()/info()==Success; in CUCG mode error is r.r and success is r.r <= tol^2.
 * x must be 16-byte aligned (the update kernels move it as 128-bit words); QDSP_ERR_INVALID otherwise. */
int qdsp_cg_solve(qdsp_cg_t s, void *x, const void *b, void *stream, int *iterations, double *error, int *success);
int qdsp_cg_solve_host(qdsp_cg_t s, void *x_host, const void *b_host, void *stream, int *iterations, double *error,
                       int *success);
/* Runs exactly `iters` iterations with the convergence exit disabled (throughput timing; not a reference call). */
int qdsp_cg_run_fixed(qdsp_cg_t s, void *x, const void *b, int iters, void *stream);
int qdsp_cg_destroy(qdsp_cg_t s);

/* ---- multi-GPU: row-block partition, one process per GPU (no reference counterpart: the reference opens device 0 only,
 * quadrants/rhi/cuda/cuda_context.cpp:16-26; design in SURVEY.md 8(e)) -------------------------------------------------
 * Rank g owns global rows and columns [row_offsets[g], row_offsets[g+1]).  Each rank builds the CSR of ITS rows with
 * global column ids (rows = n_local, cols = global size), then:
 *   qdsp_dist_localize      renumbers the columns (owned first, then from ghost_base -- n_local rounded up to a 128-byte
 *                           boundary -- the ghosts ascending by global id) and allocates the
 *                           peer-visible buffer that holds the SpMV operand [owned | ghosts];
 *   qdsp_dist_ghost_ids     the ghost ids, grouped by owner (owner_offsets): the caller ships each group to its owner
 *                           (any transport; quadrants_b200.distributed uses torch.distributed), which registers it with
 *   qdsp_dist_set_send_list (dst_offset = the requester's ghost_base + its owner_offsets[owner]; dst_offset_z likewise
 *                           from the requester's zbase);
 *   qdsp_dist_comm_handle / qdsp_dist_attach   exchange and map the CUDA IPC handles of the buffers (NVLink P2P).
 * After that qdsp_dist_spmv and a solver from qdsp_cg_create_dist run with the halo of x pushed by peer stores and the
 * CG dot products all-reduced inside the kernels over peer memory (QDSP_ALLREDUCE=nccl: by ncclAllReduce instead, after
 * qdsp_dist_nccl_init).  x, b, y are the rank's owned slices (n_local elements).  Every rank must make the same calls. */
/* Pure host helper behind qdsp_dist_localize (no CUDA needed): plans one rank from its global column ids. */
int qdsp_dist_plan_host(int rank, int world, const int64_t *row_offsets, int32_t *indices_inout, int64_t nnz,
                        int ghost_base /* local id of the first ghost; < 0: n_local */, int32_t *ghost_ids_out,
                        int64_t ghost_capacity, int64_t *n_ghost, int32_t *owner_offsets_out);
int qdsp_dist_create(int rank, int world, qdsp_dist_t *out);
int qdsp_dist_destroy(qdsp_dist_t d);
int qdsp_dist_localize(qdsp_dist_t d, qdsp_csr_t m, const int64_t *row_offsets, void *stream, int *n_local,
                       int *ghost_base, int *n_ghost, int32_t *owner_offsets_out /* world + 1 */);
int qdsp_dist_ghost_ids(qdsp_dist_t d, const int32_t **host_ids);
/* dst_offset_z: where the same rows go in the requester's SECOND ghost segment (it starts at the requester's
 * zbase = ghost_base + n_ghost rounded up to a multiple of 32 elements); CG sends z = D^-1 r through it one kernel ahead
 * of need and every rank updates its own ghost copies of p.  Pass -1 to keep the plain p halo. */
int qdsp_dist_set_send_list(qdsp_dist_t d, int peer, const int32_t *global_ids, int count, int dst_offset,
                            int dst_offset_z);
int qdsp_dist_comm_handle(qdsp_dist_t d, void *out64);
int qdsp_dist_attach(qdsp_dist_t d, const void *handles /* world x 64 bytes in rank order */);
int qdsp_dist_nccl_unique_id(void *out128);
int qdsp_dist_nccl_init(qdsp_dist_t d, const void *id128);
int qdsp_dist_spmv(qdsp_dist_t d, qdsp_csr_t m, const void *x_local, void *y_local, void *stream);
/* Like qdsp_cg_create on the localized matrix of `d` (QDSP_CG_EIGEN_JACOBI only).  iterations / error / success are
 * identical on every rank: all ranks sum the all-reduced scalars in the same order. */
int qdsp_cg_create_dist(qdsp_csr_t m, qdsp_dist_t d, int max_iters, double tol, int mode, int verbose, qdsp_cg_t *out);

/* ---- qd.algorithms.parallel_sort / PrefixSumExecutor ----------------------------------------------------------- */
/* parallel_sort (python/quadrants/algorithms/_algorithms.py:19-56): in-place ascending sort of keys, values (or
 * NULL) permuted alike.  Stable (the reference network is not; its tests never observe tie order).  value_bytes is
 * 0, 4 or 8.  tmp: caller-owned scratch of qdsp_sort_tmp_bytes(). */
size_t qdsp_sort_tmp_bytes(int64_t n, int key_kind, int value_bytes);
int qdsp_sort_pairs(void *keys, void *values, int64_t n, int key_kind, int value_bytes, void *tmp, size_t tmp_bytes,
                    void *stream);
/* PrefixSumExecutor.run (_algorithms.py:109-163): in-place inclusive i32 prefix sum, wrapping. */
size_t qdsp_scan_tmp_bytes(int64_t n);
int qdsp_inclusive_scan_i32(int32_t *inout, int64_t n, void *tmp, size_t tmp_bytes, void *stream);
/* The same prefix sum for a caller that has nowhere to get `tmp` from.  No counterpart in the reference.  Some hosts can
 * neither cudaMalloc nor carve scratch out of an allocator of their own: all device memory was given to a fixed arena
 * when the application started, or the call sits in a fixed-rate loop where the device-wide stall of cudaMalloc /
 * cudaFree is not acceptable.  For them the scan's bookkeeping (tile descriptors, ticket counter) lives in an array
 * compiled into the library: about 1 MB on every GPU that loads libqdsp.so, kept until the process exits.  The call
 * allocates and frees nothing, on the device or through any library; inout is the caller's, as above.
 *   - n may be at most qdsp_scan_static_max_n() (2^26 elements); a longer array is refused with QDSP_ERR_INVALID and a
 *     message that names the limit.  Use qdsp_inclusive_scan_i32 for those.
 *   - The first call on each device looks up where that device keeps the array (with that device current).  Make this
 *     first call once during start-up, outside any stream capture.
 *   - There is ONE array per device and every call uses it.  Each call clears what it needs first (stream-ordered), so
 *     back-to-back calls on one stream are fine; two calls that can run on the GPU at the same time (different streams,
 *     different host threads without ordering) race on it and give wrong sums. */
int64_t qdsp_scan_static_max_n(void);
int qdsp_inclusive_scan_i32_static(int32_t *inout, int64_t n, void *stream);


/* ---- qd.algorithms device-wide ops ------------------------------------------------------------ */
/* The reference's modern algorithm library: reduce_{add,min,max} (python/quadrants/algorithms/_reduce.py:448-507),
 * exclusive_scan_{add,min,max} (_scan.py:496-541), select (_select.py:124-169), reduce_by_key_add
 * (_reduce_by_key.py:145-200) and sort (_radix_sort.py:277-398).  Shared conventions, mirroring the reference's:
 *   capacity  the length of the arrays (sizes the grid and the scratch; the reference's 256**log256_max_n bound);
 *   n_dev     optional DEVICE pointer to the live i32 count, read by the kernels (no host round trip, so a call can be
 *             captured in a CUDA graph and replayed for any count <= capacity); NULL = the count is `capacity`;
 *   scratch   caller-owned, uninitialised device memory of qdsp_algo_scratch_bytes(op, capacity, elem) bytes;
 *   elem      qdsp_elem_type of the data.
 * Integer results are exact (add wraps); float add is evaluated in a FIXED tile/warp tree order, so results are
 * reproducible run to run (the reference's are not) but differ from a left-to-right host sum by rounding. */
typedef enum { QDSP_I32 = 0, QDSP_U32 = 1, QDSP_ELEM_F32 = 2, QDSP_I64 = 3, QDSP_U64 = 4, QDSP_ELEM_F64 = 5 } qdsp_elem_type;
typedef enum { QDSP_OP_ADD = 0, QDSP_OP_MIN = 1, QDSP_OP_MAX = 2 } qdsp_reduce_op;
typedef enum { QDSP_ALGO_REDUCE = 0, QDSP_ALGO_SCAN = 1, QDSP_ALGO_SELECT = 2, QDSP_ALGO_REDUCE_BY_KEY = 3,
               QDSP_ALGO_SORT = 4 } qdsp_algo;
/* Scratch bytes of one op at `capacity` elements (elem_bytes: 4 or 8; for QDSP_ALGO_SORT the key width). */
size_t qdsp_algo_scratch_bytes(int algo, int64_t capacity, int elem_bytes);
/* out[0] = op(in[0:n]); the identity of `op` when n == 0 (0 / +max,+inf / -max,-inf,0 like _reduce.py:137-203). */
int qdsp_reduce(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op, void *scratch,
                size_t scratch_bytes, void *stream);
/* out[i] = op(in[0:i]), out[0] = identity; `out` must not alias `in` (_scan.py:476-494). */
int qdsp_exclusive_scan(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op,
                        void *scratch, size_t scratch_bytes, void *stream);
/* Stream compaction (_select.py:124-147): out[k] = in[i] for the k-th i with flags[i] == 1, input order kept;
 * num_out[0] = number selected.  flags are i32 and must be exactly 0 or 1.  item_bytes: a multiple of 4 (scalars and
 * packed structs alike -- the copy is untyped). */
int qdsp_select(const void *in, const int32_t *flags, void *out, int32_t *num_out, int64_t capacity,
                const int32_t *n_dev, int item_bytes, void *scratch, size_t scratch_bytes, void *stream);
/* Run-length collapse (_reduce_by_key.py:145-177): every maximal run of equal consecutive keys becomes
 * (key, sum of its values); num_runs[0] = number of runs; only the first num_runs entries of the outputs are written.
 * key_elem / value_elem in {QDSP_I32, QDSP_U32, QDSP_ELEM_F32}; float keys compare as floats (NaN opens a run of its
 * own, +0 == -0). */
int qdsp_reduce_by_key_add(const void *keys_in, const void *values_in, void *keys_out, void *values_out,
                           int32_t *num_runs, int64_t capacity, const int32_t *n_dev, int key_elem, int value_elem,
                           void *scratch, size_t scratch_bytes, void *stream);
/* sort (_radix_sort.py:277-344): ascending stable LSD radix sort of the low `end_bit` bits of keys[0:n] (values
 * permuted alike when value_bytes is 4 or 8; 0 = keys only), ping-ponging through the caller's tmp_keys / tmp_values.
 * end_bit: positive multiple of 8 giving an EVEN number of digit passes so the result lands in keys / values.
 * key_elem: any qdsp_elem_type (numeric order: two's complement and IEEE twiddles applied on the fly; -0 < +0, NaNs
 * by bit pattern at the ends). */
int qdsp_sort(void *keys, void *tmp_keys, void *values, void *tmp_values, int64_t capacity, const int32_t *n_dev,
              int key_elem, int value_bytes, int end_bit, void *scratch, size_t scratch_bytes, void *stream);

/* ---- triplet emission ------------------------------------------------------------------------- */
/* Bulk replacement of per-triplet insert_triplet_f32/f64 (internal_functions.h:12-39: one contended atomic_add on
 * slot 0 per triplet): appends `count` triplets (row[k], col[k], +-val[k]) behind the triplets already in the buffer,
 * in array order, with NO atomic -- every thread reads slot 0, writes its triplet at slot0 + k, and a second
 * one-thread kernel bumps slot 0, all stream-ordered.  row/col: device i32; val: device f32/f64 (the builder's
 * dtype).  Overflow beyond max_triplets is detected on the device: nothing is written, slot 0 is still advanced (as the
 * reference's insert_triplet does), so qdsp_builder_count then exceeds max_triplets and qdsp_builder_build reports
 * QDSP_ERR_INVALID. */
int qdsp_builder_append(qdsp_builder_t b, const int32_t *row, const int32_t *col, const void *val, int64_t count,
                        int negate, void *stream);
/* Element-wise assembly (the FEM pattern of tests/python/test_implicit_fem.py:179-204): `n_elems` elements of `k`
 * nodes each; element e contributes the k*k triplets (nodes[e*k+a], nodes[e*k+b], mats[(e*k+a)*k+b]) in that order.
 * nodes: device i32 [n_elems*k]; mats: device f32/f64 [n_elems*k*k]. */
int qdsp_builder_append_elements(qdsp_builder_t b, const int32_t *nodes, const void *mats, int64_t n_elems, int k,
                                 void *stream);

/* ---- SparseSolver (quadrants/program/sparse_solver.h:22-136) ---------------------------------------------------------
 * make_cusparse_solver (sparse_solver.cpp:548-558) + CuSparseSolver.  solver_type: 0 "LLT", 1 "LDLT", 2 "LU";
 * ordering: 0 "AMD", 1 "COLAMD" -- both select the same symmetric minimum-degree ordering of A + A^T, as the reference's
 * arch=cuda path does (it ignores the argument and calls csrsymamdHost, sparse_solver.cpp:200).  dtype: f32 or f64 (the
 * reference's cuSOLVER path is f32 only).  LLT / LDLT read the lower triangle of the matrix (Eigen::Lower,
 * sparse_solver.h:9-13); LU pivots STATICALLY: analyze_pattern reads the values of the matrix it is given and fixes a row
 * permutation with a large entry of every row on the diagonal (qdsp_solver_lu_row_matching_host), the factorisation then
 * runs on the symmetrised pattern of the permuted matrix; a pivot that is zero or below 16 eps of its column (of L below, of U above) is reported
 * through *success / info(), never divided by.  (The reference exchanges rows during the factorisation.) */
typedef struct qdsp_solver_s *qdsp_solver_t;
int qdsp_solver_create(int dtype, int solver_type, int ordering, qdsp_solver_t *out);
/* analyze_pattern (sparse_solver.cpp:233-288): ordering + symbolic factorisation on the host, like the reference's
 * reorder() (:169-231); uploads the factor pattern.  Synchronises `stream`. */
int qdsp_solver_analyze_pattern(qdsp_solver_t s, qdsp_csr_t m, void *stream);
/* factorize (sparse_solver.cpp:290-364): numeric factorisation on the device in one launch.  *success = 0 when a pivot
 * is zero (LLT: not positive) -- the reference asserts singularity == -1 (:326-327, :359-360); *zero_pivot then holds the
 * first failing pivot in the permuted numbering (the csrcholZeroPivot convention), else -1.  Synchronises `stream`. */
int qdsp_solver_factorize(qdsp_solver_t s, qdsp_csr_t m, void *stream, int *success, int *zero_pivot);
/* solve_rf (sparse_solver.cpp:366-476): x = A^-1 b, device vectors of n elements (permutation, both triangular solves
 * and the inverse permutation on the device; the reference's LU branch round-trips through the host). */
int qdsp_solver_solve(qdsp_solver_t s, const void *b, void *x, void *stream);
/* The same through host buffers: EigenSparseSolver::solve(b) as the Python layer calls it with numpy arrays
 * (python/quadrants/linalg/sparse_solver.py:103-106). */
int qdsp_solver_solve_host(qdsp_solver_t s, const void *b_host, void *x_host, void *stream);
/* info (sparse_solver.cpp:120-123): 1 after a successful factorisation.  (QD_NOT_IMPLEMENTED on the reference's
 * arch=cuda path, sparse_solver.h:120-122.) */
int qdsp_solver_info(qdsp_solver_t s, int *success);
/* Size of the analysed factor: entries of L including the diagonal, height of the elimination tree (the length of the
 * dependency chain the device kernels walk), and sum over columns of (column length)^2 ~ multiply-adds of the
 * factorisation. */
int qdsp_solver_stats(qdsp_solver_t s, int64_t *factor_nnz, int *etree_height, double *flops);
int qdsp_solver_destroy(qdsp_solver_t s);
/* Pure host helper (no CUDA needed) behind qdsp_solver_analyze_pattern, exposed so that the ordering and the symbolic
 * factorisation are testable without a GPU.  perm_out[new] = old (n); parent_out: elimination tree (n, -1 = root);
 * colptr_out (n + 1) / rowidx_out (up to rowidx_capacity, may be NULL): the factor pattern by columns, rows ascending,
 * diagonal first.  natural_order != 0 skips the ordering (identity permutation). */
int qdsp_solver_analyze_host(int n, const int32_t *indptr, const int32_t *indices, int natural_order, int32_t *perm_out,
                             int32_t *parent_out, int32_t *colptr_out, int32_t *rowidx_out, int64_t rowidx_capacity,
                             int64_t *factor_nnz, int *etree_height);
/* Pure host helper (no CUDA needed): the row permutation qdsp_solver_analyze_pattern applies for solver_type LU.  The
 * reference's LU exchanges rows during the factorisation (Eigen SparseLU; csrluHost with pivot threshold 1.0,
 * sparse_solver.cpp:343-359); here the exchange is decided once, before the symbolic analysis (static pivoting): the
 * rows are matched to columns so that the product of the chosen entries' magnitudes, each relative to its row's largest,
 * is maximal (the MC64 objective; Hungarian method on the sparse pattern), a diagonal entry counting 10x so that rows
 * only move where it pays; a diagonal that is its row's largest entry stays outright.  abs_values[q] = |value| of
 * stored entry q; rowof_out[i] = the row of A that becomes row i; *zero_free_diagonal = 0 when the matrix is
 * structurally singular. */
int qdsp_solver_lu_row_matching_host(int n, const int32_t *indptr, const int32_t *indices, const double *abs_values,
                                     int32_t *rowof_out, int *zero_free_diagonal);

/* ---- dense vector kernels of MatrixFreeCG / MatrixFreeBICGSTAB (python/quadrants/linalg/matrixfree_cg.py:23-310) --------
 * The reference JIT-compiles `reduce`, `update_x`, `update_r`, `update_p`, ... per solve; these are their device
 * counterparts.  Vectors are dense device arrays of n elements of `dtype`; dot products are accumulated in double in a
 * fixed order (reproducible) and returned to the host (the reference's reduce kernels return host scalars too). */
/* out[0] = a.b and, when c and d are given, out[1] = c.d, in one pass. */
int qdsp_vec_dot2(const void *a, const void *b, const void *c, const void *d, int64_t n, int dtype, void *stream,
                  double *out);
/* out = sum_{i<k} coefs[i] * xs[i], 1 <= k <= 4; `out` may alias any xs[i]. */
int qdsp_vec_lincomb(void *out, int64_t n, int dtype, int k, const void *const *xs, const double *coefs, void *stream);
/* update_x + update_r + reduce(r, r) of the CG loop (matrixfree_cg.py:84-98,116-119) fused into one pass:
 * x += alpha p, r -= alpha ap, *rr = r.r. */
int qdsp_vec_cg_step(void *x, void *r, const void *p, const void *ap, double alpha, int64_t n, int dtype, void *stream,
                     double *rr);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* QDSP_H_ */
