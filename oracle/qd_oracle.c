/*
 * qd_oracle.c -- CPU restatement of the Quadrants sparse linear-algebra hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under quadrants_b200/ may import, link or
 * call this file; it is the checker used by tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs.  The product path is the
 * CUDA library (quadrants_b200/lib/libqdsp.so) and fails loudly without it.
 *
 * Parity status: PINNED.  Every function below is checked (tests/test_oracle_*.py)
 *   (1) against the reference's own golden vectors, and
 *   (2) bit-for-bit (build, SpMV, sort, scan) / to tolerance (CG) against
 *       oracle/_ref/libeigen_ref.so, i.e. the reference's vendored Eigen
 *       compiled from /root/reference/external/eigen with the reference's flags.
 *
 * Each function cites the reference file:line it restates (paths relative to
 * /root/reference).  Plain C99, no dependencies; built by oracle/Makefile with
 *   gcc -O2 -march=x86-64-v3 -ffp-contract=off
 * Floating-point contraction is therefore explicit: where the reference's
 * shipped build (-O3 -march=x86-64-v3, cmake/QuadrantsCXXFlags.cmake:123-124,
 * GCC default -ffp-contract=fast) fuses a*b+c, this file calls fmaf()/fma().
 */
#include <math.h>
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define QDO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------- */
/* A3: triplet buffer -> compressed matrix.                                   */
/* Buffer layout (quadrants/runtime/llvm/runtime_module/internal_functions.h: */
/* 12-21): slot 0 = count, then (row, col, value-bits) triples; slot width 4   */
/* bytes for f32 builders, 8 bytes for f64 (quadrants/program/                 */
/* sparse_matrix.cpp:139-152,159-166).                                         */
/* Algorithm: Eigen::internal::set_from_triplets + collapseDuplicates +       */
/* storage-order-changing assignment (external/eigen/Eigen/src/SparseCore/    */
/* SparseMatrix.h:1125-1173 and 1475-1511): bucket by the OTHER index in      */
/* input order, add later duplicates onto the first occurrence in input       */
/* order, then counting-sort into the requested order.  Explicit zeros stay.  */
/* ------------------------------------------------------------------------- */
#define DEFINE_BUILD(NAME, SLOT_T, VAL_T)                                                          \
  QDO_API int64_t NAME(const SLOT_T *buf, int rows, int cols, int row_major, int32_t *outer,       \
                       int32_t *inner, VAL_T *vals) {                                              \
    const int64_t count = (int64_t)buf[0];                                                         \
    const SLOT_T *t = buf + 1;                                                                     \
    const int out_n = row_major ? rows : cols; /* outer size of the result        */               \
    const int tr_n = row_major ? cols : rows;  /* outer size of the transposed tmp */              \
    memset(outer, 0, sizeof(int32_t) * ((size_t)out_n + 1));                                       \
    if (count <= 0)                                                                                \
      return 0;                                                                                    \
    int32_t *tr_outer = (int32_t *)calloc((size_t)tr_n + 1, sizeof(int32_t));                      \
    int32_t *tr_inner = (int32_t *)malloc(sizeof(int32_t) * (size_t)count);                        \
    VAL_T *tr_val = (VAL_T *)malloc(sizeof(VAL_T) * (size_t)count);                                \
    const int tmp_n = rows > cols ? rows : cols;                                                   \
    int32_t *tmp = (int32_t *)malloc(sizeof(int32_t) * ((size_t)tmp_n + 1));                       \
    /* pass 1: count per transposed-outer index */                                                 \
    for (int64_t k = 0; k < count; ++k) {                                                          \
      const int32_t r = (int32_t)t[3 * k], c = (int32_t)t[3 * k + 1];                              \
      if (r < 0 || r >= rows || c < 0 || c >= cols) {                                              \
        free(tr_outer); free(tr_inner); free(tr_val); free(tmp);                                   \
        return -1;                                                                                 \
      }                                                                                            \
      tr_outer[(row_major ? c : r) + 1]++;                                                         \
    }                                                                                              \
    for (int j = 0; j < tr_n; ++j) tr_outer[j + 1] += tr_outer[j]; /* std::partial_sum */          \
    memcpy(tmp, tr_outer, sizeof(int32_t) * (size_t)tr_n);                                         \
    /* pass 2: push each triplet to the back of its vector, input order */                         \
    for (int64_t k = 0; k < count; ++k) {                                                          \
      const int32_t r = (int32_t)t[3 * k], c = (int32_t)t[3 * k + 1];                              \
      const int32_t j = row_major ? c : r, i = row_major ? r : c;                                  \
      VAL_T v;                                                                                     \
      memcpy(&v, &t[3 * k + 2], sizeof(VAL_T)); /* quadrants_union_cast<T> */                      \
      const int32_t pos = tmp[j]++;                                                                \
      tr_inner[pos] = i;                                                                           \
      tr_val[pos] = v;                                                                             \
    }                                                                                              \
    /* collapseDuplicates: wi[inner] = position of the primary entry */                            \
    const int tr_inner_n = row_major ? rows : cols;                                                \
    for (int i = 0; i < tr_inner_n; ++i) tmp[i] = -1;                                              \
    int32_t cnt = 0;                                                                               \
    for (int j = 0; j < tr_n; ++j) {                                                               \
      const int32_t new_begin = cnt, end = tr_outer[j + 1];                                        \
      for (int32_t k = tr_outer[j]; k < end; ++k) {                                                \
        const int32_t i = tr_inner[k];                                                             \
        if (tmp[i] >= new_begin) {                                                                 \
          tr_val[tmp[i]] = tr_val[tmp[i]] + tr_val[k]; /* scalar_sum_op, input order */            \
        } else {                                                                                   \
          tr_inner[cnt] = i;                                                                       \
          tr_val[cnt] = tr_val[k];                                                                 \
          tmp[i] = cnt++;                                                                          \
        }                                                                                          \
      }                                                                                            \
      tr_outer[j] = new_begin;                                                                     \
    }                                                                                              \
    tr_outer[tr_n] = cnt;                                                                          \
    /* mat = trmat: counting sort by the transposed-inner index (implicit sort) */                 \
    for (int32_t k = 0; k < cnt; ++k) outer[tr_inner[k] + 1]++;                                    \
    for (int i = 0; i < out_n; ++i) outer[i + 1] += outer[i];                                      \
    memcpy(tmp, outer, sizeof(int32_t) * (size_t)out_n);                                           \
    for (int j = 0; j < tr_n; ++j) {                                                               \
      for (int32_t k = tr_outer[j]; k < tr_outer[j + 1]; ++k) {                                    \
        const int32_t pos = tmp[tr_inner[k]]++;                                                    \
        inner[pos] = j;                                                                            \
        vals[pos] = tr_val[k];                                                                     \
      }                                                                                            \
    }                                                                                              \
    free(tr_outer); free(tr_inner); free(tr_val); free(tmp);                                       \
    return (int64_t)cnt;                                                                           \
  }

DEFINE_BUILD(qdo_build_f32, int32_t, float)
DEFINE_BUILD(qdo_build_f64, int64_t, double)

/* ------------------------------------------------------------------------- */
/* A7: y = A x, EigenSparseMatrix::spmv (quadrants/program/sparse_matrix.cpp: */
/* 265-279) -> Eigen sparse_time_dense_product_impl (external/eigen/Eigen/src/ */
/* SparseCore/SparseDenseProduct.h:34-81 row-major: two alternating           */
/* accumulators, res += alpha*(a+b); :94-110 col-major: scatter per column).  */
/* y is the zero-filled result ndarray (python/quadrants/linalg/              */
/* sparse_matrix.py:157), so the function zeroes it first.                    */
/* ------------------------------------------------------------------------- */
#define DEFINE_SPMV(NAME, VAL_T, FMA_A, FMA_B, FMA)                                                              \
  QDO_API void NAME(int row_major, int rows, int cols, const int32_t *outer, const int32_t *inner, \
                    const VAL_T *vals, const VAL_T *x, VAL_T *y) {                                 \
    for (int i = 0; i < rows; ++i) y[i] = (VAL_T)0;                                                \
    if (row_major) {                                                                               \
      for (int i = 0; i < rows; ++i) {                                                             \
        VAL_T a = (VAL_T)0, b = (VAL_T)0;                                                          \
        int32_t k = outer[i];                                                                      \
        const int32_t e = outer[i + 1];                                                            \
        while (k < e) {                                                                            \
          a = FMA_A(vals[k], x[inner[k]], a);                                                        \
          ++k;                                                                                     \
          if (k < e) {                                                                             \
            b = FMA_B(vals[k], x[inner[k]], b);                                                      \
            ++k;                                                                                   \
          }                                                                                        \
        }                                                                                          \
        y[i] += (VAL_T)1 * (a + b);                                                                \
      }                                                                                            \
    } else {                                                                                       \
      for (int j = 0; j < cols; ++j) {                                                             \
        const VAL_T xj = (VAL_T)1 * x[j];                                                          \
        for (int32_t k = outer[j]; k < outer[j + 1]; ++k)                                          \
          y[inner[k]] = FMA(vals[k], xj, y[inner[k]]);                                             \
      }                                                                                            \
    }                                                                                              \
  }

#define NOFMAF(a, b, c) ((float)((float)((a) * (b)) + (c)))
#define NOFMAD(a, b, c) ((double)((double)((a) * (b)) + (c)))
/* Contraction as the reference's shipped flags produce it with GCC 13.3 (verified bit-for-bit against oracle/_ref):
 * col-major scatter: fused; row-major: accumulator a is mul+add, accumulator b is fused.  The row-major pattern is
 * a code-generation accident, not a contract; the CUDA path is held to the col-major (default storage_format)
 * sequence bit-for-bit and to BASELINE.json's 1e-5 / 1e-12 per-row tolerance otherwise. */
DEFINE_SPMV(qdo_spmv_f32, float, NOFMAF, fmaf, fmaf)
DEFINE_SPMV(qdo_spmv_f64, double, NOFMAD, fma, fma)

/* Same product without any contraction (a build of the reference for a CPU without FMA); used by tests to bound
 * the effect of contraction. */
DEFINE_SPMV(qdo_spmv_nofma_f32, float, NOFMAF, NOFMAF, NOFMAF)
DEFINE_SPMV(qdo_spmv_nofma_f64, double, NOFMAD, NOFMAD, NOFMAD)

/* ------------------------------------------------------------------------- */
/* A9: CG<EigenT,DT>::solve (quadrants/program/conjugate_gradient.h:36-49) -> */
/* Eigen::ConjugateGradient<SparseMatrix<DT>, Lower|Upper, Diagonal-          */
/* Preconditioner>: zero start (IterativeSolverBase.h:357-359), Jacobi        */
/* (BasicPreconditioners.h:62-74: 1/a_jj, or 1 where the diagonal is missing  */
/* or zero), recurrence and stopping rule of internal::conjugate_gradient     */
/* (ConjugateGradient.h:29-85), success iff error <= tol (:209).              */
/* The matrix is the col-major Eigen matrix viewed as its transpose           */
/* (ConjugateGradient.h:192-207), i.e. the compressed arrays are walked with  */
/* the ROW-major product kernel; `outer/inner/vals` are those arrays.         */
/* Reductions follow Eigen's AVX order: dot = 4 packet accumulators with      */
/* pmadd (Core/InnerProduct.h inner_product_impl<.,true>), squaredNorm = 2    */
/* packet accumulators (Core/Redux.h:275-324), predux = (lo+hi) then pairwise */
/* (arch/AVX/Reductions.h:131-135, arch/SSE/Reductions.h:206-217,267-273).    */
/* `fl` selects, per site, whether the shipped build contracts mul+add into   */
/* an fma (bit 0 spmv acc a, 1 spmv acc b, 2 x+=ap, 3 r-=a*tmp, 4 squaredNorm */
/* packets, 5 p=z+bp, 6 scalar tails); QDO_CG_FLAGS_DEFAULT is the pattern    */
/* found bit-exact against oracle/_ref (tests/test_oracle_vs_eigen.py).       */
/* Returns info (0 = Success, 2 = NoConvergence like Eigen's enum).           */
/* ------------------------------------------------------------------------- */
#define DEFINE_CG(NAME, VAL_T, PS, FMA, SQRT, MINPOS)                                              \
  static inline VAL_T NAME##_predux(const VAL_T *q) {                                              \
    if (PS == 8) {                                                                                 \
      const VAL_T h0 = q[0] + q[4], h1 = q[1] + q[5], h2 = q[2] + q[6], h3 = q[3] + q[7];          \
      return (h0 + h2) + (h1 + h3);                                                                \
    }                                                                                              \
    return (q[0] + q[2]) + (q[1] + q[3]);                                                          \
  }                                                                                                \
  static VAL_T NAME##_dot(const VAL_T *a, const VAL_T *b, int64_t size, int fl) {                  \
    if (size == 0) return 0;                                                                       \
    if (size < PS) {                                                                               \
      VAL_T r = a[0] * b[0];                                                                       \
      for (int64_t k = 1; k < size; ++k) r = FMA(a[k], b[k], r);                                   \
      return r;                                                                                    \
    }                                                                                              \
    const int64_t packet_end = size / PS * PS, quad_end = size / (4 * PS) * (4 * PS);              \
    const int64_t num_packets = size / PS, num_rem = (packet_end - quad_end) / PS;                 \
    VAL_T p[4][PS];                                                                                \
    for (int q = 0; q < 4 && q < num_packets; ++q)                                                 \
      for (int l = 0; l < PS; ++l) p[q][l] = a[q * PS + l] * b[q * PS + l];                        \
    if (num_packets >= 4) {                                                                        \
      for (int64_t k = 4 * PS; k < quad_end; k += 4 * PS)                                          \
        for (int q = 0; q < 4; ++q)                                                                \
          for (int l = 0; l < PS; ++l) p[q][l] = FMA(a[k + q * PS + l], b[k + q * PS + l], p[q][l]); \
      for (int q = 0; q < num_rem; ++q)                                                            \
        for (int l = 0; l < PS; ++l)                                                               \
          p[q][l] = FMA(a[quad_end + q * PS + l], b[quad_end + q * PS + l], p[q][l]);              \
      for (int l = 0; l < PS; ++l) p[2][l] = p[2][l] + p[3][l];                                    \
    }                                                                                              \
    if (num_packets >= 3) for (int l = 0; l < PS; ++l) p[1][l] = p[1][l] + p[2][l];                \
    if (num_packets >= 2) for (int l = 0; l < PS; ++l) p[0][l] = p[0][l] + p[1][l];                \
    VAL_T r = NAME##_predux(p[0]);                                                                 \
    for (int64_t k = packet_end; k < size; ++k) r = FMA(a[k], b[k], r);                            \
    (void)fl;                                                                                      \
    return r;                                                                                      \
  }                                                                                                \
  static VAL_T NAME##_sqnorm(const VAL_T *a, int64_t size, int fl) {                               \
    const int fp = (fl >> 4) & 1, ft = (fl >> 6) & 1;                                              \
    const int64_t aligned2 = size / (2 * PS) * (2 * PS), aligned = size / PS * PS;                 \
    VAL_T r;                                                                                       \
    if (aligned) {                                                                                 \
      VAL_T p0[PS], p1[PS];                                                                        \
      for (int l = 0; l < PS; ++l) p0[l] = a[l] * a[l];                                            \
      if (aligned > PS) {                                                                          \
        for (int l = 0; l < PS; ++l) p1[l] = a[PS + l] * a[PS + l];                                \
        for (int64_t k = 2 * PS; k < aligned2; k += 2 * PS)                                        \
          for (int l = 0; l < PS; ++l) {                                                           \
            p0[l] = fp ? FMA(a[k + l], a[k + l], p0[l]) : p0[l] + a[k + l] * a[k + l];             \
            p1[l] = fp ? FMA(a[k + PS + l], a[k + PS + l], p1[l])                                  \
                       : p1[l] + a[k + PS + l] * a[k + PS + l];                                    \
          }                                                                                        \
        for (int l = 0; l < PS; ++l) p0[l] = p0[l] + p1[l];                                        \
        if (aligned > aligned2)                                                                    \
          for (int l = 0; l < PS; ++l)                                                             \
            p0[l] = fp ? FMA(a[aligned2 + l], a[aligned2 + l], p0[l])                              \
                       : p0[l] + a[aligned2 + l] * a[aligned2 + l];                                \
      }                                                                                            \
      r = NAME##_predux(p0);                                                                       \
      for (int64_t k = aligned; k < size; ++k) r = ft ? FMA(a[k], a[k], r) : r + a[k] * a[k];      \
    } else {                                                                                       \
      r = a[0] * a[0];                                                                             \
      for (int64_t k = 1; k < size; ++k) r = ft ? FMA(a[k], a[k], r) : r + a[k] * a[k];            \
    }                                                                                              \
    return r;                                                                                      \
  }                                                                                                \
  static void NAME##_spmv(int n, const int32_t *outer, const int32_t *inner, const VAL_T *vals,    \
                          const VAL_T *x, VAL_T *y, int fl) {                                      \
    const int fa = fl & 1, fb = (fl >> 1) & 1;                                                     \
    for (int i = 0; i < n; ++i) {                                                                  \
      VAL_T a = 0, b = 0;                                                                          \
      int32_t k = outer[i];                                                                        \
      const int32_t e = outer[i + 1];                                                              \
      while (k < e) {                                                                              \
        a = fa ? FMA(vals[k], x[inner[k]], a) : a + vals[k] * x[inner[k]];                         \
        ++k;                                                                                       \
        if (k < e) {                                                                               \
          b = fb ? FMA(vals[k], x[inner[k]], b) : b + vals[k] * x[inner[k]];                       \
          ++k;                                                                                     \
        }                                                                                          \
      }                                                                                            \
      y[i] = (VAL_T)0 + (VAL_T)1 * (a + b);                                                        \
    }                                                                                              \
  }                                                                                                \
  QDO_API int NAME##_flags(int n, const int32_t *outer, const int32_t *inner, const VAL_T *vals,   \
                           const VAL_T *b, VAL_T *x, int max_iters, VAL_T tol, int fl,             \
                           int *iters_out, VAL_T *err_out) {                                       \
    VAL_T *r = (VAL_T *)malloc(sizeof(VAL_T) * (size_t)n * 5);                                     \
    VAL_T *p = r + n, *z = p + n, *tmp = z + n, *invdiag = tmp + n;                                \
    const int64_t nv = (int64_t)n / PS * PS; /* packet part; the rest is Eigen's scalar tail */    \
    const int fx = (fl >> 2) & 1, fr = (fl >> 3) & 1, fpp = (fl >> 5) & 1, ft = (fl >> 6) & 1;     \
    for (int j = 0; j < n; ++j) {                                                                  \
      int32_t k = outer[j];                                                                        \
      while (k < outer[j + 1] && inner[k] != j) ++k;                                               \
      invdiag[j] = (k < outer[j + 1] && vals[k] != (VAL_T)0) ? (VAL_T)1 / vals[k] : (VAL_T)1;      \
    }                                                                                              \
    for (int i = 0; i < n; ++i) x[i] = (VAL_T)0;                                                   \
    NAME##_spmv(n, outer, inner, vals, x, tmp, fl);                                                \
    for (int i = 0; i < n; ++i) r[i] = b[i] - tmp[i];                                              \
    const VAL_T rhs2 = NAME##_sqnorm(b, n, fl);                                                    \
    VAL_T res2 = 0, err;                                                                           \
    int it = 0;                                                                                    \
    if (rhs2 == (VAL_T)0) {                                                                        \
      err = 0;                                                                                     \
      goto done;                                                                                   \
    }                                                                                              \
    {                                                                                              \
      VAL_T thr = tol * tol * rhs2;                                                                \
      if (thr < MINPOS) thr = MINPOS;                                                              \
      res2 = NAME##_sqnorm(r, n, fl);                                                              \
      if (res2 < thr) { err = SQRT(res2 / rhs2); goto done; }                                      \
      for (int i = 0; i < n; ++i) p[i] = invdiag[i] * r[i];                                        \
      VAL_T abs_new = NAME##_dot(r, p, n, fl);                                                     \
      while (it < max_iters) {                                                                     \
        NAME##_spmv(n, outer, inner, vals, p, tmp, fl);                                            \
        const VAL_T alpha = abs_new / NAME##_dot(p, tmp, n, fl);                                   \
        for (int64_t i = 0; i < n; ++i) {                                                          \
          const int f1 = i < nv ? fx : ft, f2 = i < nv ? fr : ft;                                  \
          x[i] = f1 ? FMA(alpha, p[i], x[i]) : x[i] + alpha * p[i];                                \
          r[i] = f2 ? FMA(-alpha, tmp[i], r[i]) : r[i] - alpha * tmp[i];                           \
        }                                                                                          \
        res2 = NAME##_sqnorm(r, n, fl);                                                            \
        if (res2 < thr) break;                                                                     \
        for (int i = 0; i < n; ++i) z[i] = invdiag[i] * r[i];                                      \
        const VAL_T abs_old = abs_new;                                                             \
        abs_new = NAME##_dot(r, z, n, fl);                                                         \
        const VAL_T beta = abs_new / abs_old;                                                      \
        for (int64_t i = 0; i < n; ++i) {                                                          \
          const int f3 = i < nv ? fpp : ft;                                                        \
          p[i] = f3 ? FMA(beta, p[i], z[i]) : z[i] + beta * p[i];                                  \
        }                                                                                          \
        ++it;                                                                                      \
      }                                                                                            \
      err = SQRT(res2 / rhs2);                                                                     \
    }                                                                                              \
  done:                                                                                            \
    free(r);                                                                                       \
    *iters_out = it;                                                                               \
    *err_out = err;                                                                                \
    return err <= tol ? 0 : 2;                                                                     \
  }                                                                                                \
  QDO_API int NAME(int n, const int32_t *outer, const int32_t *inner, const VAL_T *vals,           \
                   const VAL_T *b, VAL_T *x, int max_iters, VAL_T tol, int *iters_out,             \
                   VAL_T *err_out) {                                                               \
    return NAME##_flags(n, outer, inner, vals, b, x, max_iters, tol, QDO_CG_FLAGS_DEFAULT,         \
                        iters_out, err_out);                                                       \
  }

#define QDO_CG_FLAGS_DEFAULT 0x3e /* a: mul+add, b/x/r/sqnorm/p: fma, scalar tails: mul+add */
DEFINE_CG(qdo_cg_f32, float, 8, fmaf, sqrtf, FLT_MIN)
DEFINE_CG(qdo_cg_f64, double, 4, fma, sqrt, DBL_MIN)

/* ------------------------------------------------------------------------- */
/* A10: CUCG::solve (quadrants/program/conjugate_gradient.cpp:19-83): the     */
/* reference's CUDA-arch semantics restated on the CPU in fp32 -- plain CG,   */
/* x arrives zero-filled (python/quadrants/linalg/sparse_cg.py:53), stop when */
/* r.r <= tol^2 or k > max_iters, no success flag.  `outer/inner/vals` is CSR.*/
/* Returns the number of loop bodies executed.                                */
/* ------------------------------------------------------------------------- */
QDO_API int qdo_cucg_f32(int n, const int32_t *outer, const int32_t *inner, const float *vals,
                         const float *b, float *x, int max_iters, float tol, float *r1_out) {
  float *r = (float *)malloc(sizeof(float) * (size_t)n * 3);
  float *p = r + n, *ax = p + n;
  qdo_spmv_f32(1, n, n, outer, inner, vals, x, ax);
  float r1 = 0.f, r0 = 0.f;
  for (int i = 0; i < n; ++i) { r[i] = b[i] - ax[i]; }
  for (int i = 0; i < n; ++i) r1 += r[i] * r[i];
  int k = 1;
  while (r1 > tol * tol && k <= max_iters) {
    if (k > 1) {
      const float beta = r1 / r0;
      for (int i = 0; i < n; ++i) p[i] = beta * p[i] + r[i];
    } else {
      memcpy(p, r, sizeof(float) * (size_t)n);
    }
    qdo_spmv_f32(1, n, n, outer, inner, vals, p, ax);
    float dot = 0.f;
    for (int i = 0; i < n; ++i) dot += p[i] * ax[i];
    const float a = r1 / dot;
    for (int i = 0; i < n; ++i) { x[i] += a * p[i]; r[i] -= a * ax[i]; }
    r0 = r1;
    r1 = 0.f;
    for (int i = 0; i < n; ++i) r1 += r[i] * r[i];
    ++k;
  }
  free(r);
  *r1_out = r1;
  return k - 1;
}

/* ------------------------------------------------------------------------- */
/* A13: parallel_sort.  Two restatements:                                     */
/*  (a) the exact comparator network of python/quadrants/algorithms/          */
/*      _algorithms.py:41-56 + python/quadrants/_kernels.py:227-253 (Batcher   */
/*      odd-even merge; comparators of one stage are disjoint, so a serial    */
/*      replay equals the parallel kernel).  O(N log^2 N): small N only.      */
/*  (b) a stable ascending key/value sort -- the contract pinned by the       */
/*      reference's own tests (tests/python/test_sort.py:9-62 sees only       */
/*      sortedness with values==keys; tests/python/test_algorithms.py:584-586 */
/*      pins stable order).  LSD radix, 8-bit digits, with the usual          */
/*      sign/float key twiddles so that order is numeric.                     */
/* ------------------------------------------------------------------------- */
#define DEFINE_NETWORK(NAME, KEY_T)                                                                \
  QDO_API int NAME(KEY_T *keys, int32_t *values, int use_values, int64_t N) {                      \
    int stages = 0;                                                                                \
    for (int64_t p = 1; p < N; p *= 2) {                                                           \
      for (int64_t k = p; k >= 1; k /= 2) {                                                        \
        const int64_t invocations = (N - k - k % p) / (2 * k) + 1;                                 \
        for (int64_t inv = 0; inv < invocations; ++inv) {                                          \
          const int64_t j = k % p + inv * 2 * k;                                                   \
          int64_t lim = N - j - k;                                                                 \
          if (k < lim) lim = k;                                                                    \
          for (int64_t i = 0; i < lim; ++i) {                                                      \
            const int64_t a = i + j, b2 = i + j + k;                                               \
            if (a / (p * 2) == b2 / (p * 2) && keys[a] > keys[b2]) {                               \
              KEY_T tk = keys[a]; keys[a] = keys[b2]; keys[b2] = tk;                               \
              if (use_values) { int32_t tv = values[a]; values[a] = values[b2]; values[b2] = tv; } \
            }                                                                                      \
          }                                                                                        \
        }                                                                                          \
        ++stages;                                                                                  \
      }                                                                                            \
    }                                                                                              \
    return stages;                                                                                 \
  }
DEFINE_NETWORK(qdo_network_sort_i32, int32_t)
DEFINE_NETWORK(qdo_network_sort_f32, float)

static inline uint32_t twiddle_i32(uint32_t k) { return k ^ 0x80000000u; }
static inline uint32_t twiddle_f32(uint32_t k) { return (k & 0x80000000u) ? ~k : (k | 0x80000000u); }
static inline uint32_t untwiddle_f32(uint32_t k) { return (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; }

/* key_kind: 0 = u32, 1 = i32, 2 = f32.  values may be NULL.  value_words = 1 or 2 (4/8-byte payload). */
QDO_API void qdo_stable_sort_u32(uint32_t *keys, uint32_t *values, int value_words, int64_t n, int key_kind) {
  if (n <= 0) return;
  uint32_t *kt = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n);
  uint32_t *vt = values ? (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n * value_words) : NULL;
  for (int64_t i = 0; i < n; ++i)
    keys[i] = key_kind == 1 ? twiddle_i32(keys[i]) : key_kind == 2 ? twiddle_f32(keys[i]) : keys[i];
  uint32_t *ks = keys, *kd = kt, *vs = values, *vd = vt;
  for (int pass = 0; pass < 4; ++pass) {
    int64_t hist[257] = {0};
    const int sh = pass * 8;
    for (int64_t i = 0; i < n; ++i) hist[((ks[i] >> sh) & 255u) + 1]++;
    for (int d = 0; d < 256; ++d) hist[d + 1] += hist[d];
    for (int64_t i = 0; i < n; ++i) {
      const int64_t pos = hist[(ks[i] >> sh) & 255u]++;
      kd[pos] = ks[i];
      if (vs) for (int w = 0; w < value_words; ++w) vd[pos * value_words + w] = vs[i * value_words + w];
    }
    uint32_t *t = ks; ks = kd; kd = t;
    t = vs; vs = vd; vd = t;
  }
  /* 4 passes: data is back in keys/values */
  for (int64_t i = 0; i < n; ++i)
    keys[i] = key_kind == 1 ? twiddle_i32(keys[i]) : key_kind == 2 ? untwiddle_f32(keys[i]) : keys[i];
  free(kt);
  free(vt);
}

/* u64 keys (used for the (row,col) build keys), 4-byte payload. */
QDO_API void qdo_stable_sort_u64(uint64_t *keys, uint32_t *values, int64_t n) {
  if (n <= 0) return;
  uint64_t *kt = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)n);
  uint32_t *vt = values ? (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n) : NULL;
  uint64_t *ks = keys, *kd = kt;
  uint32_t *vs = values, *vd = vt;
  for (int pass = 0; pass < 8; ++pass) {
    int64_t hist[257] = {0};
    const int sh = pass * 8;
    for (int64_t i = 0; i < n; ++i) hist[((ks[i] >> sh) & 255u) + 1]++;
    for (int d = 0; d < 256; ++d) hist[d + 1] += hist[d];
    for (int64_t i = 0; i < n; ++i) {
      const int64_t pos = hist[(ks[i] >> sh) & 255u]++;
      kd[pos] = ks[i];
      if (vs) vd[pos] = vs[i];
    }
    uint64_t *t = ks; ks = kd; kd = t;
    uint32_t *u = vs; vs = vd; vd = u;
  }
  free(kt);
  free(vt);
}

/* ------------------------------------------------------------------------- */
/* A14: PrefixSumExecutor.run (python/quadrants/algorithms/_algorithms.py:    */
/* 109-163, python/quadrants/_kernels.py:298-352): in-place INCLUSIVE i32     */
/* prefix sum with two's-complement wrap-around; the reference test           */
/* (tests/python/test_scan.py:28-31) checks it against a sequential running   */
/* sum.  Integer addition is associative, so the hierarchical scan and this   */
/* serial loop agree bit-for-bit.                                             */
/* ------------------------------------------------------------------------- */
QDO_API void qdo_inclusive_scan_i32(int32_t *arr, int64_t n) {
  uint32_t acc = 0;
  for (int64_t i = 0; i < n; ++i) {
    acc += (uint32_t)arr[i];
    arr[i] = (int32_t)acc;
  }
}

/* CuSparseMatrix::get_element-style lookup on CSR (sparse_matrix.cpp:50-64 region): linear search, 0 if absent. */
QDO_API double qdo_csr_get_f32(const int32_t *outer, const int32_t *inner, const float *vals, int r, int c) {
  for (int32_t k = outer[r]; k < outer[r + 1]; ++k)
    if (inner[k] == c) return vals[k];
  return 0.0;
}

/* ------------------------------------------------------------------------- */
/* Direct solve.  Restates Eigen's simplicial Cholesky, the  */
/* solver behind EigenSparseSolver<SimplicialLLT/LDLT<.., Lower, ..>>          */
/* (quadrants/program/sparse_solver.h:9-13, sparse_solver.cpp:86-107;          */
/* Eigen/src/SparseCholesky/SimplicialCholesky_impl.h:284-383): elimination    */
/* tree, then row-by-row ("up-looking") factorisation -- row k of L is the     */
/* solution of a sparse triangular system whose pattern is the tree paths from */
/* the entries of A(0:k, k) -- followed by L y = b, D, L^T x = y.  Only the     */
/* LOWER triangle of A (given in CSR) is read.  Natural ordering: the fill-     */
/* reducing permutation changes rounding, not the solution, and the GPU path    */
/* is compared with a tolerance.  ldlt != 0: unit L and D; else L L^T.         */
/* Returns 0, or 1 + k when pivot k is zero (LDLT) / not positive (LLT).        */
/* Dense O(n^2) workspace: small systems only (test infrastructure).           */
/* ------------------------------------------------------------------------- */
#define QDO_DIRECT(NAME, VAL_T, SQRT)                                                              \
  QDO_API int NAME(int n, const int32_t *outer, const int32_t *inner, const VAL_T *vals, int ldlt, \
                   const VAL_T *b, VAL_T *x) {                                                     \
    VAL_T *L = (VAL_T *)calloc((size_t)n * n, sizeof(VAL_T)); /* row-major dense lower factor */   \
    VAL_T *D = (VAL_T *)calloc((size_t)n, sizeof(VAL_T));                                          \
    VAL_T *y = (VAL_T *)calloc((size_t)n, sizeof(VAL_T));                                          \
    int rc = 0;                                                                                    \
    for (int k = 0; k < n && rc == 0; ++k) {                                                       \
      /* scatter row k of the lower triangle (== column k of the upper one) */                     \
      for (int i = 0; i <= k; ++i) y[i] = (VAL_T)0;                                                \
      for (int32_t p = outer[k]; p < outer[k + 1]; ++p)                                            \
        if (inner[p] <= k) y[inner[p]] += vals[p];                                                 \
      VAL_T d = y[k];                                                                              \
      /* ascending i is a topological order of the row subtree */                                  \
      for (int i = 0; i < k; ++i) {                                                                \
        VAL_T yi = y[i];                                                                           \
        VAL_T lki;                                                                                 \
        if (ldlt) lki = yi / D[i];                                                                 \
        else { lki = yi / L[(size_t)i * n + i]; yi = lki; }                                        \
        for (int r = i + 1; r < k; ++r) y[r] -= L[(size_t)r * n + i] * yi;                         \
        d -= lki * yi;                                                                             \
        L[(size_t)k * n + i] = lki;                                                                \
      }                                                                                            \
      if (ldlt) {                                                                                  \
        D[k] = d;                                                                                  \
        L[(size_t)k * n + k] = (VAL_T)1;                                                           \
        if (d == (VAL_T)0) rc = 1 + k;                                                             \
      } else {                                                                                     \
        if (!(d > (VAL_T)0)) rc = 1 + k;                                                           \
        else L[(size_t)k * n + k] = SQRT(d);                                                       \
      }                                                                                            \
    }                                                                                              \
    if (rc == 0) {                                                                                 \
      for (int i = 0; i < n; ++i) {                                                                \
        VAL_T s = b[i];                                                                            \
        for (int j = 0; j < i; ++j) s -= L[(size_t)i * n + j] * x[j];                              \
        x[i] = s / L[(size_t)i * n + i];                                                           \
      }                                                                                            \
      if (ldlt)                                                                                    \
        for (int i = 0; i < n; ++i) x[i] /= D[i];                                                  \
      for (int i = n - 1; i >= 0; --i) {                                                           \
        VAL_T s = x[i];                                                                            \
        for (int j = i + 1; j < n; ++j) s -= L[(size_t)j * n + i] * x[j];                          \
        x[i] = s / L[(size_t)i * n + i];                                                           \
      }                                                                                            \
    }                                                                                              \
    free(L); free(D); free(y);                                                                     \
    return rc;                                                                                     \
  }
QDO_DIRECT(qdo_direct_f32, float, sqrtf)
QDO_DIRECT(qdo_direct_f64, double, sqrt)

/* ------------------------------------------------------------------------- */
/* LU solve (solver_type "LU").  The reference's solver is    */
/* Eigen::SparseLU<SparseMatrix<dt>, {AMD,COLAMD}Ordering<int>>               */
/* (quadrants/program/sparse_solver.h:15-20): a supernodal left-looking LU     */
/* with PARTIAL (row) PIVOTING.  Restated here as Gaussian elimination with    */
/* partial pivoting on the dense expansion of the CSR matrix (duplicates       */
/* summed) -- the same factorisation P A = L U up to the column ordering,      */
/* which changes rounding, not the solution; the GPU path (no pivoting, static */
/* ordering) and this one are both compared with the reference's solutions     */
/* with a tolerance.  Returns 0, or 1 + k when no non-zero pivot exists in     */
/* column k.  Dense O(n^2) workspace: small systems only.                      */
/* ------------------------------------------------------------------------- */
#define QDO_LU(NAME, VAL_T, FABS)                                                                  \
  QDO_API int NAME(int n, const int32_t *outer, const int32_t *inner, const VAL_T *vals,           \
                   const VAL_T *b, VAL_T *x) {                                                     \
    VAL_T *A = (VAL_T *)calloc((size_t)n * n, sizeof(VAL_T));                                      \
    int rc = 0;                                                                                    \
    for (int r = 0; r < n; ++r)                                                                    \
      for (int32_t p = outer[r]; p < outer[r + 1]; ++p) A[(size_t)r * n + inner[p]] += vals[p];    \
    for (int i = 0; i < n; ++i) x[i] = b[i];                                                       \
    for (int k = 0; k < n && rc == 0; ++k) {                                                       \
      int piv = k;                                                                                 \
      for (int r = k + 1; r < n; ++r)                                                              \
        if (FABS(A[(size_t)r * n + k]) > FABS(A[(size_t)piv * n + k])) piv = r;                    \
      if (A[(size_t)piv * n + k] == (VAL_T)0) { rc = 1 + k; break; }                               \
      if (piv != k) {                                                                              \
        for (int c = 0; c < n; ++c) {                                                              \
          const VAL_T t = A[(size_t)k * n + c];                                                    \
          A[(size_t)k * n + c] = A[(size_t)piv * n + c];                                           \
          A[(size_t)piv * n + c] = t;                                                              \
        }                                                                                          \
        const VAL_T t = x[k]; x[k] = x[piv]; x[piv] = t;                                           \
      }                                                                                            \
      for (int r = k + 1; r < n; ++r) {                                                            \
        const VAL_T l = A[(size_t)r * n + k] / A[(size_t)k * n + k];                               \
        if (l == (VAL_T)0) continue;                                                               \
        for (int c = k + 1; c < n; ++c) A[(size_t)r * n + c] -= l * A[(size_t)k * n + c];          \
        x[r] -= l * x[k];                                                                          \
      }                                                                                            \
    }                                                                                              \
    if (rc == 0)                                                                                   \
      for (int i = n - 1; i >= 0; --i) {                                                           \
        VAL_T s = x[i];                                                                            \
        for (int c = i + 1; c < n; ++c) s -= A[(size_t)i * n + c] * x[c];                          \
        x[i] = s / A[(size_t)i * n + i];                                                           \
      }                                                                                            \
    free(A);                                                                                       \
    return rc;                                                                                     \
  }
QDO_LU(qdo_lu_f32, float, fabsf)
QDO_LU(qdo_lu_f64, double, fabs)
