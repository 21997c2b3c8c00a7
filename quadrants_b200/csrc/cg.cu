// cg.cu -- fused conjugate gradient on the device.
//
// Semantics (mode QDSP_CG_EIGEN_JACOBI) = CG<EigenT,DT>::solve (quadrants/program/conjugate_gradient.h:36-49), i.e.
// Eigen::internal::conjugate_gradient with DiagonalPreconditioner
// (external/eigen/Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:29-85, BasicPreconditioners.h:62-86):
//   x = 0; r = b - A x; thr = max(tol^2 |b|^2, min); if |r|^2 < thr stop; p = D^-1 r; absNew = r.p
//   loop i < maxIters { tmp = A p; alpha = absNew / p.tmp; x += alpha p; r -= alpha tmp; if |r|^2 < thr break;
//                       z = D^-1 r; absOld = absNew; absNew = r.z; beta = absNew/absOld; p = z + beta p; ++i }
//   error = sqrt(|r|^2 / |b|^2); success iff error <= tol.
// Mode QDSP_CG_CUCG = CUCG::solve (quadrants/program/conjugate_gradient.cpp:19-83): the same recurrence with D = I,
// started from the x passed in, stopping when r.r <= tol^2 (absolute).  The reference runs it as 1 cuSPARSE SpMV
// (+ handle/malloc/free), Sscal + 3 Saxpy, 2 blocking Sdot per iteration (~15 vector passes, >= 3 host syncs).
//
// Here one iteration is THREE kernels and no host round trip; every scalar stays on the device:
//   A  cg_spmv_dot : tmp = A p fused with the partial sums of p.tmp            (matrix + 2n)
//   B  cg_update   : x += a p, r -= a tmp, |r|^2 and r.(D^-1 r) in one pass     (5n read, 2n write)
//   C  cg_direction: p = D^-1 r + beta p                                        (3n read, 1n write)
// => nnz*(sv+4) + (n+1)*4 + 13*n*sv bytes per iteration, the model of SURVEY 8(d).
// Dot products: per-thread partials in T, block tree in double, per-block partial to memory, and the LAST block to
// finish (atomic ticket) adds the partials in a fixed order -> deterministic, no atomics on the sums, no extra launch.
// That block also advances the scalar recurrence (alpha, beta, convergence flag, iteration counter).  Once the flag
// is up every later kernel exits immediately, so the host may enqueue iterations ahead of the convergence test and
// only polls the flag every `check_every` iterations.  Iterations are replayed from a CUDA graph.
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <vector>

#include "host.hpp"
#include "spmv_kernels.cuh"

namespace qdsp {

constexpr int CG_THREADS = 256;

struct CgScalars {
  double pAp, abs_new, abs_old, res2, rhs2, thr;
  double alpha, beta;
  int iters;      // Eigen's i: completed loop bodies
  int done;       // 0 running, 1 converged, 2 budget exhausted
  int max_iters;
  int stop_le;    // 1: stop when res2 <= thr (CUCG), 0: when res2 < thr (Eigen)
  unsigned ticket[4];
};

// ---- block reduction helpers ----------------------------------------------------------------------------------------
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double *s_red /* NV * 8 */) {
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    v[j] = warp_sum(v[j]);
    if (lane == 0) s_red[j * 8 + warp] = v[j];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < CG_THREADS / 32; ++w) t += s_red[j * 8 + w];
    v[j] = t;  // every thread holds the block total
  }
  __syncthreads();
}

// Publishes this block's NV partial sums; returns true in the last block to finish, with v[] = grand totals summed in a
// fixed order (thread-strided, then the block tree above).
template <int NV>
__device__ __forceinline__ bool grid_sum_last_block(double (&v)[NV], double *partials /* NV x gridDim.x */,
                                                    unsigned *ticket, double *s_red, unsigned *s_flag) {
  block_sum<NV>(v, s_red);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int j = 0; j < NV; ++j) partials[static_cast<size_t>(j) * gridDim.x + blockIdx.x] = v[j];
    __threadfence();
    const unsigned t = atomicAdd(ticket, 1u);
    *s_flag = (t == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (*s_flag == 0u) return false;
  __threadfence();
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    double t = 0.0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += CG_THREADS)
      t += __ldcg(&partials[static_cast<size_t>(j) * gridDim.x + i]);
    v[j] = t;
  }
  block_sum<NV>(v, s_red);
  if (threadIdx.x == 0) *ticket = 0u;  // ready for the next launch
  return true;
}

// ---- setup kernels ----------------------------------------------------------------------------------------------------
// DiagonalPreconditioner::factorize: 1/a_jj, or 1 when the diagonal is missing or zero.  identity != 0 -> all ones.
template <typename T>
__global__ void __launch_bounds__(CG_THREADS)
cg_invdiag_kernel(int n, const int *__restrict__ indptr, const int *__restrict__ indices,
                  const T *__restrict__ values, int identity, T *__restrict__ dinv) {
  const int j = blockIdx.x * CG_THREADS + threadIdx.x;
  if (j >= n) return;
  T d = T(1);
  if (!identity) {
    for (int k = indptr[j]; k < indptr[j + 1]; ++k)
      if (indices[k] == j) {
        const T v = values[k];
        if (v != T(0)) d = T(1) / v;
        break;
      }
  }
  dinv[j] = d;
}

// r = b - tmp (tmp = A x0); p = D^-1 r; rhs2 = b.b, res2 = r.r, abs_new = r.p; thresholds and early exits.
template <typename T>
__global__ void __launch_bounds__(CG_THREADS)
cg_init_kernel(int n, const T *__restrict__ b, const T *__restrict__ tmp, const T *__restrict__ dinv,
               T *__restrict__ r, T *__restrict__ p, double tol, int relative, CgScalars *__restrict__ sc,
               double *__restrict__ partials) {
  __shared__ double s_red[3 * 8];
  __shared__ unsigned s_flag;
  T s0 = T(0), s1 = T(0), s2 = T(0);
  for (int i = blockIdx.x * CG_THREADS + threadIdx.x; i < n; i += gridDim.x * CG_THREADS) {
    const T bi = b[i];
    const T ri = bi - tmp[i];
    const T pi = dinv[i] * ri;
    r[i] = ri;
    p[i] = pi;
    s0 = fma_t(bi, bi, s0);
    s1 = fma_t(ri, ri, s1);
    s2 = fma_t(ri, pi, s2);
  }
  double v[3] = {static_cast<double>(s0), static_cast<double>(s1), static_cast<double>(s2)};
  if (grid_sum_last_block<3>(v, partials, &sc->ticket[0], s_red, &s_flag) && threadIdx.x == 0) {
    // scalars are kept in double but rounded through T where Eigen's RealScalar would be T
    const double rhs2 = static_cast<double>(static_cast<T>(v[0]));
    const double res2 = static_cast<double>(static_cast<T>(v[1]));
    sc->rhs2 = rhs2;
    sc->res2 = res2;
    sc->abs_new = static_cast<double>(static_cast<T>(v[2]));
    sc->abs_old = 0.0;
    sc->iters = 0;
    sc->alpha = 0.0;
    sc->beta = 0.0;
    const double tiny = sizeof(T) == 4 ? static_cast<double>(FLT_MIN) : DBL_MIN;
    double thr = relative ? static_cast<double>(static_cast<T>(static_cast<T>(tol) * static_cast<T>(tol) * static_cast<T>(rhs2)))
                          : static_cast<double>(static_cast<T>(tol) * static_cast<T>(tol));
    if (relative && thr < tiny) thr = tiny;
    if (tol < 0.0) thr = -1.0;  // fixed-iteration timing runs: never converge
    sc->thr = thr;
    int done = 0;
    if (tol >= 0.0 && relative && rhs2 == 0.0) done = 1;
    else if (sc->stop_le ? (res2 <= thr) : (res2 < thr)) done = 1;
    else if (sc->max_iters <= 0) done = 2;
    sc->done = done;
  }
}

// ---- iteration kernels --------------------------------------------------------------------------------------------------
template <typename T, bool ROW_MAJOR_ORDER>
__global__ void __launch_bounds__(SPMV_ROWS)
cg_spmv_dot_kernel(int rows, const int *__restrict__ indptr, const int *__restrict__ indices,
                   const T *__restrict__ values, const T *__restrict__ p, T *__restrict__ tmp,
                   CgScalars *__restrict__ sc, double *__restrict__ partials) {
  __shared__ SpmvSmem<T> S;
  __shared__ unsigned s_flag;
  if (sc->done) return;
  const T yi = spmv_cta<T, ROW_MAJOR_ORDER>(S, rows, indptr, indices, values, p, tmp);
  const int row = blockIdx.x * SPMV_ROWS + threadIdx.x;
  const T pi = row < rows ? p[row] : T(0);
  __syncthreads();  // S.val may still be read by the vector path
  double v[1] = {static_cast<double>(pi * yi)};
  if (grid_sum_last_block<1>(v, partials, &sc->ticket[1], S.red, &s_flag) && threadIdx.x == 0) {
    const T pap = static_cast<T>(v[0]);
    sc->pAp = static_cast<double>(pap);
    sc->alpha = static_cast<double>(static_cast<T>(sc->abs_new) / pap);
  }
}

template <typename T>
__global__ void __launch_bounds__(CG_THREADS)
cg_update_kernel(int n, const T *__restrict__ p, const T *__restrict__ tmp, const T *__restrict__ dinv,
                 T *__restrict__ x, T *__restrict__ r, CgScalars *__restrict__ sc, double *__restrict__ partials) {
  __shared__ double s_red[2 * 8];
  __shared__ unsigned s_flag;
  if (sc->done) return;
  const T alpha = static_cast<T>(sc->alpha);
  T s_rr = T(0), s_rz = T(0);
  for (int i = blockIdx.x * CG_THREADS + threadIdx.x; i < n; i += gridDim.x * CG_THREADS) {
    x[i] = fma_t(alpha, p[i], x[i]);
    const T ri = fma_t(-alpha, tmp[i], r[i]);
    r[i] = ri;
    const T zi = dinv[i] * ri;
    s_rr = fma_t(ri, ri, s_rr);
    s_rz = fma_t(ri, zi, s_rz);
  }
  double v[2] = {static_cast<double>(s_rr), static_cast<double>(s_rz)};
  if (grid_sum_last_block<2>(v, partials, &sc->ticket[2], s_red, &s_flag) && threadIdx.x == 0) {
    const double res2 = static_cast<double>(static_cast<T>(v[0]));
    sc->res2 = res2;
    if (sc->stop_le ? (res2 <= sc->thr) : (res2 < sc->thr)) {
      sc->done = 1;  // Eigen breaks before ++i
      if (sc->stop_le) sc->iters += 1;  // CUCG counts the body it just finished (k++)
    } else {
      const T abs_old = static_cast<T>(sc->abs_new);
      const T abs_new = static_cast<T>(v[1]);
      sc->abs_old = static_cast<double>(abs_old);
      sc->abs_new = static_cast<double>(abs_new);
      sc->beta = static_cast<double>(abs_new / abs_old);
      sc->iters += 1;
      if (sc->iters >= sc->max_iters) sc->done = 2;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(CG_THREADS)
cg_direction_kernel(int n, const T *__restrict__ r, const T *__restrict__ dinv, T *__restrict__ p,
                    const CgScalars *__restrict__ sc) {
  if (sc->done == 1) return;  // converged: p is dead.  (done == 2: Eigen still updates p before leaving the loop)
  const T beta = static_cast<T>(sc->beta);
  for (int i = blockIdx.x * CG_THREADS + threadIdx.x; i < n; i += gridDim.x * CG_THREADS)
    p[i] = fma_t(beta, p[i], dinv[i] * r[i]);
}

// ---- host ------------------------------------------------------------------------------------------------------------------
struct CgSolver::Impl {
  const CsrMatrix &A;
  int max_iters;
  double tol;
  int mode;
  bool verbose;
  DeviceBuffer r, p, tmp, dinv, partials, scalars;
  CgScalars *h_sc = nullptr;  // pinned mirror
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_in = nullptr, ev_out = nullptr;
  cudaGraphExec_t graph = nullptr;
  void *graph_x = nullptr;
  int graph_iters = 0;
  int vec_grid = 0, spmv_grid = 0;

  Impl(const CsrMatrix &A_, int mi, double t, int m, bool v) : A(A_), max_iters(mi), tol(t), mode(m), verbose(v) {}
};

CgSolver::CgSolver(const CsrMatrix &A, int max_iters, double tol, int mode, bool verbose, DistPlan *dist)
    : impl_(new Impl(A, max_iters, tol, mode, verbose)) {
  (void)dist;
  QDSP_REQUIRE(A.rows == A.cols, QDSP_ERR_INVALID, "CG needs a square matrix");
  QDSP_REQUIRE(mode == QDSP_CG_EIGEN_JACOBI || mode == QDSP_CG_CUCG, QDSP_ERR_INVALID, "CG: unknown mode");
  Impl &I = *impl_;
  const size_t n = static_cast<size_t>(A.rows), sv = A.value_bytes();
  I.r.alloc(n * sv); I.p.alloc(n * sv); I.tmp.alloc(n * sv); I.dinv.alloc(n * sv);
  I.spmv_grid = ceil_div(A.rows, SPMV_ROWS);
  I.vec_grid = std::max(1, std::min(ceil_div(A.rows, CG_THREADS), device_props().sm_count * 8));
  I.partials.alloc(sizeof(double) * 3 * static_cast<size_t>(std::max(I.spmv_grid, I.vec_grid)));
  I.scalars.alloc(sizeof(CgScalars));
  QDSP_CUDA(cudaMemset(I.scalars.get(), 0, sizeof(CgScalars)));
  QDSP_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&I.h_sc), sizeof(CgScalars), cudaHostAllocDefault));
  QDSP_CUDA(cudaStreamCreateWithFlags(&I.stream, cudaStreamNonBlocking));
  QDSP_CUDA(cudaEventCreateWithFlags(&I.ev_in, cudaEventDisableTiming));
  QDSP_CUDA(cudaEventCreateWithFlags(&I.ev_out, cudaEventDisableTiming));
  const int identity = mode == QDSP_CG_CUCG;
  const unsigned g = static_cast<unsigned>(ceil_div(A.rows, CG_THREADS));
  if (A.dtype == QDSP_F32)
    cg_invdiag_kernel<float><<<g, CG_THREADS, 0, I.stream>>>(A.rows, A.indptr.get<int>(), A.indices.get<int>(), A.values.get<float>(), identity, I.dinv.get<float>());
  else
    cg_invdiag_kernel<double><<<g, CG_THREADS, 0, I.stream>>>(A.rows, A.indptr.get<int>(), A.indices.get<int>(), A.values.get<double>(), identity, I.dinv.get<double>());
  QDSP_LAUNCH_CHECK();
  QDSP_CUDA(cudaStreamSynchronize(I.stream));
}

CgSolver::~CgSolver() {
  if (!impl_) return;
  Impl &I = *impl_;
  if (I.graph) cudaGraphExecDestroy(I.graph);
  if (I.ev_in) cudaEventDestroy(I.ev_in);
  if (I.ev_out) cudaEventDestroy(I.ev_out);
  if (I.stream) cudaStreamDestroy(I.stream);
  if (I.h_sc) cudaFreeHost(I.h_sc);
}

namespace {

template <typename T>
void launch_iteration(CgSolver::Impl &I, T *x) {
  const CsrMatrix &A = I.A;
  auto *sc = I.scalars.get<CgScalars>();
  double *part = I.partials.get<double>();
  const int n = A.rows;
  if (A.row_major_order)
    cg_spmv_dot_kernel<T, true><<<I.spmv_grid, SPMV_ROWS, 0, I.stream>>>(n, A.indptr.get<int>(), A.indices.get<int>(), A.values.get<T>(), I.p.get<T>(), I.tmp.get<T>(), sc, part);
  else
    cg_spmv_dot_kernel<T, false><<<I.spmv_grid, SPMV_ROWS, 0, I.stream>>>(n, A.indptr.get<int>(), A.indices.get<int>(), A.values.get<T>(), I.p.get<T>(), I.tmp.get<T>(), sc, part);
  QDSP_LAUNCH_CHECK();
  cg_update_kernel<T><<<I.vec_grid, CG_THREADS, 0, I.stream>>>(n, I.p.get<T>(), I.tmp.get<T>(), I.dinv.get<T>(), x, I.r.get<T>(), sc, part);
  QDSP_LAUNCH_CHECK();
  cg_direction_kernel<T><<<I.vec_grid, CG_THREADS, 0, I.stream>>>(n, I.r.get<T>(), I.dinv.get<T>(), I.p.get<T>(), sc);
  QDSP_LAUNCH_CHECK();
}

// (Re)captures `iters` iterations on x into an executable graph.
template <typename T>
void ensure_graph(CgSolver::Impl &I, T *x, int iters) {
  if (I.graph && I.graph_x == x && I.graph_iters == iters) return;
  if (I.graph) { cudaGraphExecDestroy(I.graph); I.graph = nullptr; }
  cudaGraph_t g = nullptr;
  const int64_t before = g_launch_count.load();
  QDSP_CUDA(cudaStreamBeginCapture(I.stream, cudaStreamCaptureModeThreadLocal));
  try {
    for (int k = 0; k < iters; ++k) launch_iteration<T>(I, x);
  } catch (...) {
    cudaStreamEndCapture(I.stream, &g);
    if (g) cudaGraphDestroy(g);
    throw;
  }
  QDSP_CUDA(cudaStreamEndCapture(I.stream, &g));
  g_launch_count.store(before);  // captured, not launched: replays are counted where the graph is launched
  QDSP_CUDA(cudaGraphInstantiate(&I.graph, g, 0));
  QDSP_CUDA(cudaGraphDestroy(g));
  I.graph_x = x;
  I.graph_iters = iters;
}

template <typename T>
void start(CgSolver::Impl &I, T *x, const T *b, int max_iters, double tol, bool never_converge) {
  const CsrMatrix &A = I.A;
  const int n = A.rows;
  auto *sc = I.scalars.get<CgScalars>();
  CgScalars init{};
  init.max_iters = max_iters;
  init.stop_le = (I.mode == QDSP_CG_CUCG) ? 1 : 0;
  QDSP_CUDA(cudaMemcpyAsync(sc, &init, sizeof(init), cudaMemcpyHostToDevice, I.stream));
  if (I.mode == QDSP_CG_EIGEN_JACOBI)  // IterativeSolverBase::_solve_impl: x.setZero()
    QDSP_CUDA(cudaMemsetAsync(x, 0, sizeof(T) * static_cast<size_t>(n), I.stream));
  cudaStream_t s = I.stream;
  // tmp = A x0 through the plain SpMV path on the solver's stream
  A.spmv(x, I.tmp.get(), s);
  const int relative = I.mode == QDSP_CG_EIGEN_JACOBI;
  cg_init_kernel<T><<<I.vec_grid, CG_THREADS, 0, s>>>(n, b, I.tmp.get<T>(), I.dinv.get<T>(), I.r.get<T>(), I.p.get<T>(),
                                                      never_converge ? -1.0 : tol, relative, sc,
                                                      I.partials.get<double>());
  QDSP_LAUNCH_CHECK();
}

template <typename T>
void solve_t(CgSolver::Impl &I, void *xv, const void *bv, cudaStream_t user, int *iterations, double *error, int *success) {
  T *x = static_cast<T *>(xv);
  const T *b = static_cast<const T *>(bv);
  QDSP_CUDA(cudaEventRecord(I.ev_in, user));
  QDSP_CUDA(cudaStreamWaitEvent(I.stream, I.ev_in, 0));
  start<T>(I, x, b, I.max_iters, I.tol, false);
  // poll the device-side flag every `batch` iterations; kernels after convergence are no-ops
  const int batch = std::max(1, std::min(I.max_iters, I.A.rows >= (1 << 20) ? 16 : 64));
  int launched = 0;
  CgScalars &h = *I.h_sc;
  while (true) {
    QDSP_CUDA(cudaMemcpyAsync(&h, I.scalars.get(), sizeof(CgScalars), cudaMemcpyDeviceToHost, I.stream));
    QDSP_CUDA(cudaStreamSynchronize(I.stream));
    if (I.verbose && launched) std::printf("iter: %d, r1: %g\n", h.iters, h.res2);
    if (h.done || launched >= I.max_iters) break;
    ensure_graph<T>(I, x, batch);
    QDSP_CUDA(cudaGraphLaunch(I.graph, I.stream));
    count_launch(3 * batch);
    launched += batch;
  }
  QDSP_CUDA(cudaEventRecord(I.ev_out, I.stream));
  QDSP_CUDA(cudaStreamWaitEvent(user, I.ev_out, 0));
  if (iterations) *iterations = h.iters;
  if (I.mode == QDSP_CG_EIGEN_JACOBI) {
    const double err = h.rhs2 == 0.0 ? 0.0 : std::sqrt(static_cast<double>(static_cast<T>(h.res2 / h.rhs2)));
    if (error) *error = err;
    if (success) *success = err <= static_cast<double>(static_cast<T>(I.tol)) ? 1 : 0;
    if (I.verbose) {
      std::printf("#iterations:     %d\n", h.iters);
      std::printf("estimated error: %g\n", err);
    }
  } else {
    if (error) *error = h.res2;
    if (success) *success = h.res2 <= h.thr ? 1 : 0;
  }
}

}  // namespace

void CgSolver::solve(void *x, const void *b, cudaStream_t stream, int *iterations, double *error, int *success) {
  QDSP_REQUIRE(x != nullptr && b != nullptr, QDSP_ERR_INVALID, "CG: null vector");
  if (impl_->A.dtype == QDSP_F32) solve_t<float>(*impl_, x, b, stream, iterations, error, success);
  else solve_t<double>(*impl_, x, b, stream, iterations, error, success);
}

void CgSolver::run_fixed(void *x, const void *b, int iters, cudaStream_t stream) {
  QDSP_REQUIRE(x != nullptr && b != nullptr && iters >= 0, QDSP_ERR_INVALID, "CG: bad arguments");
  Impl &I = *impl_;
  QDSP_CUDA(cudaEventRecord(I.ev_in, stream));
  QDSP_CUDA(cudaStreamWaitEvent(I.stream, I.ev_in, 0));
  const int batch = std::max(1, std::min(iters, 16));
  if (I.A.dtype == QDSP_F32) {
    start<float>(I, static_cast<float *>(x), static_cast<const float *>(b), iters + 1, 0.0, true);
    ensure_graph<float>(I, static_cast<float *>(x), batch);
  } else {
    start<double>(I, static_cast<double *>(x), static_cast<const double *>(b), iters + 1, 0.0, true);
    ensure_graph<double>(I, static_cast<double *>(x), batch);
  }
  int left = iters;
  while (left >= batch) {
    QDSP_CUDA(cudaGraphLaunch(I.graph, I.stream));
    count_launch(3 * batch);
    left -= batch;
  }
  for (; left > 0; --left) {
    if (I.A.dtype == QDSP_F32) launch_iteration<float>(I, static_cast<float *>(x));
    else launch_iteration<double>(I, static_cast<double *>(x));
  }
  QDSP_CUDA(cudaEventRecord(I.ev_out, I.stream));
  QDSP_CUDA(cudaStreamWaitEvent(stream, I.ev_out, 0));
}

}  // namespace qdsp
