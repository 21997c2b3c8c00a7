// solver.cu -- sparse direct solver on the device: LLT / LDLT / LU behind qd.linalg.SparseSolver.
//
// Replaces CuSparseSolver (quadrants/program/sparse_solver.cpp:149-476).  The reference's arch=cuda path is: D2H the CSR,
// csrsymamdHost + csrpermHost on the host (:169-231), then cuSOLVER's low-level csrchol{Analysis,Factor,Solve} on the
// device for LLT/LDLT (:245-265, :302-331, :379-430) -- and for LU everything on the HOST (csrlu*Host, :267-288,
// :333-364, :432-476), fp32 only.  Here:
//
//   analyze_pattern  host: approximate-minimum-degree ordering of A + A^T, elimination tree, factor pattern in CSC with a
//                    row index on top (SymbolicFactor); device: the pattern and the A -> factor slot map.
//   factorize        device, ONE launch: left-looking, one warp per column, columns claimed in ascending order from an
//                    atomic counter.  A column waits (ld.acquire on a per-column flag) for each column it depends on in
//                    ascending order, subtracts that column's contribution, then scales and publishes itself
//                    (st.release).  Ascending claims make the wait deadlock-free: the lowest unfinished column never
//                    waits, and only running warps hold claims.  The dependency order is fixed, so the factor is
//                    bit-reproducible run to run (no atomics on values).
//                    Where the tail of a dependency column has the same pattern as the target column (chains of the
//                    elimination tree / supernodes -- most of the arithmetic), slots map 1:1 and the update is a
//                    coalesced axpy; otherwise a binary search in the target's sorted rows.
//                    LU: STATIC pivoting.  The reference's Eigen SparseLU and csrluHost exchange rows while they
//                    factorise (threshold 1.0, sparse_solver.cpp:343-359); a left-looking kernel over a fixed pattern
//                    cannot, so the exchange is decided up front: analyze_pattern() reads the values it is given and
//                    picks a row permutation Q that puts a large entry of every row on the diagonal (solver_lu_row_matching_host
//                    below: the maximum-product matching of MC64 by shortest augmenting paths, with a
//                    preference for the diagonal; a diagonal entry that is its row's largest simply stays), and everything after that works on
//                    Q A.  L unit-lower by columns, U by rows in the same slots.  A pivot that is zero or vanishes
//                    against column j of L below it and of U above it (|d| < 16 eps max|column|) is reported
//                    through info() instead of being divided by.  This is weaker than partial pivoting -- the choice
//                    looks at A, not at the partly eliminated matrix -- and is stated as such in DESIGN.md 3.9.
//   solve            device, two launches: synchronisation-free triangular solves, one warp per row, the same flag
//                    protocol (forward ascending, backward descending), permutation folded into the first load and the
//                    last store.  Fixed summation order.
// Bound: latency of the dependency chain (elimination-tree height x one flag round trip), not HBM; the factor lives in L2
// for every size a direct solver is used at.  No cuSOLVER / cuSPARSE.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <numeric>

#include "solver.hpp"

namespace qdsp {

// =====================================================================================================================
// host analysis
// =====================================================================================================================
namespace {

// adjacency of A + A^T without the diagonal, sorted, duplicates removed
void symmetric_adjacency(int n, const int32_t *ip, const int32_t *ix, std::vector<int32_t> &ptr, std::vector<int32_t> &adj) {
  std::vector<int64_t> cnt(n + 1, 0);
  for (int r = 0; r < n; ++r)
    for (int q = ip[r]; q < ip[r + 1]; ++q) {
      const int c = ix[q];
      QDSP_REQUIRE(c >= 0 && c < n, QDSP_ERR_INVALID, "solver: column index out of range");
      if (c != r) { ++cnt[r + 1]; ++cnt[c + 1]; }
    }
  for (int i = 0; i < n; ++i) cnt[i + 1] += cnt[i];
  QDSP_REQUIRE(cnt[n] < INT_MAX, QDSP_ERR_INVALID, "solver: pattern too large");
  std::vector<int32_t> tmp(static_cast<size_t>(cnt[n]));
  std::vector<int64_t> next(cnt.begin(), cnt.end() - 1);
  for (int r = 0; r < n; ++r)
    for (int q = ip[r]; q < ip[r + 1]; ++q) {
      const int c = ix[q];
      if (c != r) { tmp[next[r]++] = c; tmp[next[c]++] = r; }
    }
  ptr.assign(n + 1, 0);
  adj.clear();
  adj.reserve(tmp.size());
  for (int i = 0; i < n; ++i) {
    auto b = tmp.begin() + cnt[i], e = tmp.begin() + cnt[i + 1];
    std::sort(b, e);
    e = std::unique(b, e);
    adj.insert(adj.end(), b, e);
    ptr[i + 1] = static_cast<int32_t>(adj.size());
  }
}

// degree buckets: doubly linked lists
struct Buckets {
  std::vector<int32_t> head, next, prev;
  explicit Buckets(int n) : head(n + 1, -1), next(n, -1), prev(n, -1) {}
  void insert(int i, int d) {
    next[i] = head[d];
    prev[i] = -1;
    if (head[d] >= 0) prev[head[d]] = i;
    head[d] = i;
  }
  void remove(int i, int d) {
    if (prev[i] >= 0) next[prev[i]] = next[i];
    else head[d] = next[i];
    if (next[i] >= 0) prev[next[i]] = prev[i];
  }
};

}  // namespace

void solver_order_host(int n, const int32_t *indptr, const int32_t *indices, std::vector<int32_t> &perm) {
  perm.assign(n, 0);
  if (n == 0) return;
  std::vector<int32_t> aptr, aadj;
  symmetric_adjacency(n, indptr, indices, aptr, aadj);
  // quotient graph: a live variable i has variable neighbours adjv[i] and element neighbours adje[i]; an element e (named
  // by its pivot) has the variable list elem[e].  state: 0 variable, 1 live element, 2 absorbed element.
  std::vector<std::vector<int32_t>> adjv(n), adje(n), elem(n);
  std::vector<int32_t> deg(n), mark(n, -1);
  std::vector<uint8_t> state(n, 0);
  std::vector<int64_t> w(n, 0);
  int64_t wflg = 1;
  Buckets bk(n);
  for (int i = 0; i < n; ++i) {
    adjv[i].assign(aadj.begin() + aptr[i], aadj.begin() + aptr[i + 1]);
    deg[i] = static_cast<int32_t>(adjv[i].size());
  }
  for (int i = n - 1; i >= 0; --i) bk.insert(i, deg[i]);  // ties: lowest index first
  std::vector<int32_t>().swap(aadj);
  int mindeg = 0;
  std::vector<int32_t> lp, keep;
  for (int step = 0; step < n; ++step) {
    while (bk.head[mindeg] < 0) ++mindeg;
    const int p = bk.head[mindeg];
    bk.remove(p, mindeg);
    perm[step] = p;
    // the new element: L_p = (A_p u U_{e in E_p} L_e) \ {p}
    lp.clear();
    mark[p] = step;
    for (int v : adjv[p])
      if (state[v] == 0 && mark[v] != step) { mark[v] = step; lp.push_back(v); }
    for (int e : adje[p])
      if (state[e] == 1) {
        for (int v : elem[e])
          if (state[v] == 0 && mark[v] != step) { mark[v] = step; lp.push_back(v); }
        state[e] = 2;
        std::vector<int32_t>().swap(elem[e]);
      }
    state[p] = 1;
    std::vector<int32_t>().swap(adjv[p]);
    std::vector<int32_t>().swap(adje[p]);
    // w[e] - wflg = |L_e \ L_p| for every live element e that touches L_p
    for (int i : lp)
      for (int e : adje[i])
        if (state[e] == 1) {
          if (w[e] < wflg) w[e] = wflg + static_cast<int64_t>(elem[e].size());
          --w[e];
        }
    const int lpn = static_cast<int>(lp.size());
    for (int i : lp) {
      int64_t sum = 0;
      keep.clear();
      for (int e : adje[i])
        if (state[e] == 1) {
          const int64_t ext = w[e] - wflg;
          if (ext > 0) { sum += ext; keep.push_back(e); }
          else { state[e] = 2; std::vector<int32_t>().swap(elem[e]); }  // aggressive absorption: L_e is inside L_p
        }
      keep.push_back(p);
      adje[i].assign(keep.begin(), keep.end());
      keep.clear();
      for (int v : adjv[i])
        if (state[v] == 0 && mark[v] != step) keep.push_back(v);
      adjv[i].assign(keep.begin(), keep.end());
      const int64_t d = std::min<int64_t>(std::min<int64_t>(n - step - 1, static_cast<int64_t>(deg[i]) + lpn - 1),
                                          static_cast<int64_t>(adjv[i].size()) + lpn - 1 + sum);
      bk.remove(i, deg[i]);
      deg[i] = static_cast<int32_t>(std::max<int64_t>(d, 0));
      bk.insert(i, deg[i]);
      if (deg[i] < mindeg) mindeg = deg[i];
    }
    elem[p] = lp;
    wflg += n + 1;
  }
}

void solver_symbolic_host(int n, const int32_t *indptr, const int32_t *indices, const std::vector<int32_t> &perm,
                          SymbolicFactor &out) {
  out = SymbolicFactor();
  out.n = n;
  out.perm = perm;
  out.pinv.assign(n, -1);
  for (int i = 0; i < n; ++i) {
    QDSP_REQUIRE(perm[i] >= 0 && perm[i] < n && out.pinv[perm[i]] < 0, QDSP_ERR_INVALID, "solver: not a permutation");
    out.pinv[perm[i]] = i;
  }
  // strictly-lower adjacency of the permuted symmetric pattern: low[i] = {k < i : B(i,k) != 0}
  std::vector<int32_t> aptr, aadj;
  symmetric_adjacency(n, indptr, indices, aptr, aadj);
  std::vector<int32_t> lptr(n + 1, 0), low;
  for (int i = 0; i < n; ++i) {
    const int o = perm[i];
    int c = 0;
    for (int q = aptr[o]; q < aptr[o + 1]; ++q) c += out.pinv[aadj[q]] < i;
    lptr[i + 1] = lptr[i] + c;
  }
  low.resize(lptr[n]);
  for (int i = 0; i < n; ++i) {
    const int o = perm[i];
    int at = lptr[i];
    for (int q = aptr[o]; q < aptr[o + 1]; ++q) {
      const int k = out.pinv[aadj[q]];
      if (k < i) low[at++] = k;
    }
  }
  // elimination tree (Liu, path compression)
  out.parent.assign(n, -1);
  {
    std::vector<int32_t> anc(n, -1);
    for (int i = 0; i < n; ++i)
      for (int q = lptr[i]; q < lptr[i + 1]; ++q) {
        int r = low[q];
        while (r != -1 && r < i) {
          const int nx = anc[r];
          anc[r] = i;
          if (nx == -1) out.parent[r] = i;
          r = nx;
        }
      }
  }
  // row patterns by walking the tree from each B(i,k) up to the first marked node (row subtree); the walk yields a row in
  // no particular order, which is all the column fill below needs
  std::vector<int32_t> mark(n, -1), colcount(n, 1);
  std::vector<int32_t> walk_ptr(n + 1, 0), walk;
  int64_t total = n;
  for (int i = 0; i < n; ++i) {
    mark[i] = i;
    for (int q = lptr[i]; q < lptr[i + 1]; ++q)
      for (int r = low[q]; mark[r] != i; r = out.parent[r]) {
        mark[r] = i;
        walk.push_back(r);
        ++colcount[r];
      }
    total = n + static_cast<int64_t>(walk.size());
    QDSP_REQUIRE(total < INT_MAX, QDSP_ERR_INVALID, "solver: the factor has more than 2^31 entries");
    walk_ptr[i + 1] = static_cast<int32_t>(walk.size());
  }
  out.factor_nnz = total;
  out.colptr.assign(n + 1, 0);
  for (int j = 0; j < n; ++j) out.colptr[j + 1] = out.colptr[j] + colcount[j];
  out.rowidx.assign(static_cast<size_t>(total), 0);
  std::vector<int32_t> next(n);
  for (int j = 0; j < n; ++j) {
    out.rowidx[out.colptr[j]] = j;  // the diagonal leads its column
    next[j] = out.colptr[j] + 1;
    out.flops += static_cast<double>(colcount[j]) * colcount[j];
  }
  // columns: rows arrive in ascending order because the rows are visited in ascending order
  for (int i = 0; i < n; ++i)
    for (int q = walk_ptr[i]; q < walk_ptr[i + 1]; ++q) out.rowidx[next[walk[q]]++] = i;
  // rows, sorted by column without a sort: transpose the columns (visited in ascending order) back into the rows
  out.rowptr = walk_ptr;
  std::vector<int32_t>().swap(walk);
  out.rowcol.assign(static_cast<size_t>(total - n), 0);
  out.rowpos.assign(static_cast<size_t>(total - n), 0);
  for (int i = 0; i < n; ++i) next[i] = out.rowptr[i];
  for (int k = 0; k < n; ++k)
    for (int p = out.colptr[k] + 1; p < out.colptr[k + 1]; ++p) {
      const int q = next[out.rowidx[p]]++;
      out.rowcol[q] = k;
      out.rowpos[q] = p;
    }
  std::vector<int32_t> depth(n, 1);
  for (int j = 0; j < n; ++j) {  // parents have larger indices
    out.height = std::max(out.height, depth[j]);
    if (out.parent[j] >= 0) depth[out.parent[j]] = std::max(depth[out.parent[j]], depth[j] + 1);
  }
}

// =====================================================================================================================
// device
// =====================================================================================================================
namespace {

constexpr int SOLVER_THREADS = 256;
constexpr int SOLVER_WARPS = SOLVER_THREADS / 32;
// flag block: [0, n) per-column flags, then the control words
constexpr int CTL_COUNTER = 0, CTL_ERROR = 1, CTL_ABORT = 2, CTL_WORDS = 4;
constexpr unsigned long long WAIT_TIMEOUT_NS = 3000000000ull;

struct FactorDev {
  int n;
  const int *colptr, *rowidx, *rowptr, *rowcol, *rowpos, *perm;
  const int *bperm;  // where row j of the permuted system takes its right-hand side from (= perm unless LU moved rows)
  int *flags;  // n + CTL_WORDS
  void *lx, *ux;
};

__device__ __forceinline__ int ld_acquire_s32(const int *p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_s32(int *p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ int ld_relaxed_s32(const int *p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Spins until flags[k] is set.  Polls are relaxed loads with an exponential back-off (thousands of warps wait at any time;
// acquire polls would invalidate the SM's L1 on every probe and crowd the L2 slice of the flag the critical path is about
// to write); one acquire load after the flag was seen orders the reads of the published data.  A wait that lasts longer
// than WAIT_TIMEOUT_NS (it cannot, unless the GPU is shared with a kernel that starves the claimed columns' warps) raises
// the abort word, after which every wait returns at once.
__device__ __forceinline__ void wait_flag(const int *flags, int n, int k) {
  if (ld_acquire_s32(flags + k) != 0) return;
  unsigned long long t0 = 0;
  unsigned ns = 32;
  for (unsigned spin = 1;; ++spin) {
    if (ld_relaxed_s32(flags + k) != 0) {
      (void)ld_acquire_s32(flags + k);
      return;
    }
    __nanosleep(ns);
    if (ns < 256) ns <<= 1;
    if ((spin & 255u) == 0) {
      if (ld_relaxed_s32(flags + n + CTL_ABORT) != 0) return;
      const unsigned long long t = global_ns();
      if (t0 == 0) t0 = t;
      else if (t - t0 > WAIT_TIMEOUT_NS) {
        atomicExch(const_cast<int *>(flags) + n + CTL_ABORT, 1);
        return;
      }
    }
  }
}

__device__ __forceinline__ int lower_bound_dev(const int *a, int len, int key) {
  int lo = 0, hi = len;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < key) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// A's values into the (zeroed) factor slots (amap: >= 0 slot of lx, <= -2 slot -2-v of ux, -1 not used by this
// factorisation).  Added, not stored: a matrix adopted through qdsp_csr_from_device_csr may hold duplicate entries, which
// Eigen's solvers sum as well (canonical matrices have one entry per slot, so the order of the additions never matters).
template <typename T>
__global__ void solver_scatter_kernel(const T *__restrict__ aval, const int *__restrict__ amap, int64_t nnz, T *lx, T *ux) {
  for (int64_t q = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; q < nnz;
       q += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int s = amap[q];
    if (s >= 0) atomicAdd(lx + s, aval[q]);
    else if (s <= -2) atomicAdd(ux + (-2 - s), aval[q]);
  }
}

template <typename T, int MODE>
__global__ void __launch_bounds__(SOLVER_THREADS) solver_factor_kernel(FactorDev f) {
  const int lane = threadIdx.x & 31;
  const int n = f.n;
  T *lx = static_cast<T *>(f.lx);
  T *ux = static_cast<T *>(f.ux);
  int *flags = f.flags;
  for (;;) {
    int j = 0;
    if (lane == 0) j = atomicAdd(flags + n + CTL_COUNTER, 1);
    j = __shfl_sync(FULL, j, 0);
    if (j >= n) return;
    const int cj = __ldg(f.colptr + j), lenj = __ldg(f.colptr + j + 1) - cj;
    const int r0 = __ldg(f.rowptr + j), cnt = __ldg(f.rowptr + j + 1) - r0;
    for (int m = 0; m < cnt; ++m) {
      const int k = __ldg(f.rowcol + r0 + m);
      const int pjk = __ldg(f.rowpos + r0 + m);  // slot of (j, k) in column k
      if (lane == 0) wait_flag(flags, n, k);
      __syncwarp();
      const int ck = __ldg(f.colptr + k), ek = __ldg(f.colptr + k + 1);
      // the multiplier of column k in column j.  Own-column slots are only touched by this warp; other columns' values
      // were published before their flag, and are read through L2 (ld.cg) so that no stale L1 line is seen.
      T u;
      if (MODE == SOLVER_LU) u = __ldcg(ux + pjk);
      else {
        u = __ldcg(lx + pjk);
        if (MODE == SOLVER_LDLT) u *= __ldcg(lx + ck);
      }
      // rows >= j of column k  ->  column j
      const int tail = ek - pjk;
      if (tail == lenj) {
        for (int t = lane; t < lenj; t += 32) __stcg(lx + cj + t, __ldcg(lx + cj + t) - __ldcg(lx + pjk + t) * u);
      } else {
        for (int t = lane; t < tail; t += 32) {
          const int pos = lower_bound_dev(f.rowidx + cj, lenj, __ldg(f.rowidx + pjk + t));
          __stcg(lx + cj + pos, __ldcg(lx + cj + pos) - __ldcg(lx + pjk + t) * u);
        }
      }
      if (MODE == SOLVER_LU) {
        // rows strictly between k and j of column k  ->  U(i, j), i.e. the slots of the later entries of row j
        const int nup = pjk - ck - 1, rest = cnt - 1 - m;
        if (nup == rest) {
          for (int t = lane; t < nup; t += 32) {
            const int s = __ldg(f.rowpos + r0 + m + 1 + t);
            __stcg(ux + s, __ldcg(ux + s) - __ldcg(lx + ck + 1 + t) * u);
          }
        } else {
          for (int t = lane; t < nup; t += 32) {
            const int mm = lower_bound_dev(f.rowcol + r0 + m + 1, rest, __ldg(f.rowidx + ck + 1 + t));
            const int s = __ldg(f.rowpos + r0 + m + 1 + mm);
            __stcg(ux + s, __ldcg(ux + s) - __ldcg(lx + ck + 1 + t) * u);
          }
        }
      }
      __syncwarp();
    }
    // pivot and scaling
    T d = __ldcg(lx + cj);
    bool bad;
    if (MODE == SOLVER_LLT) bad = !(d > T(0));
    else bad = !(d != T(0)) || d != d;
    if (MODE == SOLVER_LU) {
      // static pivoting has no second candidate: a pivot that all but cancelled against its column is a failure to
      // report, not a number to divide by
      T cmax = T(0);
      for (int t = 1 + lane; t < lenj; t += 32) cmax = fmax(cmax, fabs(__ldcg(lx + cj + t)));
      // ... and against column j of U above it (the slots of row j, final by now): the last column has nothing below
      for (int m = lane; m < cnt; m += 32) cmax = fmax(cmax, fabs(__ldcg(ux + __ldg(f.rowpos + r0 + m))));
      for (int o = 16; o > 0; o >>= 1) cmax = fmax(cmax, __shfl_xor_sync(FULL, cmax, o));
      const T eps = sizeof(T) == 8 ? T(2.220446049250313e-16) : T(1.1920929e-07);
      bad = bad || fabs(d) < T(16) * eps * cmax;
    }
    if (bad) {
      if (lane == 0) atomicMin(flags + n + CTL_ERROR, j);
      d = T(1);
    }
    if (MODE == SOLVER_LLT) {
      d = sqrt(d);
      if (lane == 0) __stcg(lx + cj, d);
    }
    for (int t = 1 + lane; t < lenj; t += 32) __stcg(lx + cj + t, __ldcg(lx + cj + t) / d);
    __syncwarp();
    if (lane == 0) {
      __threadfence();
      st_release_s32(flags + j, 1);
    }
  }
}

// ---- triangular solves ------------------------------------------------------------------------------------------------------
// The chain row -> row is the critical path of a solve (elimination-tree height x the latency of one hand-over), so the
// hand-over is ONE memory word: the solution entry itself.  y (and, for the backward pass, the permuted x) starts as the
// all-ones bit pattern -- a NaN no arithmetic produces (hardware NaNs are 0x7ff8.. / 0x7fc0..) -- and a consumer simply
// polls the entry with relaxed gpu-scope loads until it is something else: an aligned 4/8-byte store is single-copy atomic,
// so no flag, no fence and no second load sit between producer and consumer.
template <typename T> struct SolveWord;
template <> struct SolveWord<float> {
  static __device__ __forceinline__ bool pending(float v) { return __float_as_uint(v) == 0xffffffffu; }
  static __device__ __forceinline__ float load(const float *p) {
    float v;
    asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
  }
  static __device__ __forceinline__ void store(float *p, float v) {
    asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
  }
};
template <> struct SolveWord<double> {
  static __device__ __forceinline__ bool pending(double v) { return __double_as_longlong(v) == -1ll; }
  static __device__ __forceinline__ double load(const double *p) {
    double v;
    asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
  }
  static __device__ __forceinline__ void store(double *p, double v) {
    asm volatile("st.relaxed.gpu.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
  }
};

// Polls *p until it holds a value (see above); same time-out / abort protocol as wait_flag.
template <typename T>
__device__ __forceinline__ T wait_value(const T *p, int *ctl) {
  T v = SolveWord<T>::load(p);
  if (!SolveWord<T>::pending(v)) return v;
  unsigned long long t0 = 0;
  for (unsigned spin = 1;; ++spin) {
    v = SolveWord<T>::load(p);
    if (!SolveWord<T>::pending(v)) return v;
    if ((spin & 63u) == 0) __nanosleep(64);
    if ((spin & 1023u) == 0) {
      if (ld_relaxed_s32(ctl + CTL_ABORT) != 0) return T(0);
      const unsigned long long t = global_ns();
      if (t0 == 0) t0 = t;
      else if (t - t0 > WAIT_TIMEOUT_NS) {
        atomicExch(ctl + CTL_ABORT, 1);
        return T(0);
      }
    }
  }
}

// forward substitution  L y = P b  (LLT: L;  LDLT, LU: unit L), one warp per row, rows claimed ascending
template <typename T, int MODE>
__global__ void __launch_bounds__(SOLVER_THREADS) solver_forward_kernel(FactorDev f, const T *__restrict__ b, T *y) {
  const int lane = threadIdx.x & 31;
  const int n = f.n;
  const T *lx = static_cast<const T *>(f.lx);
  int *ctl = f.flags + n;
  for (;;) {
    int j = 0;
    if (lane == 0) j = atomicAdd(ctl + CTL_COUNTER, 1);
    j = __shfl_sync(FULL, j, 0);
    if (j >= n) return;
    const int r0 = __ldg(f.rowptr + j), cnt = __ldg(f.rowptr + j + 1) - r0;
    // everything that does not depend on other rows is read BEFORE the waits
    T bj = T(0), dj = T(1);
    if (lane == 0) {
      bj = b[__ldg(f.bperm + j)];
      if (MODE == SOLVER_LLT) dj = __ldcg(lx + __ldg(f.colptr + j));
    }
    T acc = T(0);
    for (int m = lane; m < cnt; m += 32) {
      const int k = __ldg(f.rowcol + r0 + m);
      const T l = __ldcg(lx + __ldg(f.rowpos + r0 + m));
      acc += l * wait_value(y + k, ctl);
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      T v = bj - acc;
      if (MODE == SOLVER_LLT) v /= dj;
      if (SolveWord<T>::pending(v)) v = T(0) / T(0);  // (cannot happen; keeps the sentinel out of the data)
      SolveWord<T>::store(y + j, v);
    }
  }
}

// backward substitution, rows claimed descending; xp = x in the permuted numbering (the hand-over words), x = the caller's
template <typename T, int MODE>
__global__ void __launch_bounds__(SOLVER_THREADS) solver_backward_kernel(FactorDev f, const T *__restrict__ y, T *xp, T *x) {
  const int lane = threadIdx.x & 31;
  const int n = f.n;
  const T *lx = static_cast<const T *>(f.lx);
  const T *vx = MODE == SOLVER_LU ? static_cast<const T *>(f.ux) : lx;
  int *ctl = f.flags + n;
  for (;;) {
    int c = 0;
    if (lane == 0) c = atomicAdd(ctl + CTL_COUNTER, 1);
    c = __shfl_sync(FULL, c, 0);
    if (c >= n) return;
    const int j = n - 1 - c;
    const int cj = __ldg(f.colptr + j), lenj = __ldg(f.colptr + j + 1) - cj;
    T d = T(1), yj = T(0);
    int dst = 0;
    if (lane == 0) {
      d = __ldcg(lx + cj);
      yj = __ldcg(y + j);  // from the forward pass (previous launch)
      dst = __ldg(f.perm + j);
    }
    T acc = T(0);
    for (int t = 1 + lane; t < lenj; t += 32) {
      const int i = __ldg(f.rowidx + cj + t);
      const T l = __ldcg(vx + cj + t);
      acc += l * wait_value(xp + i, ctl);
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      T v;
      if (MODE == SOLVER_LDLT) v = yj / d - acc;
      else v = (yj - acc) / d;
      if (SolveWord<T>::pending(v)) v = T(0) / T(0);
      SolveWord<T>::store(xp + j, v);
      x[dst] = v;
    }
  }
}

template <typename K>
int solver_grid(K kernel, int n) {
  int occ = 0;
  QDSP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, SOLVER_THREADS, 0));
  const int cap = std::max(1, occ) * device_props().sm_count;
  return std::max(1, std::min(cap, ceil_div(n, SOLVER_WARPS)));
}


}  // namespace

// Row permutation for LU (see the header): the matching of rows to columns that maximises the PRODUCT of the chosen
// entries' magnitudes relative to their row maxima (the objective of Duff & Koster's MC64), by the Hungarian method on the
// sparse pattern: cost(r, c) = log(rowmax_r / |a_rc|) >= 0, shortest augmenting paths (Dijkstra on reduced costs) with row
// / column potentials.  A diagonal entry is given a 10x bonus (cost lowered by log 10, floored at 0), so rows move where it
// buys something, not for a factor of two; a diagonal that is its row's largest entry is matched up front, which leaves
// diagonally dominant and the usual SPD systems untouched at no cost.
// rowof[i] = the row of A that becomes row i of Q A; w = |value| per stored entry (duplicates of a non-canonical CSR are
// looked at one by one, which can only make the choice less good, never invalid).  Returns false when the matrix is
// structurally singular (no zero-free diagonal exists); rowof is then still a permutation and the factorisation reports
// the zero pivot.
bool solver_lu_row_matching_host(int n, const int32_t *ip, const int32_t *ix, const std::vector<double> &w,
                                 std::vector<int32_t> &rowof) {
  std::vector<int32_t> col_of_row(n, -1);
  rowof.assign(n, -1);
  std::vector<double> rowmax(n, 0.0);
  for (int r = 0; r < n; ++r)
    for (int q = ip[r]; q < ip[r + 1]; ++q) rowmax[r] = std::max(rowmax[r], w[q]);
  const double bonus = std::log(10.0);
  auto cost = [&](int r, int q) {
    const double c = std::log(rowmax[r] / w[q]);  // w[q] > 0 is checked by the callers
    return ix[q] == r ? std::max(0.0, c - bonus) : c;
  };
  // start: zero-cost entries (the potentials start at zero, so these are the tight edges).  Diagonals first.
  for (int r = 0; r < n; ++r) {
    double d = 0.0;
    for (int q = ip[r]; q < ip[r + 1]; ++q)
      if (ix[q] == r) d = std::max(d, w[q]);
    if (d > 0.0 && d >= rowmax[r]) { col_of_row[r] = r; rowof[r] = r; }
  }
  for (int r = 0; r < n; ++r) {
    if (col_of_row[r] >= 0) continue;
    for (int q = ip[r]; q < ip[r + 1]; ++q)
      if (w[q] > 0.0 && rowof[ix[q]] < 0 && cost(r, q) == 0.0) { col_of_row[r] = ix[q]; rowof[ix[q]] = r; break; }
  }
  // the rest: one shortest augmenting path per free row
  bool complete = true;
  std::vector<double> u(n, 0.0), v(n, 0.0), dist(n, 0.0);
  std::vector<int32_t> via(n, -1), stamp(n, -1), done(n, -1), touched;
  using Item = std::pair<double, int32_t>;
  std::vector<Item> heap;
  auto cmp = [](const Item &a, const Item &b) { return a.first > b.first; };
  for (int r0 = 0; r0 < n; ++r0) {
    if (col_of_row[r0] >= 0) continue;
    heap.clear();
    touched.clear();
    auto relax = [&](int r, double base) {
      for (int q = ip[r]; q < ip[r + 1]; ++q) {
        if (!(w[q] > 0.0)) continue;
        const int c = ix[q];
        if (done[c] == r0) continue;
        const double nd = base + std::max(0.0, cost(r, q) - u[r] - v[c]);  // (max: rounding may leave -1e-17)
        if (stamp[c] != r0 || nd < dist[c]) {
          if (stamp[c] != r0) { stamp[c] = r0; touched.push_back(c); }
          dist[c] = nd;
          via[c] = r;
          heap.push_back({nd, c});
          std::push_heap(heap.begin(), heap.end(), cmp);
        }
      }
    };
    relax(r0, 0.0);
    int found = -1;
    double D = 0.0;
    while (!heap.empty()) {
      std::pop_heap(heap.begin(), heap.end(), cmp);
      const Item it = heap.back();
      heap.pop_back();
      const int c = it.second;
      if (done[c] == r0 || it.first > dist[c]) continue;  // stale entry
      done[c] = r0;
      if (rowof[c] < 0) { found = c; D = it.first; break; }
      relax(rowof[c], it.first);
    }
    if (found < 0) { complete = false; continue; }
    // potentials: finalised columns (and their rows) move by (D - dist); the new row by D.  Reduced costs stay >= 0 and
    // every matched edge stays tight.
    for (int c : touched)
      if (done[c] == r0 && c != found) {
        const double delta = D - dist[c];
        v[c] -= delta;
        u[rowof[c]] += delta;
      }
    u[r0] += D;
    for (int c = found; c >= 0;) {  // flip the path
      const int r = via[c], prev = col_of_row[r];
      col_of_row[r] = c;
      rowof[c] = r;
      c = r == r0 ? -1 : prev;
    }
  }
  if (!complete) {  // leftovers in index order: still a permutation
    int c = 0;
    for (int r = 0; r < n; ++r) {
      if (col_of_row[r] >= 0) continue;
      while (rowof[c] >= 0) ++c;
      rowof[c] = r;
      col_of_row[r] = c;
    }
  }
  return complete;
}

DirectSolver::DirectSolver(int dtype, int solver_type, int ordering) : dtype_(dtype), type_(solver_type), ordering_(ordering) {
  QDSP_REQUIRE(dtype == QDSP_F32 || dtype == QDSP_F64, QDSP_ERR_INVALID, "Not supported sparse solver data type");
  QDSP_REQUIRE(solver_type >= SOLVER_LLT && solver_type <= SOLVER_LU, QDSP_ERR_INVALID, "Not supported sparse solver type");
  QDSP_REQUIRE(ordering == 0 || ordering == 1, QDSP_ERR_INVALID, "Not supported sparse solver ordering");
}
DirectSolver::~DirectSolver() = default;

void DirectSolver::analyze_pattern(const CsrMatrix &A, cudaStream_t stream) {
  QDSP_REQUIRE(A.rows == A.cols, QDSP_ERR_INVALID, "solver: the matrix must be square");
  QDSP_REQUIRE(A.dtype == dtype_, QDSP_ERR_INVALID, "solver: dtype of the matrix differs from the solver's");
  const int n = A.rows;
  std::vector<int32_t> ip(n + 1), ix(static_cast<size_t>(A.nnz));
  QDSP_CUDA(cudaMemcpyAsync(ip.data(), A.indptr.get(), sizeof(int32_t) * (n + 1), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaMemcpyAsync(ix.data(), A.indices.get(), sizeof(int32_t) * A.nnz, cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  // LU: rows are exchanged up front (static pivoting); from here on `ip`/`ix` describe Q A and row r of A is row rinv[r]
  std::vector<int32_t> rowof(n), rinv(n);
  std::iota(rowof.begin(), rowof.end(), 0);
  std::iota(rinv.begin(), rinv.end(), 0);
  const std::vector<int32_t> ip_a = ip;  // the rows of A as stored (the slot map below walks A, not Q A)
  const std::vector<int32_t> ix_a = ix;
  if (type_ == SOLVER_LU && n > 0 && A.nnz > 0) {
    std::vector<double> w(static_cast<size_t>(A.nnz));
    if (dtype_ == QDSP_F64) {
      std::vector<double> v(w.size());
      QDSP_CUDA(cudaMemcpyAsync(v.data(), A.values.get(), sizeof(double) * A.nnz, cudaMemcpyDeviceToHost, stream));
      QDSP_CUDA(cudaStreamSynchronize(stream));
      for (size_t q = 0; q < w.size(); ++q) w[q] = std::fabs(v[q]);
    } else {
      std::vector<float> v(w.size());
      QDSP_CUDA(cudaMemcpyAsync(v.data(), A.values.get(), sizeof(float) * A.nnz, cudaMemcpyDeviceToHost, stream));
      QDSP_CUDA(cudaStreamSynchronize(stream));
      for (size_t q = 0; q < w.size(); ++q) w[q] = std::fabs(static_cast<double>(v[q]));
    }
    for (double &x : w)
      if (!(x == x)) x = 0.0;  // NaN entries never attract a pivot
    solver_lu_row_matching_host(n, ip.data(), ix.data(), w, rowof);
    bool identity = true;
    for (int i = 0; i < n; ++i) {
      rinv[rowof[i]] = i;
      identity = identity && rowof[i] == i;
    }
    if (!identity) {
      std::vector<int32_t> ipq(n + 1, 0), ixq(ix.size());
      for (int i = 0; i < n; ++i) ipq[i + 1] = ipq[i] + (ip_a[rowof[i] + 1] - ip_a[rowof[i]]);
      for (int i = 0; i < n; ++i)
        std::copy(ix_a.begin() + ip_a[rowof[i]], ix_a.begin() + ip_a[rowof[i] + 1], ixq.begin() + ipq[i]);
      ip.swap(ipq);
      ix.swap(ixq);
    }
  }
  std::vector<int32_t> perm;
  // both ordering names select the same symmetric minimum-degree ordering, like the reference's arch=cuda path, which
  // ignores the argument and calls csrsymamdHost (make_cusparse_solver, sparse_solver.cpp:548-558)
  solver_order_host(n, ip.data(), ix.data(), perm);
  solver_symbolic_host(n, ip.data(), ix.data(), perm, sym_);
  // where each entry of A lands.  LLT / LDLT read the LOWER triangle of A only (Eigen::Lower, sparse_solver.h:9-13).
  std::vector<int32_t> amap(static_cast<size_t>(A.nnz), -1);
  for (int r = 0; r < n; ++r)
    for (int q = ip_a[r]; q < ip_a[r + 1]; ++q) {
      const int c = ix_a[q];
      if (type_ != SOLVER_LU && c > r) continue;
      const int pr = sym_.pinv[rinv[r]], pc = sym_.pinv[c];
      int hi = pr, lo = pc;
      bool upper = false;
      if (pr < pc) { hi = pc; lo = pr; upper = type_ == SOLVER_LU; }
      const int32_t *b = sym_.rowidx.data() + sym_.colptr[lo], *e = sym_.rowidx.data() + sym_.colptr[lo + 1];
      const int32_t *it = std::lower_bound(b, e, hi);
      QDSP_REQUIRE(it != e && *it == hi, QDSP_ERR_INVALID, "solver: entry outside the analysed pattern");
      const int slot = static_cast<int>(it - sym_.rowidx.data());
      amap[q] = upper ? -2 - slot : slot;
    }
  auto up = [&](DeviceBuffer &d, const std::vector<int32_t> &h) {
    d.alloc(std::max<size_t>(h.size(), 1) * sizeof(int32_t));
    if (!h.empty()) QDSP_CUDA(cudaMemcpyAsync(d.get(), h.data(), h.size() * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
  };
  up(d_colptr_, sym_.colptr);
  up(d_rowidx_, sym_.rowidx);
  up(d_rowptr_, sym_.rowptr);
  up(d_rowcol_, sym_.rowcol);
  up(d_rowpos_, sym_.rowpos);
  up(d_perm_, sym_.perm);
  std::vector<int32_t> bperm(n);
  for (int j = 0; j < n; ++j) bperm[j] = rowof[sym_.perm[j]];  // (Q b)[i] = b[rowof[i]], then the symmetric permutation
  up(d_bperm_, bperm);
  up(d_amap_, amap);
  const size_t vb = dtype_ == QDSP_F64 ? 8 : 4;
  d_lx_.alloc(static_cast<size_t>(sym_.factor_nnz) * vb);
  if (type_ == SOLVER_LU) d_ux_.alloc(static_cast<size_t>(sym_.factor_nnz) * vb);
  d_flags_.alloc(static_cast<size_t>(n + CTL_WORDS) * sizeof(int));
  d_work_.alloc(2 * static_cast<size_t>(std::max(n, 1)) * vb);  // y | permuted x
  QDSP_CUDA(cudaStreamSynchronize(stream));  // the host vectors above go out of scope
  nnz_a_ = A.nnz;
  analyzed_ = true;
  factorized_ = false;
  zero_pivot_ = -1;
}

namespace {
void reset_flags(int *flags, int n, cudaStream_t stream) {
  QDSP_CUDA(cudaMemsetAsync(flags, 0, static_cast<size_t>(n + CTL_WORDS) * sizeof(int), stream));
  const int big = INT_MAX;
  QDSP_CUDA(cudaMemcpyAsync(flags + n + CTL_ERROR, &big, sizeof(int), cudaMemcpyHostToDevice, stream));
}
void reset_ctl(int *flags, int n, cudaStream_t stream) {
  static const int init[CTL_WORDS] = {0, INT_MAX, 0, 0};
  QDSP_CUDA(cudaMemcpyAsync(flags + n, init, sizeof(init), cudaMemcpyHostToDevice, stream));
}
void read_ctl(const int *flags, int n, cudaStream_t stream, int *ctl) {
  QDSP_CUDA(cudaMemcpyAsync(ctl, flags + n, CTL_WORDS * sizeof(int), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  QDSP_REQUIRE(ctl[CTL_ABORT] == 0, QDSP_ERR_CUDA, "solver: a dependency wait timed out on the device");
}
}  // namespace

template <typename T, int MODE>
static void factor_launch(const FactorDev &f, cudaStream_t stream) {
  const int grid = solver_grid(solver_factor_kernel<T, MODE>, f.n);
  solver_factor_kernel<T, MODE><<<grid, SOLVER_THREADS, 0, stream>>>(f);
  QDSP_LAUNCH_CHECK();
}
template <typename T, int MODE>
static void solve_launch(const FactorDev &f, const T *b, T *y, T *x, cudaStream_t stream) {
  // y and the permuted x (y + n) start as the all-ones "pending" pattern; only the control words of the flag block are used
  QDSP_CUDA(cudaMemsetAsync(y, 0xff, 2 * static_cast<size_t>(f.n) * sizeof(T), stream));
  reset_ctl(f.flags, f.n, stream);
  solver_forward_kernel<T, MODE><<<solver_grid(solver_forward_kernel<T, MODE>, f.n), SOLVER_THREADS, 0, stream>>>(f, b, y);
  QDSP_LAUNCH_CHECK();
  QDSP_CUDA(cudaMemsetAsync(f.flags + f.n + CTL_COUNTER, 0, sizeof(int), stream));
  solver_backward_kernel<T, MODE><<<solver_grid(solver_backward_kernel<T, MODE>, f.n), SOLVER_THREADS, 0, stream>>>(f, y, y + f.n, x);
  QDSP_LAUNCH_CHECK();
}

#define SOLVER_DISPATCH(FN, ...)                                                             \
  do {                                                                                       \
    if (dtype_ == QDSP_F64) {                                                                \
      if (type_ == SOLVER_LLT) FN<double, SOLVER_LLT>(__VA_ARGS__);                          \
      else if (type_ == SOLVER_LDLT) FN<double, SOLVER_LDLT>(__VA_ARGS__);                   \
      else FN<double, SOLVER_LU>(__VA_ARGS__);                                               \
    } else {                                                                                 \
      if (type_ == SOLVER_LLT) FN<float, SOLVER_LLT>(__VA_ARGS__);                           \
      else if (type_ == SOLVER_LDLT) FN<float, SOLVER_LDLT>(__VA_ARGS__);                    \
      else FN<float, SOLVER_LU>(__VA_ARGS__);                                                \
    }                                                                                        \
  } while (0)

void DirectSolver::factorize(const CsrMatrix &A, cudaStream_t stream) {
  QDSP_REQUIRE(analyzed_, QDSP_ERR_STATE, "solver: analyze_pattern() must be called before factorize()");
  QDSP_REQUIRE(A.rows == sym_.n && A.cols == sym_.n && A.nnz == nnz_a_ && A.dtype == dtype_, QDSP_ERR_INVALID,
               "solver: the matrix does not match the analysed pattern");
  const int n = sym_.n;
  factorized_ = false;
  zero_pivot_ = -1;
  if (n == 0) { factorized_ = true; return; }
  const size_t vb = dtype_ == QDSP_F64 ? 8 : 4;
  QDSP_CUDA(cudaMemsetAsync(d_lx_.get(), 0, static_cast<size_t>(sym_.factor_nnz) * vb, stream));
  if (type_ == SOLVER_LU) QDSP_CUDA(cudaMemsetAsync(d_ux_.get(), 0, static_cast<size_t>(sym_.factor_nnz) * vb, stream));
  const int sgrid = std::max(1, std::min(device_props().sm_count * 8, ceil_div(A.nnz, 256)));
  if (dtype_ == QDSP_F64)
    solver_scatter_kernel<double><<<sgrid, 256, 0, stream>>>(A.values.get<double>(), d_amap_.get<int>(), A.nnz,
                                                              d_lx_.get<double>(), d_ux_.get<double>());
  else
    solver_scatter_kernel<float><<<sgrid, 256, 0, stream>>>(A.values.get<float>(), d_amap_.get<int>(), A.nnz,
                                                             d_lx_.get<float>(), d_ux_.get<float>());
  QDSP_LAUNCH_CHECK();
  reset_flags(d_flags_.get<int>(), n, stream);
  FactorDev f{n, d_colptr_.get<int>(), d_rowidx_.get<int>(), d_rowptr_.get<int>(), d_rowcol_.get<int>(),
              d_rowpos_.get<int>(), d_perm_.get<int>(), d_bperm_.get<int>(), d_flags_.get<int>(), d_lx_.get(), d_ux_.get()};
  SOLVER_DISPATCH(factor_launch, f, stream);
  int ctl[CTL_WORDS];
  read_ctl(d_flags_.get<int>(), n, stream, ctl);
  zero_pivot_ = ctl[CTL_ERROR] == INT_MAX ? -1 : ctl[CTL_ERROR];
  factorized_ = true;
}

void DirectSolver::solve(const void *b, void *x, cudaStream_t stream) {
  QDSP_REQUIRE(factorized_, QDSP_ERR_STATE, "solver: factorize() must be called before solve()");
  QDSP_REQUIRE(b != nullptr && x != nullptr, QDSP_ERR_INVALID, "solver: null vector");
  const int n = sym_.n;
  if (n == 0) return;
  FactorDev f{n, d_colptr_.get<int>(), d_rowidx_.get<int>(), d_rowptr_.get<int>(), d_rowcol_.get<int>(),
              d_rowpos_.get<int>(), d_perm_.get<int>(), d_bperm_.get<int>(), d_flags_.get<int>(), d_lx_.get(), d_ux_.get()};
  if (dtype_ == QDSP_F64) {
    const double *bb = static_cast<const double *>(b);
    double *yy = d_work_.get<double>(), *xx = static_cast<double *>(x);
    if (type_ == SOLVER_LLT) solve_launch<double, SOLVER_LLT>(f, bb, yy, xx, stream);
    else if (type_ == SOLVER_LDLT) solve_launch<double, SOLVER_LDLT>(f, bb, yy, xx, stream);
    else solve_launch<double, SOLVER_LU>(f, bb, yy, xx, stream);
  } else {
    const float *bb = static_cast<const float *>(b);
    float *yy = d_work_.get<float>(), *xx = static_cast<float *>(x);
    if (type_ == SOLVER_LLT) solve_launch<float, SOLVER_LLT>(f, bb, yy, xx, stream);
    else if (type_ == SOLVER_LDLT) solve_launch<float, SOLVER_LDLT>(f, bb, yy, xx, stream);
    else solve_launch<float, SOLVER_LU>(f, bb, yy, xx, stream);
  }
  int ctl[CTL_WORDS];
  read_ctl(d_flags_.get<int>(), n, stream, ctl);
}

void DirectSolver::solve_host(const void *b_host, void *x_host, cudaStream_t stream) {
  QDSP_REQUIRE(b_host != nullptr && x_host != nullptr, QDSP_ERR_INVALID, "solver: null vector");
  const size_t bytes = static_cast<size_t>(sym_.n) * (dtype_ == QDSP_F64 ? 8 : 4);
  DeviceBuffer db(bytes), dx(bytes);
  QDSP_CUDA(cudaMemcpyAsync(db.get(), b_host, bytes, cudaMemcpyHostToDevice, stream));
  solve(db.get(), dx.get(), stream);
  QDSP_CUDA(cudaMemcpyAsync(x_host, dx.get(), bytes, cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
}

}  // namespace qdsp
