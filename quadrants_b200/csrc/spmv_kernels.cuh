// spmv_kernels.cuh -- row-length-binned CSR SpMV, y = A x  (fp32 and fp64), optionally fused with the p.Ap dot of CG.
//
// Replaces CuSparseMatrix::spmv (quadrants/program/sparse_matrix.cpp:654-686: cusparseCreate + SpMV_bufferSize +
// cuMemAlloc + cusparseSpMV(CSR_ALG1, f32) + destroy + cuMemFree on EVERY call) and is held to the arithmetic of
// EigenSparseMatrix::spmv (:265-279).
//
// HBM-bound.  Algorithmic bytes per SpMV: nnz*(sv+4) + (n+1)*4 + m*sv + n*sv  (SURVEY 8(d)); the matrix stream
// (indices + values) is ~82 % of that for the 7-point Poisson matrix, so the kernel is organised around streaming it:
//
//   * persistent CTAs (a multiple of the SM count), each owning a CONTIGUOUS run of row blocks, so the x windows of
//     neighbouring blocks are re-used from L1 and every CTA reads one sequential stream of the matrix;
//   * a row block = R = 256 / L rows, L = lanes per row chosen ONCE per matrix from its mean row length (the bins:
//     L = 1, 2, 4, ... 32 for mean lengths up to 7.2, 14, 29, ... 230);
//   * warp-specialised: one producer thread walks the blocks and moves each block's slice of indptr / indices / values
//     from HBM into a ring of shared-memory stages with TMA bulk copies (cp.async.bulk + mbarrier complete_tx, L2
//     evict-first: the matrix is read once, x should stay in L2); eight consumer warps wait on the stage's "full"
//     barrier, walk their rows in shared memory, gather x through the read-only path (for banded matrices the k-th
//     entries of 32 consecutive rows are 32 consecutive x: a coalesced gather), and release the stage;
//   * L = 1: one thread per row adding left to right with fma -- Eigen's order, so the result is BIT-IDENTICAL to the
//     reference's product (col_major: one accumulator; row_major: two alternating accumulators).  L > 1: the L lanes
//     of a row stride it and combine with warp shuffles (held to the 1e-5 / 1e-12 per-row tolerance);
//   * a block whose slice does not fit a stage (rows far longer than the mean) is streamed straight from global
//     memory, one warp per row.
//   Tensor cores are deliberately unused: 2 flop per 8-12 bytes.
#pragma once
#include "common.cuh"
#include "dist_dev.cuh"
#include "tma.cuh"

namespace qdsp {

constexpr int SPMV_CONSUMERS = 256;                 // consumer threads per CTA (8 warps)
constexpr int SPMV_THREADS = SPMV_CONSUMERS + 32;   // + the producer warp
constexpr int SPMV_CAP = 2048;                      // non-zeros per stage
constexpr int SPMV_STAGES = 3;
constexpr int SPMV_STREAM_STAGES = 2;                // stream path: gather latency, not the TMA depth, is what needs
                                                    // hiding, so fewer stages -> more resident CTAs per SM
constexpr int SPMV_PTR_SLOTS = 264;                 // indptr slice (<= 3 + 257 entries) + the 4-word block descriptor
constexpr int SPMV_DESC_SLOT = 260;                 // stream path: where the producer leaves the block descriptor
constexpr int SPMV_STREAM_CAP = SPMV_CAP - 4;       // non-zeros per block of the stream path (16-byte alignment slack)
constexpr int SPMV_STREAM_THREAD_ROWS = 64;         // blocks with more rows than this sum one row per thread

__device__ __forceinline__ float fma_t(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fma_t(double a, double b, double c) { return fma(a, b, c); }
// unfused multiply-add (the compiler may not contract these intrinsics)
__device__ __forceinline__ float mad_unfused(float a, float b, float c) { return __fadd_rn(__fmul_rn(a, b), c); }
__device__ __forceinline__ double mad_unfused(double a, double b, double c) { return __dadd_rn(__dmul_rn(a, b), c); }

// lanes per row for a matrix of this mean row length (host side; the bins of the north star's "row-length-binned")
inline int spmv_lanes_per_row(int64_t nnz, int rows) {
  const double mean = rows > 0 ? static_cast<double>(nnz) / rows : 0.0;
  int L = 1;
  while (L < 32 && (SPMV_CONSUMERS / L) * mean * 1.1 + 16 > SPMV_CAP) L <<= 1;
  return L;
}

// AUX: per-row side arrays (R <= 256 entries of T each) that travel with the block through the same TMA ring -- the
// one-reduction CG kernel needs r[row] and D^-1[row] next to the product, and a plain load there would expose a full
// DRAM latency per block
// CAP: non-zeros per stage.  The merged CG kernel trades 192 of them for its AUX arrays so that four CTAs still fit an SM.
constexpr int SPMV_CAP_MERGED = 1856;
template <typename T, int AUX = 0, int CAP = SPMV_CAP>
__host__ __device__ constexpr size_t spmv_stage_bytes() {
  return CAP * sizeof(int) + CAP * sizeof(T) + SPMV_PTR_SLOTS * sizeof(int) + AUX * SPMV_CONSUMERS * sizeof(T);
}
template <typename T, int AUX = 0, int CAP = SPMV_CAP>
__host__ __device__ constexpr size_t spmv_smem_bytes() {
  return SPMV_STAGES * spmv_stage_bytes<T, AUX, CAP>() + 2 * SPMV_STAGES * sizeof(unsigned long long) + 128 /*alignment slack*/;
}

// Stream path, hot columns: a matrix with a skewed column distribution (power-law graphs / meshes: a few columns carry a
// large share of the non-zeros) gathers those x entries from SHARED memory.  The plan (spmv.cu) picks the SPMV_HOT_BYTES /
// sizeof(T) most referenced columns and keeps a copy of `indices` in which they appear as ~slot (negative); every CTA loads
// the hot x entries once per launch.  Why: the irregular gather is bound by the SM's outstanding-miss capacity (each
// 8-byte gather is its own 32-byte L2 sector request; ncu: L1 hit 7 %, nothing else above 45 %), so every gather served
// on-chip lifts throughput in proportion.
constexpr int SPMV_HOT_BYTES = 16384;
constexpr int SPMV_HOT_OFFSET = 64;  // behind the mbarriers (2 * stages * 8 bytes), 16-byte aligned
template <typename T>
__host__ __device__ constexpr size_t spmv_stream_smem_bytes(bool hot = false) {
  return SPMV_STREAM_STAGES * spmv_stage_bytes<T>() + 2 * SPMV_STREAM_STAGES * sizeof(unsigned long long) + 128 +
         (hot ? SPMV_HOT_BYTES : 0);
}

// The operand gather.  Stand-alone kernels read x through the read-only (non-coherent) path.  Inside the persistent CG
// kernel (cg.cu, cg_fused_kernel) the operand is REWRITTEN between two products of the same launch, by other CTAs and,
// multi-GPU, by peer ranks: there the gather is a plain weak load (ld.global.ca), which the acquire of the grid barrier /
// halo wait orders after those writes (the acquire invalidates the SM's L1); .nc loads are outside the memory model.
template <bool COHERENT, typename T>
__device__ __forceinline__ T ld_x(const T *p) {
  // -DQDSP_GATHER_COHERENT makes every gather a coherent load, which is what the memory model asks for where peers write
  // ghost entries during a stand-alone kernel's lifetime (amode / lazy halo).  Measured on one B200, same box, round 3
  // (profiles/r03e_gather_ab.txt): C2 CG and both SpMV paths unchanged, but the C5 (irregular, fp64) CG drops from 4714
  // to 4374 it/s, so the default stays .nc for the stand-alone kernels; the halo waits end in an acquire (load or
  // fence, dist_dev.cuh), which on sm_100 invalidates the L1 lines an .nc load would otherwise hit.
#ifdef QDSP_GATHER_COHERENT
  return __ldca(p);
#else
  if constexpr (COHERENT) return __ldca(p);
  else return __ldg(p);
#endif
}

// Explicit shared-memory accesses for the stream path.  The stage pointers are computed from an aligned byte base, so the
// compiler has lost the address space and emits GENERIC loads (LD.E): those queue behind the SM's outstanding divergent
// gathers in the global-memory pipe (ncu, profiles/r02ae: a single descriptor read sampled at ~800 cycles) and serialised
// the eight index reads of a gather batch.  ld.shared / st.shared take the short path.
__device__ __forceinline__ int lds_i32(const int *p) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ int lds_u8(const unsigned char *p) {
  unsigned v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)));
  return static_cast<int>(v);
}
__device__ __forceinline__ int4 lds_i32x4(const int *p) {
  int4 v;
  asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ float lds_val(const float *p) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ double lds_val(const double *p) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ void sts_val(float *p, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_u32(p)), "f"(v) : "memory");
}
__device__ __forceinline__ void sts_val(double *p, double v) {
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(smem_u32(p)), "d"(v) : "memory");
}

// mbarrier ring of a persistent SpMV: `full` (1 arrival + tx bytes) and `empty` (one arrival per consumer warp) per stage
template <int STAGES>
__device__ __forceinline__ void spmv_ring_init(unsigned long long *full, unsigned long long *empty) {
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], SPMV_CONSUMERS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
}

template <typename T>
struct SpmvStage {
  int *idx;
  T *val;
  int *ptr;
  T *aux;  // AUX arrays of SPMV_CONSUMERS entries each
};
template <typename T, int AUX = 0, int CAP = SPMV_CAP>
__device__ __forceinline__ SpmvStage<T> spmv_stage(unsigned char *base, int s) {
  unsigned char *p = base + static_cast<size_t>(s) * spmv_stage_bytes<T, AUX, CAP>();
  SpmvStage<T> st;
  st.val = reinterpret_cast<T *>(p);
  st.idx = reinterpret_cast<int *>(p + CAP * sizeof(T));
  st.ptr = reinterpret_cast<int *>(p + CAP * sizeof(T) + CAP * sizeof(int));
  st.aux = reinterpret_cast<T *>(p + CAP * sizeof(T) + CAP * sizeof(int) + SPMV_PTR_SLOTS * sizeof(int));
  return st;
}

// Irregular matrices (row lengths far from the mean: FEM power-law tails) use the "stream" path instead of fixed-size
// row blocks.  The plan (host-built once per matrix, spmv.cu) cuts the rows into blocks of <= SPMV_STREAM_CAP
// non-zeros and <= 256 rows; a row longer than that becomes several single-row SEGMENT blocks whose partial sums are
// added in segment order by the last CTA to finish (deterministic).  CSR-adaptive in spirit (Greathouse & Daga).
struct SpmvPlanDev {
  const int4 *blocks = nullptr;    // {row0, nrows (> 0) or -(segment id + 1), nz0, nz1}
  int nblk = 0;
  int n_long = 0;
  const int2 *long_rows = nullptr; // n_long + 1 entries {row, first segment id}; the last one is a sentinel
  void *partials = nullptr;        // one T per segment
  unsigned *ticket = nullptr;
  // hot columns (0 = off): hot_cols[slot] = column id; hot_idx = copy of `indices` with hot columns stored as ~slot
  int hot_n = 0;
  const int *hot_cols = nullptr;
  const int *hot_idx = nullptr;
};

// Work distribution of the uniform path.  work == nullptr: every CTA owns a contiguous run of ceil(nblk / grid) row blocks.
// Otherwise CTA c owns the contiguous run [c * static_blocks, (c + 1) * static_blocks) and the blocks from
// grid * static_blocks on are handed out in chunks of `chunk` through the device counter *work -- the CTAs of an SM do not
// get equal shares of HBM bandwidth (measured: the first CTA of a static partition was done after 176 us of a 229 us
// product), so the last quarter of the matrix goes to whoever is free.  The kernel resets both counters itself (the last
// producer thread to run out of work), so consecutive launches need no memset.
struct SpmvSched {
  unsigned *work = nullptr;  // next dynamic block offset
  unsigned *done = nullptr;  // producers that have finished
  int static_blocks = 0;
  int chunk = 2;
  // Diagonal-coded column indices (uniform path, one thread per row; spmv.cu spmv_plan_codes): when the matrix has at
  // most 256 distinct values of (column - row) -- stencils, banded and block-banded matrices, also after the multi-GPU
  // renumbering -- the plan keeps ONE BYTE per non-zero, codes[k] = position of (col - row) in the sorted table `tab`,
  // and the kernel streams that instead of the 4-byte index: nnz * (sv + 1) instead of nnz * (sv + 4) bytes per product
  // (fp32 7-point stencil: 0.63 of the canonical traffic).  Same entries, same order, same arithmetic => the same bits;
  // the canonical `indices` array stays (export, oversized blocks).  nullptr = canonical indices.
  const unsigned char *codes = nullptr;
  const int *tab = nullptr;
  int ntab = 0;
};

__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(SPMV_CONSUMERS) : "memory"); }

// ---- row sinks: what a kernel does with each finished y[row] besides storing it ----------------------------------------
// `pre(row, stage aux, row in block)` is called before the row's products (its loads overlap them), `row(row, y, pre)` once per finished row by
// the thread that stores it, `segment(row, partial)` for one piece of a long row of the stream path (linear sinks only).
struct NoSink {
  static constexpr int AUX = 0;
  static constexpr int CAP = SPMV_CAP;
  struct Pre {};
  template <typename T> __device__ __forceinline__ Pre pre(int, const T *, int) const { return Pre{}; }
  template <typename T> __device__ __forceinline__ void row(int, T, Pre) {}
  template <typename T> __device__ __forceinline__ void segment(int, T) {}
};
// sum_i w[i] * y[i] in double (CG: p.Ap)
// (COHERENT: w is rewritten between two products of the same launch -- see ld_x)
template <typename T, bool COHERENT = false>
struct DotSink {
  const T *w;
  double acc = 0.0;
  static constexpr int AUX = 0;
  static constexpr int CAP = SPMV_CAP;
  struct Pre { T w; };
  __device__ __forceinline__ Pre pre(int i, const T *, int) const { return Pre{ld_x<COHERENT>(w + i)}; }
  __device__ __forceinline__ void row(int, T y, Pre q) { acc += static_cast<double>(q.w * y); }
  __device__ __forceinline__ void segment(int i, T part) { acc += static_cast<double>(ld_x<COHERENT>(w + i) * part); }
};
// The seven inner products of the one-reduction CG iteration (cg.cu), t = A p, z = D^-1 r:
//   v0 = p.t  v1 = t.z  v2 = t.D^-1 t  v3 = r.t  v4 = t.t  v5 = r.z  v6 = r.r      (all accumulated in double)
template <typename T>
struct MergedSink {
  const T *__restrict__ p;
  const T *__restrict__ r;
  const T *__restrict__ dinv;
  double v[7] = {0, 0, 0, 0, 0, 0, 0};
  static constexpr int AUX = 2;  // r and D^-1 of the block's rows ride the TMA ring (aux_src); p[row] is an L1 hit
  static constexpr int CAP = SPMV_CAP_MERGED;
  struct Pre { T p, r, d; };
  __device__ __forceinline__ const T *aux_src(int k) const { return k == 0 ? r : dinv; }
  __device__ __forceinline__ Pre pre(int i, const T *aux, int j) const {
    return Pre{p[i], aux[j], aux[SPMV_CONSUMERS + j]};
  }
  __device__ __forceinline__ void row(int, T y, Pre q) {
    const double t = static_cast<double>(y), pi = static_cast<double>(q.p), ri = static_cast<double>(q.r);
    const double di = static_cast<double>(q.d), zi = static_cast<double>(static_cast<T>(q.d * q.r));
    v[0] += pi * t; v[1] += t * zi; v[2] += t * di * t; v[3] += ri * t; v[4] += t * t; v[5] += ri * zi; v[6] += ri * ri;
  }
};

// The whole SpMV of one persistent CTA.  LPR = lanes per row.  Every finished row is also handed to `sink` (per-thread
// state: each consumer thread accumulates over the rows it stored; producer-warp threads never call it).
// Multi-GPU (dist.world > 1): x = [owned | ghosts]; the ghosts are stored by the neighbour ranks.  Only row blocks
// flagged in dist.ghost_blocks read a ghost, so the producer waits for the halo flags right before it hands the first
// such block to the consumers (its acquire reaches them through the stage's mbarrier) and interior blocks start at
// once.  The ghost segment begins on a 128-byte boundary, so no L1 line holds both owned and ghost entries.
// `indptr`, `indices`, `values` must be readable 16 bytes past their ends (the matrices of this library carry 64).
// FUSED (the persistent CG kernel): the caller initialised the ring once (spmv_ring_init) and calls this function once
// per CG iteration `iter` from the producer thread and from the consumers INDEPENDENTLY (the producer streams the constant
// matrix arrays and may run ahead of the consumers by the depth of the ring); ring positions continue across calls;
// nothing here synchronises the whole CTA; the halo wait moves to the consumers; x is read coherently (ld_x).
template <typename T, int LPR, bool ROW_MAJOR_ORDER, typename Sink, bool FUSED = false, bool CODED = false>
__device__ __forceinline__ void spmv_persistent(unsigned char *smem_raw, int rows, const int *__restrict__ indptr,
                                                const int *__restrict__ indices, const T *__restrict__ values,
                                                const T *x, T *y, Sink &sink, const DistDev &dist, int iter = 0,
                                                const SpmvSched sched = SpmvSched{}) {
  constexpr int R = SPMV_CONSUMERS / LPR;  // rows per block
  constexpr int AUX = Sink::AUX;
  constexpr int CAP = Sink::CAP;
  unsigned char *base = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  auto *full = reinterpret_cast<unsigned long long *>(base + SPMV_STAGES * spmv_stage_bytes<T, AUX, CAP>());
  auto *empty = full + SPMV_STAGES;
  const int t = threadIdx.x;
  __shared__ int s_tab[256];  // column - row of every code (diagonal-coded indices, see SpmvSched)
  static_assert(!CODED || (LPR == 1 && !FUSED), "diagonal codes: one thread per row, stand-alone kernels only");
  constexpr bool coded = CODED;  // a separate instantiation: the canonical kernel's code is untouched by the coded path
  if (coded)  // all 256 entries are defined: the row body looks codes up before it knows whether the slot is in the row
    for (int i = t; i < 256; i += SPMV_THREADS) s_tab[i] = i < sched.ntab ? sched.tab[i] : 0;
  if constexpr (!FUSED) spmv_ring_init<SPMV_STAGES>(full, empty);  // (its __syncthreads also publishes s_tab)
  // no per-block ghost map for this block size: every CTA waits for the halo before its first gather
  const bool lazy_halo = dist.world > 1 && dist.ghost_blocks != nullptr && dist.ghost_block_rows == R;
  if constexpr (!FUSED) {
    if (dist.world > 1 && !lazy_halo) {
      if (t == 0) dist_halo_wait(dist);
      __syncthreads();
    }
  }

  const int nblk = (rows + R - 1) / R;
  // work items of this CTA.  Plain: a contiguous run of blocks.  Ordered (multi-GPU): a contiguous run of the interior
  // blocks, then every gridDim.x-th ghost block; the producer leaves the block id in the stage for the consumers.
  // Ordered AND dynamic: the CTAs take their static runs and the dynamic chunks from the same interior-first list.
  const bool ordered = lazy_halo && dist.block_order != nullptr;
  const bool dyn = !FUSED && sched.work != nullptr;  // static run, then chunks from the shared counter
  const bool split = ordered && !dyn;                // interior run + strided ghost tail (the static multi-GPU order)
  const int n_first = split ? dist.n_interior : nblk;
  const int per_cta = dyn ? sched.static_blocks : (n_first + gridDim.x - 1) / gridDim.x;
  const int blk0 = min(n_first, static_cast<int>(blockIdx.x) * per_cta);
  const int n_run = min(n_first, blk0 + per_cta) - blk0;
  const int n_ghost_blk = split ? nblk - n_first : 0;
  const int n_tail = n_ghost_blk > static_cast<int>(blockIdx.x) ? (n_ghost_blk - 1 - blockIdx.x) / gridDim.x + 1 : 0;
  const int n_items = n_run + n_tail;  // (dyn: only the static run; the dynamic part ends with a sentinel)
  const int ring0 = FUSED ? iter * n_items : 0;  // ring position of this call's first item
  const int dyn_base = dyn ? min(nblk, static_cast<int>(gridDim.x) * per_cta) : nblk;

  if (t >= SPMV_CONSUMERS) {
    // ---- producer: one thread ----
    if (t == SPMV_CONSUMERS && (n_items > 0 || dyn)) {
      const unsigned long long pol = l2_evict_first_policy();
      bool halo_pending = lazy_halo && !FUSED;
      int s_next = 0, blk_prev = -2;
      int dyn_next = 0, dyn_end = 0;  // current chunk of the dynamic part
      for (int it = 0; dyn || it < n_items; ++it) {
        int blk = blk0 + it;
        if (split) blk = dist.block_order[it < n_run ? blk0 + it : n_first + blockIdx.x + (it - n_run) * gridDim.x];
        if (dyn && it >= n_items) {
          if (dyn_next == dyn_end) {
            dyn_next = dyn_base + static_cast<int>(atomicAdd(sched.work, static_cast<unsigned>(sched.chunk)));
            dyn_end = min(nblk, dyn_next + sched.chunk);
          }
          if (dyn_next >= nblk) {
            // out of work: sentinel for the consumers, then the completion ticket (the last producer re-arms the counters)
            const int stage = it % SPMV_STAGES;
            mbar_wait(&empty[stage], ((it / SPMV_STAGES) & 1) ^ 1u);
            spmv_stage<T, AUX, CAP>(base, stage).ptr[SPMV_DESC_SLOT] = -1;
            mbar_arrive(&full[stage]);
            if (atomicAdd(sched.done, 1u) == gridDim.x - 1) {
              *sched.work = 0u;
              *sched.done = 0u;
            }
            break;
          }
          blk = dyn_next++;
        }
        if (ordered && dyn) blk = dist.block_order[blk];  // position in the interior-first list -> block
        if (halo_pending && (split ? it >= n_run : dist.ghost_blocks[blk] != 0)) {
          dist_halo_wait(dist);
          halo_pending = false;
        }
        const int stage = (ring0 + it) % SPMV_STAGES;
        const unsigned phase = ((ring0 + it) / SPMV_STAGES) & 1;
        const int r0 = blk * R;
        const int nr = min(R, rows - r0);
        const int s = (blk == blk_prev + 1) ? s_next : indptr[r0];  // consecutive blocks: one indptr load per block
        const int e = indptr[r0 + nr];
        s_next = e;
        blk_prev = blk;
        if (!mbar_wait_b<FUSED>(&empty[stage], phase ^ 1u)) return;
        const SpmvStage<T> st = spmv_stage<T, AUX, CAP>(base, stage);
        if (ordered || dyn) st.ptr[SPMV_DESC_SLOT] = blk;  // published by the arrive below (release)
        const int a0 = s & ~3;
        const int cnt = e - a0;
        const unsigned ptr_bytes = static_cast<unsigned>(((nr + 1) * 4 + 15) & ~15);
        const unsigned aux_bytes = static_cast<unsigned>((nr * static_cast<int>(sizeof(T)) + 15) & ~15);
        if constexpr (AUX > 0) {
          // (expect_tx may follow the copies it accounts for: the phase cannot complete before this thread's arrive)
#pragma unroll
          for (int k = 0; k < AUX; ++k)
            tma_bulk_g2s(st.aux + k * SPMV_CONSUMERS, sink.aux_src(k) + r0, aux_bytes, &full[stage], pol);
        }
        if (cnt <= CAP && cnt > 0) {
          const int a0c = s & ~15;  // codes: one byte each, 16 to a TMA unit
          const unsigned ib = coded ? static_cast<unsigned>((e - a0c + 15) & ~15) : static_cast<unsigned>((cnt * 4 + 15) & ~15);
          const unsigned vb = static_cast<unsigned>((cnt * static_cast<int>(sizeof(T)) + 15) & ~15);
          mbar_expect_tx(&full[stage], ptr_bytes + ib + vb + AUX * aux_bytes);
          tma_bulk_g2s(st.ptr, indptr + r0, ptr_bytes, &full[stage], pol);
          if (coded) tma_bulk_g2s(st.idx, sched.codes + a0c, ib, &full[stage], pol);
          else tma_bulk_g2s(st.idx, indices + a0, ib, &full[stage], pol);
          tma_bulk_g2s(st.val, values + a0, vb, &full[stage], pol);
        } else {
          mbar_expect_tx(&full[stage], ptr_bytes + AUX * aux_bytes);
          tma_bulk_g2s(st.ptr, indptr + r0, ptr_bytes, &full[stage], pol);
        }
      }
    }
  } else {
    // ---- consumers ----
    const int lane = t & 31, warp = t >> 5;
    const int row_in_blk = t / LPR, q = t % LPR;
    if constexpr (!FUSED) {
      if (dist.amode) {
        // this kernel sends the halo of its own operand: the CTAs whose chunk of x holds a halo row store it into the
        // neighbours' ghost segments now; the NVLink round trip runs beside the interior blocks of every rank
        const int per = ((rows + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x) + 1023) & ~1023;
        const int lo = min(rows, static_cast<int>(blockIdx.x) * per);
        dist_halo_push_impl<T, false, SPMV_CONSUMERS>(dist, x, nullptr, lo, min(rows, lo + per), per);
      }
    }
    bool halo_pending = FUSED && dist.world > 1;
    for (int it = 0; dyn || it < n_items; ++it) {
      const int stage = (ring0 + it) % SPMV_STAGES;
      const unsigned phase = ((ring0 + it) / SPMV_STAGES) & 1;
      const SpmvStage<T> st = spmv_stage<T, AUX, CAP>(base, stage);
      if (!mbar_wait_b<FUSED>(&full[stage], phase)) return;
      const int blk = (ordered || dyn) ? st.ptr[SPMV_DESC_SLOT] : blk0 + it;
      if (dyn && blk < 0) break;  // the producer's sentinel
      if constexpr (FUSED) {
        // (uniform over the consumers: every thread evaluates the same block)
        if (halo_pending && (!lazy_halo || (ordered ? it >= n_run : dist.ghost_blocks[blk] != 0))) {
          if (t == 0) dist_halo_wait(dist);
          consumer_bar();
          halo_pending = false;
        }
      }
      const int r0 = blk * R;
      const int nr = min(R, rows - r0);
      // (coded instantiation: explicit ld.shared for the row pointers as well; the canonical kernel is left as measured)
      const int s = CODED ? lds_i32(st.ptr) : st.ptr[0], e = CODED ? lds_i32(st.ptr + nr) : st.ptr[nr];
      const int a0 = s & ~3;
      if (e - a0 <= CAP) {
        T result = T(0);
        typename Sink::Pre pre{};
        if (row_in_blk < nr) {
          if (q == 0) pre = sink.pre(r0 + row_in_blk, st.aux, row_in_blk);
          const int lo = (CODED ? lds_i32(st.ptr + row_in_blk) : st.ptr[row_in_blk]) - a0,
                    hi = (CODED ? lds_i32(st.ptr + row_in_blk + 1) : st.ptr[row_in_blk + 1]) - a0;
          if constexpr (CODED) {
            // one byte per entry: column = row + s_tab[code].  Same batch-of-eight shape as the canonical body (the eight
            // codes, then the eight table look-ups, then the eight gathers are issued together; the products are added in
            // Eigen's order), with explicit ld.shared -- the stage pointers are generic to the compiler and generic loads
            // of shared memory wait in the global-memory pipe (ncu, profiles/r02as: the dependent chain code -> offset ->
            // x -> fma was 65 % long-scoreboard stalls).  The coded stream moves 0.63 of the canonical bytes, so issue
            // slots, not HBM, are what this body has to watch.
            const unsigned char *code = reinterpret_cast<const unsigned char *>(st.idx) + (a0 - (s & ~15));
            const T *xr = x + (r0 + row_in_blk);
            T acc_a = T(0), acc_b = T(0);
            for (int k = lo; k < hi; k += 8) {
              int cd[8];
              T v[8], xv[8];
#pragma unroll
              // (the clamp keeps the eight look-ups on this row's own codes; dropping it -- the slots past the row are
              //  discarded anyway -- saved eight instructions but measured 0.154 instead of 0.142 ms stand-alone)
              for (int j = 0; j < 8; ++j) cd[j] = lds_u8(code + min(k + j, hi - 1));
#pragma unroll
              for (int j = 0; j < 8; ++j) cd[j] = s_tab[cd[j]];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const bool ok = k + j < hi;
                v[j] = ok ? lds_val(st.val + k + j) : T(0);
                xv[j] = ok ? ld_x<FUSED>(xr + cd[j]) : T(0);
              }
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                if (k + j < hi) {
                  if (ROW_MAJOR_ORDER) {
                    if ((j & 1) == 0) acc_a = mad_unfused(v[j], xv[j], acc_a);
                    else acc_b = fma_t(v[j], xv[j], acc_b);
                  } else {
                    acc_a = fma_t(v[j], xv[j], acc_a);
                  }
                }
              }
            }
            result = ROW_MAJOR_ORDER ? (T(0) + (acc_a + acc_b)) : acc_a;
          } else if (LPR == 1) {
            T acc_a = T(0), acc_b = T(0);
            for (int k = lo; k < hi; k += 8) {
              T v[8], xv[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const bool ok = k + j < hi;
                const int c = ok ? st.idx[k + j] : 0;
                v[j] = ok ? st.val[k + j] : T(0);
                xv[j] = ok ? ld_x<FUSED>(x + c) : T(0);
              }
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                if (k + j < hi) {
                  if (ROW_MAJOR_ORDER) {
                    // Eigen row-major processRow: entries 0,2,4.. of the row -> accumulator a, 1,3,5.. -> b.  The
                    // reference's shipped build (GCC 13, -O3 -march=x86-64-v3) leaves a as mul+add and fuses b;
                    // mirrored so that the result equals oracle/_ref bit for bit (tests/test_oracle_vs_eigen.py).
                    if ((j & 1) == 0) acc_a = mad_unfused(v[j], xv[j], acc_a);
                    else acc_b = fma_t(v[j], xv[j], acc_b);
                  } else {
                    // Eigen col-major scatter: y_i += a_ij * x_j for j ascending, one fused multiply-add each
                    acc_a = fma_t(v[j], xv[j], acc_a);
                  }
                }
              }
            }
            result = ROW_MAJOR_ORDER ? (T(0) + (acc_a + acc_b)) : acc_a;
          } else {
            T acc = T(0);
            for (int k = lo + q; k < hi; k += 4 * LPR) {
              T v[4], xv[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const bool ok = k + j * LPR < hi;
                const int c = ok ? st.idx[k + j * LPR] : 0;
                v[j] = ok ? st.val[k + j * LPR] : T(0);
                xv[j] = ok ? ld_x<FUSED>(x + c) : T(0);
              }
#pragma unroll
              for (int j = 0; j < 4; ++j) acc = fma_t(v[j], xv[j], acc);
            }
            result = acc;
          }
        }
        if (LPR > 1) {
#pragma unroll
          for (int o = LPR / 2; o > 0; o >>= 1) result += __shfl_xor_sync(FULL, result, o);
        }
        if (row_in_blk < nr && q == 0) {
          y[r0 + row_in_blk] = result;
          sink.row(r0 + row_in_blk, result, pre);
        }
      } else {
        // ---- oversized block: warp per row straight from global memory ----
        for (int r = warp; r < nr; r += SPMV_CONSUMERS / 32) {
          const int a = st.ptr[r], b = st.ptr[r + 1];
          T acc = T(0);
          for (int k = a + lane; k < b; k += 32) acc = fma_t(values[k], ld_x<FUSED>(x + indices[k]), acc);
          acc = warp_sum(acc);
          if (lane == 0) {
            y[r0 + r] = acc;
            sink.row(r0 + r, acc, sink.pre(r0 + r, st.aux, r));
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
    }
  }
}

// ---- stream path ---------------------------------------------------------------------------------------------------------
// Same persistent, warp-specialised TMA pipeline; the work unit is a plan block.  Consumers: (1) all 256 threads form the
// products val[k] * x[idx[k]] in place in shared memory -- perfectly balanced whatever the row lengths --, (2) after a
// consumer-only barrier the rows of the block are summed from shared memory, one row per thread (many short rows) or
// one row per warp (few long rows); a SEGMENT block reduces to one partial sum.  Not bit-identical to Eigen's order
// (held to the 1e-5 / 1e-12 per-row tolerance).
// FUSED: as for spmv_persistent (ring initialised by the caller, producer and consumers call independently once per CG
// iteration `iter`, halo wait on the consumer side, coherent gather).
template <typename T, typename Sink, bool FUSED = false>
__device__ __forceinline__ void spmv_stream(unsigned char *smem_raw, const int *__restrict__ indptr,
                                            const int *__restrict__ indices, const T *__restrict__ values,
                                            const T *x, T *y, Sink &sink, const SpmvPlanDev &plan, const DistDev &dist,
                                            int iter = 0) {
  unsigned char *base = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  auto *full = reinterpret_cast<unsigned long long *>(base + SPMV_STREAM_STAGES * spmv_stage_bytes<T>());
  auto *empty = full + SPMV_STREAM_STAGES;
  __shared__ unsigned s_last;
  const int t = threadIdx.x;
  static_assert(2 * SPMV_STREAM_STAGES * sizeof(unsigned long long) <= SPMV_HOT_OFFSET, "hot array overlaps the barriers");
  T *hot = reinterpret_cast<T *>(base + SPMV_STREAM_STAGES * spmv_stage_bytes<T>() + SPMV_HOT_OFFSET);
  const bool use_hot = !FUSED && plan.hot_n > 0;  // (the persistent CG kernel rewrites x between products: no cache there)
  if constexpr (!FUSED) spmv_ring_init<SPMV_STREAM_STAGES>(full, empty);
  if (dist.world > 1) {  // no per-block ghost map on this path: wait for the halo before the first gather
    if constexpr (FUSED) {
      if (t < SPMV_CONSUMERS) {
        if (t == 0) dist_halo_wait(dist);
        consumer_bar();
      }
    } else {
      if (t == 0) dist_halo_wait(dist);
      __syncthreads();
    }
  }
  const int per_cta = (plan.nblk + gridDim.x - 1) / gridDim.x;
  const int blk0 = min(plan.nblk, static_cast<int>(blockIdx.x) * per_cta);
  const int blk1 = min(plan.nblk, blk0 + per_cta);
  const int ring0 = FUSED ? iter * (blk1 - blk0) : 0;  // ring position of this call's first block
  if constexpr (!FUSED) {
    if (use_hot) {
      if (blk0 < blk1)
        for (int i = t; i < plan.hot_n; i += SPMV_THREADS) hot[i] = ld_x<FUSED>(x + __ldg(plan.hot_cols + i));
      __syncthreads();
    }
    if (use_hot) indices = plan.hot_idx;
  }

  if (t >= SPMV_CONSUMERS) {
    if (t == SPMV_CONSUMERS && blk0 < blk1) {
      const unsigned long long pol = l2_evict_first_policy();
      int4 next = plan.blocks[blk0];
      for (int blk = blk0, it = 0; blk < blk1; ++blk, ++it) {
        const int4 d = next;
        if (blk + 1 < blk1) next = plan.blocks[blk + 1];
        const int stage = (ring0 + it) % SPMV_STREAM_STAGES;
        const unsigned phase = ((ring0 + it) / SPMV_STREAM_STAGES) & 1;
        if (!mbar_wait_b<FUSED>(&empty[stage], phase ^ 1u)) return;
        const SpmvStage<T> st = spmv_stage<T>(base, stage);
        *reinterpret_cast<int4 *>(st.ptr + SPMV_DESC_SLOT) = d;  // published by the arrive below (release)
        const int a0 = d.z & ~3;
        const int cnt = d.w - a0;
        unsigned bytes = 0;
        const unsigned ib = static_cast<unsigned>((cnt * 4 + 15) & ~15);
        const unsigned vb = static_cast<unsigned>((cnt * static_cast<int>(sizeof(T)) + 15) & ~15);
        unsigned pb = 0;
        if (d.y > 0) pb = static_cast<unsigned>((((d.x & 3) + d.y + 1) * 4 + 15) & ~15);
        if (cnt > 0) bytes += ib + vb;
        bytes += pb;
        if (bytes == 0) {
          mbar_arrive(&full[stage]);
          continue;
        }
        mbar_expect_tx(&full[stage], bytes);
        if (pb) tma_bulk_g2s(st.ptr, indptr + (d.x & ~3), pb, &full[stage], pol);
        if (cnt > 0) {
          tma_bulk_g2s(st.idx, indices + a0, ib, &full[stage], pol);
          tma_bulk_g2s(st.val, values + a0, vb, &full[stage], pol);
        }
      }
    }
  } else {
    const int lane = t & 31, warp = t >> 5;
    for (int blk = blk0, it = 0; blk < blk1; ++blk, ++it) {
      const int stage = (ring0 + it) % SPMV_STREAM_STAGES;
      const unsigned phase = ((ring0 + it) / SPMV_STREAM_STAGES) & 1;
      const SpmvStage<T> st = spmv_stage<T>(base, stage);
      if (!mbar_wait_b<FUSED>(&full[stage], phase)) return;
      const int4 d = lds_i32x4(st.ptr + SPMV_DESC_SLOT);
      const int a0 = d.z & ~3;
      const int lo = d.z - a0, cnt = d.w - a0;
      // (1) products, in place
      // (8 independent gathers in flight per thread: the x of an irregular matrix comes from L2, not L1.  The eight column
      // ids are read first, the eight gathers issued back to back, the eight products stored last.)
      for (int k = lo + t; k < cnt; k += 8 * SPMV_CONSUMERS) {
        int c[8];
        T xv[8], av[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) c[j] = lds_i32(st.idx + min(k + j * SPMV_CONSUMERS, cnt - 1));
#pragma unroll
        for (int j = 0; j < 8; ++j) xv[j] = (!FUSED && c[j] < 0) ? lds_val(hot + (~c[j])) : ld_x<FUSED>(x + c[j]);
#pragma unroll
        for (int j = 0; j < 8; ++j) av[j] = lds_val(st.val + min(k + j * SPMV_CONSUMERS, cnt - 1));
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int kk = k + j * SPMV_CONSUMERS;
          if (kk < cnt) sts_val(st.val + kk, av[j] * xv[j]);
        }
      }
      consumer_bar();
      // (2) row sums
      if (d.y > 0) {
        const int nr = d.y, poff = d.x & 3;
        if (nr > SPMV_STREAM_THREAD_ROWS) {
          if (t < nr) {
            const int b = lds_i32(st.ptr + poff + t) - a0, e = lds_i32(st.ptr + poff + t + 1) - a0;
            T acc = T(0);
            for (int k = b; k < e; ++k) acc += lds_val(st.val + k);
            y[d.x + t] = acc;
            sink.row(d.x + t, acc, sink.pre(d.x + t, static_cast<const T *>(nullptr), 0));
          }
        } else {
          for (int r = warp; r < nr; r += SPMV_CONSUMERS / 32) {
            const int b = lds_i32(st.ptr + poff + r) - a0, e = lds_i32(st.ptr + poff + r + 1) - a0;
            T acc = T(0);
            for (int k = b + lane; k < e; k += 32) acc += lds_val(st.val + k);
            acc = warp_sum(acc);
            if (lane == 0) {
              y[d.x + r] = acc;
              sink.row(d.x + r, acc, sink.pre(d.x + r, static_cast<const T *>(nullptr), 0));
            }
          }
        }
      } else {
        // one segment of a long row: block-wide sum -> partials[segment]
        T acc = T(0);
        for (int k = lo + t; k < cnt; k += SPMV_CONSUMERS) acc += lds_val(st.val + k);
        acc = warp_sum(acc);
        T *red = reinterpret_cast<T *>(st.ptr);  // the indptr slice is unused for a segment
        if (lane == 0) sts_val(red + warp, acc);
        consumer_bar();
        if (t == 0) {
          T tot = T(0);
          for (int w = 0; w < SPMV_CONSUMERS / 32; ++w) tot += lds_val(red + w);
          static_cast<T *>(plan.partials)[-d.y - 1] = tot;
          sink.segment(d.x, tot);
        }
      }
      // the stage was written through the generic proxy (products); order that before the next TMA fill
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
    }
    // long rows: the last CTA to get here adds each row's segment sums in segment order
    if (plan.n_long > 0) {
      consumer_bar();
      if (t == 0) {
        __threadfence();
        s_last = (atomicAdd(plan.ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
      }
      consumer_bar();
      if (s_last) {
        __threadfence();
        const T *part = static_cast<const T *>(plan.partials);
        for (int i = t; i < plan.n_long; i += SPMV_CONSUMERS) {
          const int2 a = plan.long_rows[i];
          const int end = plan.long_rows[i + 1].y;
          T acc = T(0);
          for (int sgm = a.y; sgm < end; ++sgm) acc += __ldcg(part + sgm);
          y[a.x] = acc;
        }
        if (t == 0) *plan.ticket = 0u;
      }
    }
  }
}

}  // namespace qdsp
