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

namespace qdsp {

constexpr int SPMV_CONSUMERS = 256;                 // consumer threads per CTA (8 warps)
constexpr int SPMV_THREADS = SPMV_CONSUMERS + 32;   // + the producer warp
constexpr int SPMV_CAP = 2048;                      // non-zeros per stage
constexpr int SPMV_STAGES = 3;
constexpr int SPMV_PTR_SLOTS = 260;                 // R + 1 <= 257, rounded up to a multiple of 4

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

template <typename T>
__host__ __device__ constexpr size_t spmv_stage_bytes() {
  return SPMV_CAP * sizeof(int) + SPMV_CAP * sizeof(T) + SPMV_PTR_SLOTS * sizeof(int);
}
template <typename T>
__host__ __device__ constexpr size_t spmv_smem_bytes() {
  return SPMV_STAGES * spmv_stage_bytes<T>() + 2 * SPMV_STAGES * sizeof(unsigned long long) + 128 /*alignment slack*/;
}

// ---- mbarrier / TMA bulk-copy primitives (sm_90+; SASS: SYNCS.*, UBLKCP) -------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ unsigned long long l2_evict_first_policy() {
  unsigned long long pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16; completes on `bar`
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar,
                                             unsigned long long policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

template <typename T>
struct SpmvStage {
  int *idx;
  T *val;
  int *ptr;
};
template <typename T>
__device__ __forceinline__ SpmvStage<T> spmv_stage(unsigned char *base, int s) {
  unsigned char *p = base + static_cast<size_t>(s) * spmv_stage_bytes<T>();
  SpmvStage<T> st;
  st.val = reinterpret_cast<T *>(p);
  st.idx = reinterpret_cast<int *>(p + SPMV_CAP * sizeof(T));
  st.ptr = reinterpret_cast<int *>(p + SPMV_CAP * sizeof(T) + SPMV_CAP * sizeof(int));
  return st;
}

// The whole SpMV of one persistent CTA.  LPR = lanes per row.  Returns, per consumer thread, the sum over the rows it
// owned of xdot[row] * y[row] in double (0 when xdot == nullptr); producer-warp threads return 0.
// Multi-GPU (dist.world > 1): x = [owned | ghosts]; the ghosts are stored by the neighbour ranks.  Only row blocks
// flagged in dist.ghost_blocks read a ghost, so the producer waits for the halo flags right before it hands the first
// such block to the consumers (its acquire reaches them through the stage's mbarrier) and interior blocks start at
// once.  The ghost segment begins on a 128-byte boundary, so no L1 line holds both owned and ghost entries.
// `indptr`, `indices`, `values` must be readable 16 bytes past their ends (the matrices of this library carry 64).
template <typename T, int LPR, bool ROW_MAJOR_ORDER>
__device__ __forceinline__ double spmv_persistent(unsigned char *smem_raw, int rows, const int *__restrict__ indptr,
                                                  const int *__restrict__ indices, const T *__restrict__ values,
                                                  const T *__restrict__ x, T *__restrict__ y,
                                                  const T *__restrict__ xdot, const DistDev &dist) {
  constexpr int R = SPMV_CONSUMERS / LPR;  // rows per block
  unsigned char *base = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  auto *full = reinterpret_cast<unsigned long long *>(base + SPMV_STAGES * spmv_stage_bytes<T>());
  auto *empty = full + SPMV_STAGES;
  const int t = threadIdx.x;
  if (t == 0) {
    for (int s = 0; s < SPMV_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], SPMV_CONSUMERS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // no per-block ghost map for this block size: every CTA waits for the halo before its first gather
  const bool lazy_halo = dist.world > 1 && dist.ghost_blocks != nullptr && dist.ghost_block_rows == R;
  if (dist.world > 1 && !lazy_halo) {
    if (t == 0) dist_halo_wait(dist);
    __syncthreads();
  }

  const int nblk = (rows + R - 1) / R;
  const int per_cta = (nblk + gridDim.x - 1) / gridDim.x;
  const int blk0 = blockIdx.x * per_cta;
  const int blk1 = min(nblk, blk0 + per_cta);
  double dot = 0.0;

  if (t >= SPMV_CONSUMERS) {
    // ---- producer: one thread ----
    if (t == SPMV_CONSUMERS && blk0 < blk1) {
      const unsigned long long pol = l2_evict_first_policy();
      int s_next = indptr[blk0 * R];
      bool halo_pending = lazy_halo;
      for (int blk = blk0, it = 0; blk < blk1; ++blk, ++it) {
        if (halo_pending && dist.ghost_blocks[blk]) {
          dist_halo_wait(dist);
          halo_pending = false;
        }
        const int stage = it % SPMV_STAGES;
        const unsigned phase = (it / SPMV_STAGES) & 1;
        const int r0 = blk * R;
        const int nr = min(R, rows - r0);
        const int s = s_next;
        const int e = indptr[r0 + nr];
        s_next = e;
        mbar_wait(&empty[stage], phase ^ 1u);
        const SpmvStage<T> st = spmv_stage<T>(base, stage);
        const int a0 = s & ~3;
        const int cnt = e - a0;
        const unsigned ptr_bytes = static_cast<unsigned>(((nr + 1) * 4 + 15) & ~15);
        if (cnt <= SPMV_CAP && cnt > 0) {
          const unsigned ib = static_cast<unsigned>((cnt * 4 + 15) & ~15);
          const unsigned vb = static_cast<unsigned>((cnt * static_cast<int>(sizeof(T)) + 15) & ~15);
          mbar_expect_tx(&full[stage], ptr_bytes + ib + vb);
          tma_bulk_g2s(st.ptr, indptr + r0, ptr_bytes, &full[stage], pol);
          tma_bulk_g2s(st.idx, indices + a0, ib, &full[stage], pol);
          tma_bulk_g2s(st.val, values + a0, vb, &full[stage], pol);
        } else {
          mbar_expect_tx(&full[stage], ptr_bytes);
          tma_bulk_g2s(st.ptr, indptr + r0, ptr_bytes, &full[stage], pol);
        }
      }
    }
  } else {
    // ---- consumers ----
    const int lane = t & 31, warp = t >> 5;
    const int row_in_blk = t / LPR, q = t % LPR;
    for (int blk = blk0, it = 0; blk < blk1; ++blk, ++it) {
      const int stage = it % SPMV_STAGES;
      const unsigned phase = (it / SPMV_STAGES) & 1;
      const int r0 = blk * R;
      const int nr = min(R, rows - r0);
      const SpmvStage<T> st = spmv_stage<T>(base, stage);
      mbar_wait(&full[stage], phase);
      const int s = st.ptr[0], e = st.ptr[nr];
      const int a0 = s & ~3;
      if (e - a0 <= SPMV_CAP) {
        T result = T(0);
        if (row_in_blk < nr) {
          const int lo = st.ptr[row_in_blk] - a0, hi = st.ptr[row_in_blk + 1] - a0;
          if (LPR == 1) {
            T acc_a = T(0), acc_b = T(0);
            for (int k = lo; k < hi; k += 8) {
              T v[8], xv[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const bool ok = k + j < hi;
                const int c = ok ? st.idx[k + j] : 0;
                v[j] = ok ? st.val[k + j] : T(0);
                xv[j] = ok ? __ldg(x + c) : T(0);
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
                xv[j] = ok ? __ldg(x + c) : T(0);
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
          if (xdot) dot += static_cast<double>(xdot[r0 + row_in_blk] * result);
        }
      } else {
        // ---- oversized block: warp per row straight from global memory ----
        for (int r = warp; r < nr; r += SPMV_CONSUMERS / 32) {
          const int a = st.ptr[r], b = st.ptr[r + 1];
          T acc = T(0);
          for (int k = a + lane; k < b; k += 32) acc = fma_t(values[k], __ldg(x + indices[k]), acc);
          acc = warp_sum(acc);
          if (lane == 0) {
            y[r0 + r] = acc;
            if (xdot) dot += static_cast<double>(xdot[r0 + r] * acc);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
    }
  }
  return dot;
}

}  // namespace qdsp
