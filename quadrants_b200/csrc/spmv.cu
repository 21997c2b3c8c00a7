// spmv.cu -- CsrMatrix: SpMV launcher and the small host-visible accessors of CuSparseMatrix
// (quadrants/program/sparse_matrix.cpp:654-780).
#include <cstdlib>
#include <cstring>
#include <string>
#include <fstream>
#include <algorithm>
#include <climits>
#include <vector>

#include "host.hpp"
#include "spmv_kernels.cuh"

namespace qdsp {

std::atomic<int64_t> g_launch_count{0};

const DeviceProps &device_props() {
  static thread_local DeviceProps cached;
  int dev = 0;
  QDSP_CUDA(cudaGetDevice(&dev));
  if (cached.device != dev) {
    cudaDeviceProp p;
    QDSP_CUDA(cudaGetDeviceProperties(&p, dev));
    cached.device = dev;
    cached.sm_count = p.multiProcessorCount;
    cached.cc_major = p.major;
    cached.cc_minor = p.minor;
    cached.l2_bytes = static_cast<size_t>(p.l2CacheSize);
    cached.hbm_bytes = p.totalGlobalMem;
  }
  return cached;
}

template <typename T, int LPR, bool ROW_MAJOR_ORDER, bool CODED = false>
__global__ void __launch_bounds__(SPMV_THREADS)
spmv_kernel(int rows, const int *__restrict__ indptr, const int *__restrict__ indices, const T *__restrict__ values,
            const T *__restrict__ x, T *__restrict__ y, DistDev d, SpmvSched sched) {
  extern __shared__ unsigned char smem_raw[];
  NoSink sink;
  spmv_persistent<T, LPR, ROW_MAJOR_ORDER, NoSink, false, CODED>(smem_raw, rows, indptr, indices, values, x, y, sink, d, 0, sched);
}

// Static run per CTA = 3/4 of an even share; the rest is handed out two blocks at a time.  Small launches (< 8 blocks per
// CTA) stay static: nothing to balance.  QDSP_SPMV_DYNAMIC=0 switches the dynamic tail off (measurement aid).
SpmvSched spmv_sched_for(int nblk, int grid, unsigned *counters) {
  SpmvSched s;
  static const bool on = [] { const char *e = std::getenv("QDSP_SPMV_DYNAMIC"); return !(e && e[0] == '0'); }();
  if (!on || counters == nullptr || grid <= 0 || nblk < 8 * grid) return s;
  s.work = counters;
  s.done = counters + 1;
  s.static_blocks = static_cast<int>((static_cast<int64_t>(nblk) / grid) * 3 / 4);
  s.chunk = 2;
  return s;
}

template <typename T>
__global__ void __launch_bounds__(SPMV_THREADS)
spmv_stream_kernel(const int *__restrict__ indptr, const int *__restrict__ indices, const T *__restrict__ values,
                   const T *__restrict__ x, T *__restrict__ y, SpmvPlanDev plan, DistDev d) {
  extern __shared__ unsigned char smem_raw[];
  NoSink sink;
  spmv_stream<T>(smem_raw, indptr, indices, values, x, y, sink, plan, d);
}

// out[0] = longest row; out[1] = the largest stage footprint (non-zeros, counted from the 16-byte-aligned start) of a
// uniform block of R rows
__global__ void __launch_bounds__(256) row_stats_kernel(int rows, int R, const int *__restrict__ indptr, int *__restrict__ out) {
  int m = 0, mb = 0;
  for (int r = blockIdx.x * 256 + threadIdx.x; r < rows; r += gridDim.x * 256) {
    m = max(m, indptr[r + 1] - indptr[r]);
    if (r % R == 0) mb = max(mb, indptr[min(rows, r + R)] - (indptr[r] & ~3));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m = max(m, __shfl_xor_sync(FULL, m, o));
    mb = max(mb, __shfl_xor_sync(FULL, mb, o));
  }
  if (lane_id() == 0) {
    if (m > 0) atomicMax(out, m);
    if (mb > 0) atomicMax(out + 1, mb);
  }
}

// ---- diagonal-coded indices of the uniform path (spmv_kernels.cuh SpmvSched) ----------------------------------------------------
constexpr int CODE_SLOTS = 1024;          // open-addressing set of distinct (col - row); more than 256 entries = give up
constexpr int CODE_EMPTY = INT_MIN;
__global__ void __launch_bounds__(256) code_collect_kernel(int rows, const int *__restrict__ indptr, const int *__restrict__ idx,
                                                           int *__restrict__ set, int *__restrict__ count) {
  for (int r = blockIdx.x * 256 + threadIdx.x; r < rows; r += gridDim.x * 256) {
    if (*reinterpret_cast<volatile int *>(count) > 256) return;  // too many distinct offsets: the host gives up anyway
    for (int k = indptr[r]; k < indptr[r + 1]; ++k) {
      const int off = idx[k] - r;
      unsigned h = (static_cast<unsigned>(off) * 2654435761u) >> 22;  // 10 bits
      for (int probe = 0; probe < CODE_SLOTS; ++probe, h = (h + 1) & (CODE_SLOTS - 1)) {
        const int cur = set[h];
        if (cur == off) break;
        if (cur == CODE_EMPTY) {
          const int old = atomicCAS(set + h, CODE_EMPTY, off);
          if (old == CODE_EMPTY) { atomicAdd(count, 1); break; }
          if (old == off) break;
        }
      }
    }
  }
}
__global__ void __launch_bounds__(256) code_encode_kernel(int rows, const int *__restrict__ indptr, const int *__restrict__ idx,
                                                          const int *__restrict__ tab, int ntab, unsigned char *__restrict__ codes,
                                                          int *__restrict__ failed) {
  __shared__ int s_tab[256];
  for (int i = threadIdx.x; i < ntab; i += 256) s_tab[i] = tab[i];
  __syncthreads();
  for (int r = blockIdx.x * 256 + threadIdx.x; r < rows; r += gridDim.x * 256) {
    for (int k = indptr[r]; k < indptr[r + 1]; ++k) {
      const int off = idx[k] - r;
      int lo = 0, hi = ntab;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s_tab[mid] < off) lo = mid + 1;
        else hi = mid;
      }
      if (lo < ntab && s_tab[lo] == off) codes[k] = static_cast<unsigned char>(lo);
      else *failed = 1;
    }
  }
}

// Makes the one-byte codes when the matrix has <= 256 distinct (col - row).  QDSP_SPMV_COMPRESS=0 switches it off
// process-wide; qdsp_csr_spmv_config per matrix.
static void spmv_plan_codes(const CsrMatrix &A, cudaStream_t stream) {
  A.spmv_ntab = 0;
  static const bool env_on = [] { const char *e = std::getenv("QDSP_SPMV_COMPRESS"); return !(e && e[0] == '0'); }();
  const bool on = A.spmv_compress < 0 ? env_on : A.spmv_compress != 0;
  if (!on || A.nnz == 0 || A.spmv_lpr != 1) return;
  if (A.spmv_ntab_built > 0) {  // made before and switched off in between: the arrays are still there (see spmv_set_compress)
    A.spmv_ntab = A.spmv_ntab_built;
    return;
  }
  if (A.spmv_ntab_built < 0) return;  // tried before: the matrix has too many distinct offsets
  A.spmv_ntab_built = -1;
  DeviceBuffer set(sizeof(int) * (CODE_SLOTS + 2));
  std::vector<int> h(CODE_SLOTS + 2, CODE_EMPTY);
  h[CODE_SLOTS] = h[CODE_SLOTS + 1] = 0;
  QDSP_CUDA(cudaMemcpyAsync(set.get(), h.data(), sizeof(int) * h.size(), cudaMemcpyHostToDevice, stream));
  const int grid = std::max(1, std::min(device_props().sm_count * 8, ceil_div(A.rows, 256)));
  code_collect_kernel<<<grid, 256, 0, stream>>>(A.rows, A.indptr.get<int>(), A.indices.get<int>(), set.get<int>(),
                                                set.get<int>() + CODE_SLOTS);
  QDSP_LAUNCH_CHECK();
  QDSP_CUDA(cudaMemcpyAsync(h.data(), set.get(), sizeof(int) * h.size(), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  if (h[CODE_SLOTS] > 256 || h[CODE_SLOTS] <= 0) return;
  std::vector<int> tab;
  for (int i = 0; i < CODE_SLOTS; ++i)
    if (h[i] != CODE_EMPTY) tab.push_back(h[i]);
  std::sort(tab.begin(), tab.end());
  if (tab.size() > 256) return;
  DeviceBuffer d_tab(sizeof(int) * 256), d_codes(static_cast<size_t>(A.nnz) + 64);  // TMA reads whole 16-byte units
  QDSP_CUDA(cudaMemsetAsync(d_codes.get(), 0, d_codes.bytes(), stream));
  QDSP_CUDA(cudaMemcpyAsync(d_tab.get(), tab.data(), sizeof(int) * tab.size(), cudaMemcpyHostToDevice, stream));
  int *failed = set.get<int>() + CODE_SLOTS + 1;
  code_encode_kernel<<<grid, 256, 0, stream>>>(A.rows, A.indptr.get<int>(), A.indices.get<int>(), d_tab.get<int>(),
                                               static_cast<int>(tab.size()), d_codes.get<unsigned char>(), failed);
  QDSP_LAUNCH_CHECK();
  int bad = 0;
  QDSP_CUDA(cudaMemcpyAsync(&bad, failed, sizeof(int), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  if (bad) return;  // (cannot happen: the set holds every offset)
  A.spmv_code_tab = std::move(d_tab);
  A.spmv_codes = std::move(d_codes);
  A.spmv_ntab = A.spmv_ntab_built = static_cast<int>(tab.size());
}

// qdsp_csr_spmv_config.  Only the choice "codes or canonical indices" changes: the rest of the plan and, above all, the
// code arrays themselves stay where they are -- CG solvers made earlier hold their addresses in captured graphs and keep
// running the path they were created with.
void spmv_set_compress(const CsrMatrix &A, int compress, cudaStream_t stream) {
  A.spmv_compress = compress < 0 ? -1 : (compress ? 1 : 0);
  if (A.spmv_mode == 1) spmv_plan_codes(A, stream);
}

void spmv_attach_codes(const CsrMatrix &A, SpmvSched &sched) {
  if (A.spmv_ntab > 0) {
    sched.codes = A.spmv_codes.get<unsigned char>();
    sched.tab = A.spmv_code_tab.get<int>();
    sched.ntab = A.spmv_ntab;
  }
}

// ---- hot columns of the stream path (spmv_kernels.cuh) ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) col_hist_kernel(int64_t nnz, int cols, const int *__restrict__ idx, int *__restrict__ counts) {
  for (int64_t k = blockIdx.x * int64_t{256} + threadIdx.x; k < nnz; k += static_cast<int64_t>(gridDim.x) * 256) {
    const int c = idx[k];
    if (c >= 0 && c < cols) atomicAdd(counts + c, 1);
  }
}
__global__ void __launch_bounds__(256) hot_remap_kernel(int64_t nnz, int cols, const int *__restrict__ idx,
                                                        const int *__restrict__ slot_of, int *__restrict__ out) {
  for (int64_t k = blockIdx.x * int64_t{256} + threadIdx.x; k < nnz; k += static_cast<int64_t>(gridDim.x) * 256) {
    const int c = idx[k];
    const int s = (c >= 0 && c < cols) ? slot_of[c] : -1;
    out[k] = s >= 0 ? ~s : c;
  }
}

// Picks the K = SPMV_HOT_BYTES / sizeof(T) most referenced columns (ties: lower column id) when they carry at least 15 %
// of the non-zeros.  QDSP_SPMV_HOT=0 switches the cache off (measurement aid).
static void spmv_plan_hot(const CsrMatrix &A, cudaStream_t stream) {
  A.spmv_plan_hot_n = 0;
  A.spmv_plan_hot_coverage = 0.0;
  static const bool on = [] { const char *e = std::getenv("QDSP_SPMV_HOT"); return !(e && e[0] == '0'); }();
  if (!on || A.nnz == 0 || A.cols == 0) return;
  const int K = std::min<int>(A.cols, SPMV_HOT_BYTES / static_cast<int>(A.value_bytes()));
  DeviceBuffer counts(sizeof(int) * static_cast<size_t>(A.cols));
  QDSP_CUDA(cudaMemsetAsync(counts.get(), 0, counts.bytes(), stream));
  const int grid = std::max(1, std::min(device_props().sm_count * 8, ceil_div(A.nnz, 256)));
  col_hist_kernel<<<grid, 256, 0, stream>>>(A.nnz, A.cols, A.indices.get<int>(), counts.get<int>());
  QDSP_LAUNCH_CHECK();
  std::vector<int> h(A.cols);
  QDSP_CUDA(cudaMemcpyAsync(h.data(), counts.get(), sizeof(int) * h.size(), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  std::vector<int> order(A.cols);
  for (int i = 0; i < A.cols; ++i) order[i] = i;
  auto hotter = [&](int a, int b) { return h[a] != h[b] ? h[a] > h[b] : a < b; };
  if (K < A.cols) std::nth_element(order.begin(), order.begin() + K, order.end(), hotter);
  order.resize(K);
  std::sort(order.begin(), order.end());  // ascending ids: the per-CTA load of the hot x entries walks x forwards
  int64_t covered = 0;
  for (int c : order) covered += h[c];
  A.spmv_plan_hot_coverage = static_cast<double>(covered) / static_cast<double>(A.nnz);
  if (A.spmv_plan_hot_coverage < 0.15) return;
  std::vector<int> slot_of(A.cols, -1);
  for (int s = 0; s < K; ++s) slot_of[order[s]] = s;
  DeviceBuffer d_slot(sizeof(int) * static_cast<size_t>(A.cols));
  A.spmv_plan_hot_cols.alloc(sizeof(int) * static_cast<size_t>(K));
  A.spmv_plan_hot_idx.alloc(sizeof(int32_t) * static_cast<size_t>(A.nnz) + 64);  // TMA reads whole 16-byte units
  QDSP_CUDA(cudaMemcpyAsync(d_slot.get(), slot_of.data(), sizeof(int) * slot_of.size(), cudaMemcpyHostToDevice, stream));
  QDSP_CUDA(cudaMemcpyAsync(A.spmv_plan_hot_cols.get(), order.data(), sizeof(int) * order.size(), cudaMemcpyHostToDevice, stream));
  QDSP_CUDA(cudaMemsetAsync(A.spmv_plan_hot_idx.get(), 0, A.spmv_plan_hot_idx.bytes(), stream));
  hot_remap_kernel<<<grid, 256, 0, stream>>>(A.nnz, A.cols, A.indices.get<int>(), d_slot.get<int>(), A.spmv_plan_hot_idx.get<int>());
  QDSP_LAUNCH_CHECK();
  QDSP_CUDA(cudaStreamSynchronize(stream));
  A.spmv_plan_hot_n = K;
}

// ---- plan ----------------------------------------------------------------------------------------------------------------
// Pure host: cuts the rows into stream-path blocks.  blocks: {row0, nrows | -(segment + 1), nz0, nz1};
// long_rows: {row, first segment} + sentinel.
void spmv_plan_host(const int32_t *indptr, int rows, std::vector<int4> &blocks, std::vector<int2> &long_rows,
                    int &n_segments) {
  blocks.clear();
  long_rows.clear();
  n_segments = 0;
  int r = 0;
  while (r < rows) {
    const int len = indptr[r + 1] - indptr[r];
    if (len > SPMV_STREAM_CAP) {
      long_rows.push_back(make_int2(r, n_segments));
      for (int s = indptr[r]; s < indptr[r + 1];) {
        // segment starts stay 4-aligned relative to nothing in particular: the kernel copies from nz0 & ~3
        const int e = std::min(indptr[r + 1], (s & ~3) + SPMV_STREAM_CAP);
        blocks.push_back(make_int4(r, -(n_segments + 1), s, e));
        ++n_segments;
        s = e;
      }
      ++r;
      continue;
    }
    const int r0 = r;
    const int a0 = indptr[r0] & ~3;
    while (r < rows && r - r0 < SPMV_CONSUMERS) {
      const int l = indptr[r + 1] - indptr[r];
      if (l > SPMV_STREAM_CAP || indptr[r + 1] - a0 > SPMV_CAP) break;
      ++r;
    }
    blocks.push_back(make_int4(r0, r - r0, indptr[r0], indptr[r]));
  }
  long_rows.push_back(make_int2(rows, n_segments));
}

void spmv_prepare(const CsrMatrix &A, cudaStream_t stream) {
  if (A.spmv_mode != 0 || A.rows == 0) return;
  A.spmv_ntab = A.spmv_ntab_built = 0;  // a new plan (first use, or the column renumbering of qdsp_dist_localize): codes are made afresh
  // uniform row blocks (bit-exact Eigen order when one lane per row) iff no block of R rows overflows a stage
  const int L = spmv_lanes_per_row(A.nnz, A.rows);
  DeviceBuffer mx(2 * sizeof(int));
  QDSP_CUDA(cudaMemsetAsync(mx.get(), 0, 2 * sizeof(int), stream));
  const int g = std::max(1, std::min(ceil_div(A.rows, 256), device_props().sm_count * 8));
  row_stats_kernel<<<g, 256, 0, stream>>>(A.rows, SPMV_CONSUMERS / L, A.indptr.get<int>(), mx.get<int>());
  QDSP_LAUNCH_CHECK();
  int stats[2] = {0, 0};
  QDSP_CUDA(cudaMemcpyAsync(stats, mx.get(), 2 * sizeof(int), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  A.spmv_max_row_len = stats[0];
  A.spmv_max_block_nnz = stats[1];
  const char *force = std::getenv("QDSP_SPMV_PATH");  // "uniform" | "stream": testing aid
  const bool uniform = force ? std::string(force) != "stream" : stats[1] <= SPMV_CAP;
  A.spmv_lpr = L;
  if (uniform) {
    A.spmv_sched_buf.alloc(16 * 2 * sizeof(unsigned));
    QDSP_CUDA(cudaMemsetAsync(A.spmv_sched_buf.get(), 0, A.spmv_sched_buf.bytes(), stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
    A.spmv_mode = 1;
    spmv_plan_codes(A, stream);
    return;
  }
  std::vector<int32_t> ip(static_cast<size_t>(A.rows) + 1);
  QDSP_CUDA(cudaMemcpyAsync(ip.data(), A.indptr.get(), sizeof(int32_t) * ip.size(), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  std::vector<int4> blocks;
  std::vector<int2> long_rows;
  int nseg = 0;
  spmv_plan_host(ip.data(), A.rows, blocks, long_rows, nseg);
  A.spmv_plan_blocks.alloc(sizeof(int4) * blocks.size());
  A.spmv_plan_long.alloc(sizeof(int2) * long_rows.size());
  A.spmv_plan_partials.alloc(A.value_bytes() * static_cast<size_t>(std::max(1, nseg)));
  A.spmv_plan_ticket.alloc(sizeof(unsigned));
  QDSP_CUDA(cudaMemcpyAsync(A.spmv_plan_blocks.get(), blocks.data(), sizeof(int4) * blocks.size(), cudaMemcpyHostToDevice, stream));
  QDSP_CUDA(cudaMemcpyAsync(A.spmv_plan_long.get(), long_rows.data(), sizeof(int2) * long_rows.size(), cudaMemcpyHostToDevice, stream));
  QDSP_CUDA(cudaMemsetAsync(A.spmv_plan_ticket.get(), 0, sizeof(unsigned), stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  A.spmv_plan_nblk = static_cast<int>(blocks.size());
  A.spmv_plan_nlong = static_cast<int>(long_rows.size()) - 1;
  A.spmv_mode = 2;
  spmv_plan_hot(A, stream);
}

SpmvPlanDev spmv_plan_dev(const CsrMatrix &A, bool allow_hot) {
  SpmvPlanDev p;
  if (allow_hot && A.spmv_plan_hot_n > 0) {
    p.hot_n = A.spmv_plan_hot_n;
    p.hot_cols = A.spmv_plan_hot_cols.get<int>();
    p.hot_idx = A.spmv_plan_hot_idx.get<int>();
  }
  p.blocks = A.spmv_plan_blocks.get<int4>();
  p.nblk = A.spmv_plan_nblk;
  p.n_long = A.spmv_plan_nlong;
  p.long_rows = A.spmv_plan_long.get<int2>();
  p.partials = A.spmv_plan_partials.get();
  p.ticket = A.spmv_plan_ticket.get<unsigned>();
  return p;
}

namespace {

template <typename T>
void launch_spmv_stream(const CsrMatrix &A, const T *x, T *y, const DistDev &d, cudaStream_t stream) {
  static int occs[16][2] = {};  // per device (function attributes and occupancy are per device), per smem size
  const bool hot = A.spmv_plan_hot_n > 0;
  const size_t smem = spmv_stream_smem_bytes<T>(hot);
  int &occ = occs[device_props().device & 15][hot ? 1 : 0];
  if (occ == 0) {
    QDSP_CUDA(cudaFuncSetAttribute(spmv_stream_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   static_cast<int>(spmv_stream_smem_bytes<T>(true))));
    int o = 0;
    QDSP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, spmv_stream_kernel<T>, SPMV_THREADS, smem));
    occ = std::max(1, o);
  }
  const int grid = std::max(1, std::min(A.spmv_plan_nblk, device_props().sm_count * occ));
  spmv_stream_kernel<T><<<grid, SPMV_THREADS, smem, stream>>>(A.indptr.get<int>(), A.indices.get<int>(), A.values.get<T>(), x,
                                                              y, spmv_plan_dev(A), d);
  QDSP_LAUNCH_CHECK();
}

// persistent grid = SM count x resident CTAs of this instantiation (queried once per instantiation)
template <typename T, int LPR>
int spmv_ctas_per_sm() {
  static int occs[16] = {};  // per device
  int &occ = occs[device_props().device & 15];
  if (occ == 0) {
    const size_t smem = spmv_smem_bytes<T>();
    QDSP_CUDA(cudaFuncSetAttribute(spmv_kernel<T, LPR, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    QDSP_CUDA(cudaFuncSetAttribute(spmv_kernel<T, LPR, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int o = 0;
    QDSP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, spmv_kernel<T, LPR, false>, SPMV_THREADS, smem));
    occ = std::max(1, o);
  }
  return occ;
}

template <typename T, int LPR>
void launch_spmv_t(const CsrMatrix &A, const T *x, T *y, const DistDev &d, cudaStream_t stream) {
  constexpr int R = SPMV_CONSUMERS / LPR;
  const int grid = std::max(1, std::min(ceil_div(A.rows, R), device_props().sm_count * spmv_ctas_per_sm<T, LPR>()));
  const size_t smem = spmv_smem_bytes<T>();
  const int *ip = A.indptr.get<int>(), *ix = A.indices.get<int>();
  unsigned *counters = A.spmv_sched_buf.get<unsigned>();
  if (counters) counters += 2 * (A.spmv_sched_next.fetch_add(1u, std::memory_order_relaxed) % 16);
  SpmvSched sched = spmv_sched_for(ceil_div(A.rows, R), grid, counters);
  if constexpr (LPR == 1) {
    spmv_attach_codes(A, sched);
    if (sched.codes != nullptr) {  // the coded instantiation (same shared-memory footprint, same grid)
      static bool attr[16] = {};
      if (!attr[device_props().device & 15]) {
        QDSP_CUDA(cudaFuncSetAttribute(spmv_kernel<T, 1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        QDSP_CUDA(cudaFuncSetAttribute(spmv_kernel<T, 1, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        attr[device_props().device & 15] = true;
      }
      if (A.row_major_order)
        spmv_kernel<T, 1, true, true><<<grid, SPMV_THREADS, smem, stream>>>(A.rows, ip, ix, A.values.get<T>(), x, y, d, sched);
      else
        spmv_kernel<T, 1, false, true><<<grid, SPMV_THREADS, smem, stream>>>(A.rows, ip, ix, A.values.get<T>(), x, y, d, sched);
      QDSP_LAUNCH_CHECK();
      return;
    }
  }
  if (A.row_major_order && LPR == 1)
    spmv_kernel<T, LPR, true><<<grid, SPMV_THREADS, smem, stream>>>(A.rows, ip, ix, A.values.get<T>(), x, y, d, sched);
  else
    spmv_kernel<T, LPR, false><<<grid, SPMV_THREADS, smem, stream>>>(A.rows, ip, ix, A.values.get<T>(), x, y, d, sched);
  QDSP_LAUNCH_CHECK();
}

template <typename T>
void launch_spmv_lpr(const CsrMatrix &A, const void *x, void *y, const DistDev &d, cudaStream_t stream) {
  const T *xv = static_cast<const T *>(x);
  T *yv = static_cast<T *>(y);
  switch (A.spmv_lpr) {
    case 1: launch_spmv_t<T, 1>(A, xv, yv, d, stream); break;
    case 2: launch_spmv_t<T, 2>(A, xv, yv, d, stream); break;
    case 4: launch_spmv_t<T, 4>(A, xv, yv, d, stream); break;
    case 8: launch_spmv_t<T, 8>(A, xv, yv, d, stream); break;
    case 16: launch_spmv_t<T, 16>(A, xv, yv, d, stream); break;
    default: launch_spmv_t<T, 32>(A, xv, yv, d, stream); break;
  }
}

}  // namespace

void launch_spmv(const CsrMatrix &A, const void *x, void *y, const DistDev &d, cudaStream_t stream) {
  if (A.rows == 0) return;
  spmv_prepare(A, stream);
  if (A.spmv_mode == 2) {
    if (A.dtype == QDSP_F32) launch_spmv_stream<float>(A, static_cast<const float *>(x), static_cast<float *>(y), d, stream);
    else launch_spmv_stream<double>(A, static_cast<const double *>(x), static_cast<double *>(y), d, stream);
    return;
  }
  if (A.dtype == QDSP_F32) launch_spmv_lpr<float>(A, x, y, d, stream);
  else launch_spmv_lpr<double>(A, x, y, d, stream);
}

void CsrMatrix::spmv(const void *x, void *y, cudaStream_t stream) const {
  QDSP_REQUIRE(x != nullptr && y != nullptr, QDSP_ERR_INVALID, "spmv: null vector");
  QDSP_REQUIRE(x != y, QDSP_ERR_INVALID, "spmv: x and y must not alias");
  launch_spmv(*this, x, y, DistDev{}, stream);
}

void CsrMatrix::to_host(std::vector<int32_t> &ip, std::vector<int32_t> &ix, std::vector<char> &vals) const {
  ip.resize(static_cast<size_t>(rows) + 1);
  ix.resize(static_cast<size_t>(nnz));
  vals.resize(static_cast<size_t>(nnz) * value_bytes());
  QDSP_CUDA(cudaMemcpy(ip.data(), indptr.get(), ip.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (nnz) {
    QDSP_CUDA(cudaMemcpy(ix.data(), indices.get(), ix.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    QDSP_CUDA(cudaMemcpy(vals.data(), values.get(), vals.size(), cudaMemcpyDeviceToHost));
  }
}

// CuSparseMatrix::get_element copies the whole CSR to the host on every call (sparse_matrix.cpp:688-712); here only
// the two indptr entries and the row's slice travel.
double CsrMatrix::get_element(int r, int c) const {
  QDSP_REQUIRE(r >= 0 && r < rows && c >= 0 && c < cols, QDSP_ERR_INVALID, "get_element: index out of range");
  int32_t be[2];
  QDSP_CUDA(cudaMemcpy(be, indptr.get<int32_t>() + r, sizeof(be), cudaMemcpyDeviceToHost));
  const int len = be[1] - be[0];
  if (len <= 0) return 0.0;
  std::vector<int32_t> ix(len);
  QDSP_CUDA(cudaMemcpy(ix.data(), indices.get<int32_t>() + be[0], sizeof(int32_t) * len, cudaMemcpyDeviceToHost));
  for (int k = 0; k < len; ++k) {
    if (ix[k] == c) {
      if (dtype == QDSP_F32) {
        float v;
        QDSP_CUDA(cudaMemcpy(&v, values.get<float>() + be[0] + k, sizeof(v), cudaMemcpyDeviceToHost));
        return v;
      }
      double v;
      QDSP_CUDA(cudaMemcpy(&v, values.get<double>() + be[0] + k, sizeof(v), cudaMemcpyDeviceToHost));
      return v;
    }
  }
  return 0.0;
}

// "%%MatrixMarket matrix coordinate real general", 1-based, row by row (sparse_matrix.cpp:748-780)
void CsrMatrix::mmwrite(const std::string &path) const {
  std::vector<int32_t> ip, ix;
  std::vector<char> vals;
  to_host(ip, ix, vals);
  std::ofstream f(path);
  QDSP_REQUIRE(f.good(), QDSP_ERR_INVALID, "mmwrite: cannot open " + path);
  f << "%%MatrixMarket matrix coordinate real general\n%" << std::endl;
  f << rows << " " << cols << " " << nnz << std::endl;
  for (int r = 0; r < rows; ++r)
    for (int k = ip[r]; k < ip[r + 1]; ++k) {
      f << r + 1 << " " << ix[k] + 1 << " ";
      if (dtype == QDSP_F32) f << reinterpret_cast<const float *>(vals.data())[k];
      else f << reinterpret_cast<const double *>(vals.data())[k];
      f << std::endl;
    }
}

}  // namespace qdsp
