// abi.cu -- the extern "C" surface declared in include/qdsp.h.  Thin: argument checks, handle casts, and the
// exception -> (status, thread-local message) translation.  No exception crosses this file.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "dist.hpp"
#include "host.hpp"
#include "solver.hpp"

using namespace qdsp;

// ---- scratch pool (common.cuh) -----------------------------------------------------------------------------------------------
namespace qdsp {
namespace {
struct PoolBlock { void *p; size_t bytes; int device; };
std::mutex g_pool_mutex;
std::vector<PoolBlock> g_pool;  // free blocks, oldest first
size_t g_pool_bytes = 0;
size_t pool_cap_bytes() {
  static const size_t cap = [] {
    const char *e = std::getenv("QDSP_SCRATCH_POOL_MB");
    const long long mb = e ? std::atoll(e) : 16384;
    return static_cast<size_t>(mb < 0 ? 0 : mb) << 20;
  }();
  return cap;
}
// What the pool may hold right now: the configured cap, but never more than a quarter of the device memory that would be
// free without it.  Other allocators on the device (torch's caching allocator in the host application) cannot reclaim
// blocks parked here, so the pool shrinks as the device fills up instead of sitting on gigabytes next to an OOM.
// (cudaMemGetInfo is a driver round trip: the answer is reused for 20 ms, a build releases half a dozen blocks in a row.)
size_t pool_cap_now(size_t pool_bytes, bool refresh = false) {
  size_t cap = pool_cap_bytes();
  if (cap == 0) return 0;
  static std::chrono::steady_clock::time_point stamp{};
  static size_t other_free = ~size_t{0};  // free + pooled at the last look (callers hold g_pool_mutex)
  const auto now = std::chrono::steady_clock::now();
  if (refresh || other_free == ~size_t{0} || now - stamp > std::chrono::milliseconds(20)) {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
      other_free = free_b + pool_bytes;
      stamp = now;
    } else {
      (void)cudaGetLastError();
    }
  }
  if (other_free != ~size_t{0}) cap = std::min(cap, other_free / 4);
  return cap;
}
constexpr size_t POOL_MIN_BYTES = size_t{1} << 20;  // smaller requests go straight to cudaMalloc
constexpr size_t POOL_GRANULE = size_t{2} << 20;
}  // namespace

void *scratch_pool_acquire(size_t bytes, size_t *got) {
  int dev = 0;
  QDSP_CUDA(cudaGetDevice(&dev));
  if (bytes >= POOL_MIN_BYTES && pool_cap_bytes() > 0) {
    std::lock_guard<std::mutex> lock(g_pool_mutex);
    int best = -1;
    for (int i = 0; i < static_cast<int>(g_pool.size()); ++i) {
      const PoolBlock &b = g_pool[i];
      // best fit, and never burn a block more than twice the request
      if (b.device == dev && b.bytes >= bytes && b.bytes <= 2 * bytes + POOL_GRANULE &&
          (best < 0 || b.bytes < g_pool[best].bytes))
        best = i;
    }
    if (best >= 0) {
      PoolBlock b = g_pool[best];
      g_pool.erase(g_pool.begin() + best);
      g_pool_bytes -= b.bytes;
      *got = b.bytes;
      return b.p;
    }
  }
  const size_t rounded = bytes >= POOL_MIN_BYTES ? (bytes + POOL_GRANULE - 1) / POOL_GRANULE * POOL_GRANULE : bytes;
  void *p = nullptr;
  cudaError_t e = cudaMalloc(&p, rounded);
  if (e == cudaErrorMemoryAllocation) {
    (void)cudaGetLastError();
    scratch_pool_trim();
    e = cudaMalloc(&p, rounded);
  }
  QDSP_CUDA(e);
  *got = rounded;
  return p;
}

void scratch_pool_release(void *p, size_t bytes) {
  if (!p) return;
  int dev = 0;
  if (bytes < POOL_MIN_BYTES || bytes > pool_cap_bytes() || cudaGetDevice(&dev) != cudaSuccess) {
    cudaFree(p);
    return;
  }
  std::vector<void *> drop;
  {
    std::lock_guard<std::mutex> lock(g_pool_mutex);
    g_pool.push_back({p, bytes, dev});
    g_pool_bytes += bytes;
    const size_t cap_now = pool_cap_now(g_pool_bytes);
    while (g_pool_bytes > cap_now && !g_pool.empty()) {  // evict the oldest
      drop.push_back(g_pool.front().p);
      g_pool_bytes -= g_pool.front().bytes;
      g_pool.erase(g_pool.begin());
    }
  }
  for (void *q : drop) cudaFree(q);
}

void scratch_pool_info(size_t *cached, size_t *cap_now) {
  std::lock_guard<std::mutex> lock(g_pool_mutex);
  if (cached) *cached = g_pool_bytes;
  if (cap_now) *cap_now = pool_cap_now(g_pool_bytes, true);
}

void scratch_pool_trim() {
  std::vector<PoolBlock> all;
  {
    std::lock_guard<std::mutex> lock(g_pool_mutex);
    all.swap(g_pool);
    g_pool_bytes = 0;
  }
  for (const PoolBlock &b : all) cudaFree(b.p);
}
}  // namespace qdsp

struct qdsp_builder_s { std::unique_ptr<Builder> impl; };
struct qdsp_csr_s { std::unique_ptr<CsrMatrix> impl; };
struct qdsp_cg_s { std::unique_ptr<CgSolver> impl; qdsp_csr_s *matrix; };
struct qdsp_dist_s { std::unique_ptr<DistContext> impl; };
struct qdsp_solver_s { std::unique_ptr<DirectSolver> impl; };

namespace {
thread_local std::string g_last_error;

template <typename F>
int guarded(F &&f) {
  try {
    f();
    return QDSP_OK;
  } catch (const Error &e) {
    g_last_error = e.what();
    return e.status;
  } catch (const std::bad_alloc &) {
    g_last_error = "out of host memory";
    return QDSP_ERR_INVALID;
  } catch (const std::exception &e) {
    g_last_error = e.what();
    return QDSP_ERR_INVALID;
  } catch (...) {
    g_last_error = "unknown error";
    return QDSP_ERR_INVALID;
  }
}
inline cudaStream_t S(void *s) { return static_cast<cudaStream_t>(s); }
#define NOT_NULL(p) QDSP_REQUIRE((p) != nullptr, QDSP_ERR_INVALID, std::string(__func__) + ": null " #p)
}  // namespace

extern "C" {

const char *qdsp_last_error(void) { return g_last_error.c_str(); }
int qdsp_abi_version(void) { return QDSP_ABI_VERSION; }
int64_t qdsp_launch_count(void) { return g_launch_count.load(); }
int qdsp_scratch_info(size_t *cached_bytes, size_t *cap_bytes) {
  return guarded([&] { scratch_pool_info(cached_bytes, cap_bytes); });
}
int qdsp_release_scratch(void) {
  return guarded([&] { scratch_pool_trim(); });
}

int qdsp_init(int device) {
  return guarded([&] {
    int n = 0;
    QDSP_CUDA(cudaGetDeviceCount(&n));
    QDSP_REQUIRE(n > 0, QDSP_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
    QDSP_REQUIRE(device >= 0 && device < n, QDSP_ERR_INVALID, "qdsp_init: no such device");
    QDSP_CUDA(cudaSetDevice(device));
    const DeviceProps &p = device_props();
    QDSP_REQUIRE(p.cc_major >= 10, QDSP_ERR_CUDA,
                 "qdsp is built for sm_100a only; device reports compute capability " + std::to_string(p.cc_major) +
                     "." + std::to_string(p.cc_minor));
  });
}

int qdsp_device_info(int *device, int *sm_count, int *cc_major, int *cc_minor, size_t *l2_bytes, size_t *hbm_bytes) {
  return guarded([&] {
    const DeviceProps &p = device_props();
    if (device) *device = p.device;
    if (sm_count) *sm_count = p.sm_count;
    if (cc_major) *cc_major = p.cc_major;
    if (cc_minor) *cc_minor = p.cc_minor;
    if (l2_bytes) *l2_bytes = p.l2_bytes;
    if (hbm_bytes) *hbm_bytes = p.hbm_bytes;
  });
}

// ---- builder ---------------------------------------------------------------------------------------------------------
int qdsp_builder_create(int rows, int cols, int64_t max_triplets, int dtype, const char *storage_format,
                        qdsp_builder_t *out) {
  return guarded([&] {
    NOT_NULL(out);
    auto *h = new qdsp_builder_s();
    try {
      h->impl = std::make_unique<Builder>(rows, cols, max_triplets, dtype, storage_format ? storage_format : "col_major");
    } catch (...) { delete h; throw; }
    *out = h;
  });
}
int qdsp_builder_buffer_ptr(qdsp_builder_t b, void **device_ptr) {
  return guarded([&] { NOT_NULL(b); NOT_NULL(device_ptr); *device_ptr = b->impl->buffer_ptr(); });
}
int qdsp_builder_count(qdsp_builder_t b, void *stream, int64_t *count) {
  return guarded([&] { NOT_NULL(b); NOT_NULL(count); *count = b->impl->count(S(stream)); });
}
int qdsp_builder_fill_from_host(qdsp_builder_t b, const void *host_triplets, int64_t count, void *stream) {
  return guarded([&] {
    NOT_NULL(b);
    QDSP_REQUIRE(count == 0 || host_triplets != nullptr, QDSP_ERR_INVALID, "fill_from_host: null triplets");
    b->impl->fill_from_host(host_triplets, count, S(stream));
  });
}
int qdsp_builder_clear(qdsp_builder_t b, void *stream) {
  return guarded([&] { NOT_NULL(b); b->impl->clear(S(stream)); });
}
int qdsp_builder_build(qdsp_builder_t b, void *stream, qdsp_csr_t *out) {
  return guarded([&] {
    NOT_NULL(b); NOT_NULL(out);
    auto m = b->impl->build(S(stream));
    auto *h = new qdsp_csr_s();
    h->impl = std::move(m);
    *out = h;
  });
}
int qdsp_builder_print_triplets(qdsp_builder_t b) {
  return guarded([&] { NOT_NULL(b); b->impl->print_triplets(); });
}
int qdsp_builder_destroy(qdsp_builder_t b) {
  return guarded([&] { delete b; });
}

// ---- matrix -----------------------------------------------------------------------------------------------------------
static qdsp_csr_t wrap(std::unique_ptr<CsrMatrix> m) {
  auto *h = new qdsp_csr_s();
  h->impl = std::move(m);
  return h;
}

int qdsp_csr_from_device_coo(int rows, int cols, int64_t nnz, const int32_t *row, const int32_t *col, const void *val,
                             int dtype, void *stream, qdsp_csr_t *out) {
  return guarded([&] {
    NOT_NULL(out);
    QDSP_REQUIRE(dtype == QDSP_F32 || dtype == QDSP_F64, QDSP_ERR_INVALID, "Unsupported sparse matrix data type!");
    QDSP_REQUIRE(nnz == 0 || (row && col && val), QDSP_ERR_INVALID, "build_csr_from_coo: null array");
    *out = wrap(CsrMatrix::from_device_coo(rows, cols, nnz, row, col, val, dtype, S(stream)));
  });
}
int qdsp_csr_from_device_csr(int rows, int cols, int64_t nnz, const int32_t *indptr, const int32_t *indices,
                             const void *values, int dtype, int order_row_major, void *stream, qdsp_csr_t *out) {
  return guarded([&] {
    NOT_NULL(out); NOT_NULL(indptr);
    QDSP_REQUIRE(rows >= 0 && cols >= 0 && nnz >= 0, QDSP_ERR_INVALID, "from_device_csr: bad shape");
    QDSP_REQUIRE(dtype == QDSP_F32 || dtype == QDSP_F64, QDSP_ERR_INVALID, "Unsupported sparse matrix data type!");
    QDSP_REQUIRE(nnz == 0 || (indices && values), QDSP_ERR_INVALID, "from_device_csr: null array");
    *out = wrap(CsrMatrix::from_device_csr(rows, cols, nnz, indptr, indices, values, dtype, order_row_major != 0, S(stream)));
  });
}
int qdsp_csr_info(qdsp_csr_t m, int *rows, int *cols, int64_t *nnz, int *dtype) {
  return guarded([&] {
    NOT_NULL(m);
    if (rows) *rows = m->impl->rows;
    if (cols) *cols = m->impl->cols;
    if (nnz) *nnz = m->impl->nnz;
    if (dtype) *dtype = m->impl->dtype;
  });
}
int qdsp_csr_export(qdsp_csr_t m, const int32_t **indptr, const int32_t **indices, const void **values) {
  return guarded([&] {
    NOT_NULL(m);
    if (indptr) *indptr = m->impl->indptr.get<int32_t>();
    if (indices) *indices = m->impl->indices.get<int32_t>();
    if (values) *values = m->impl->values.get();
  });
}
int qdsp_csr_spmv(qdsp_csr_t m, const void *x, void *y, void *stream) {
  return guarded([&] { NOT_NULL(m); m->impl->spmv(x, y, S(stream)); });
}
int qdsp_csr_spmv_host(qdsp_csr_t m, const void *x_host, void *y_host, void *stream) {
  return guarded([&] {
    NOT_NULL(m); NOT_NULL(x_host); NOT_NULL(y_host);
    const CsrMatrix &A = *m->impl;
    const size_t sv = A.value_bytes();
    DeviceBuffer x(sv * A.cols), y(sv * A.rows);
    QDSP_CUDA(cudaMemcpyAsync(x.get(), x_host, sv * A.cols, cudaMemcpyHostToDevice, S(stream)));
    A.spmv(x.get(), y.get(), S(stream));
    QDSP_CUDA(cudaMemcpyAsync(y_host, y.get(), sv * A.rows, cudaMemcpyDeviceToHost, S(stream)));
    QDSP_CUDA(cudaStreamSynchronize(S(stream)));
  });
}
int qdsp_csr_get_element(qdsp_csr_t m, int row, int col, double *out) {
  return guarded([&] { NOT_NULL(m); NOT_NULL(out); *out = m->impl->get_element(row, col); });
}
int qdsp_csr_add(qdsp_csr_t a, qdsp_csr_t b, double alpha, double beta, void *stream, qdsp_csr_t *out) {
  return guarded([&] { NOT_NULL(a); NOT_NULL(b); NOT_NULL(out); *out = wrap(a->impl->add(*b->impl, alpha, beta, S(stream))); });
}
int qdsp_csr_scale(qdsp_csr_t a, double alpha, void *stream, qdsp_csr_t *out) {
  return guarded([&] { NOT_NULL(a); NOT_NULL(out); *out = wrap(a->impl->scale(alpha, S(stream))); });
}
int qdsp_csr_transpose(qdsp_csr_t a, void *stream, qdsp_csr_t *out) {
  return guarded([&] { NOT_NULL(a); NOT_NULL(out); *out = wrap(a->impl->transpose(S(stream))); });
}
int qdsp_csr_matmul(qdsp_csr_t a, qdsp_csr_t b, void *stream, qdsp_csr_t *out) {
  return guarded([&] { NOT_NULL(a); NOT_NULL(b); NOT_NULL(out); *out = wrap(a->impl->matmul(*b->impl, S(stream))); });
}
int qdsp_csr_mmwrite(qdsp_csr_t m, const char *path) {
  return guarded([&] { NOT_NULL(m); NOT_NULL(path); m->impl->mmwrite(path); });
}
int qdsp_csr_destroy(qdsp_csr_t m) {
  return guarded([&] { delete m; });
}

// ---- CG -----------------------------------------------------------------------------------------------------------------
int qdsp_cg_create(qdsp_csr_t m, int max_iters, double tol, int mode, int verbose, qdsp_cg_t *out) {
  return guarded([&] {
    NOT_NULL(m); NOT_NULL(out);
    auto *h = new qdsp_cg_s();
    try {
      h->impl = std::make_unique<CgSolver>(*m->impl, max_iters, tol, mode, verbose != 0);
    } catch (...) { delete h; throw; }
    h->matrix = m;
    *out = h;
  });
}
int qdsp_cg_solve(qdsp_cg_t s, void *x, const void *b, void *stream, int *iterations, double *error, int *success) {
  return guarded([&] { NOT_NULL(s); s->impl->solve(x, b, S(stream), iterations, error, success); });
}
int qdsp_cg_solve_host(qdsp_cg_t s, void *x_host, const void *b_host, void *stream, int *iterations, double *error,
                       int *success) {
  return guarded([&] { NOT_NULL(s); s->impl->solve_host(x_host, b_host, S(stream), iterations, error, success); });
}
int qdsp_cg_run_fixed(qdsp_cg_t s, void *x, const void *b, int iters, void *stream) {
  return guarded([&] { NOT_NULL(s); s->impl->run_fixed(x, b, iters, S(stream)); });
}
int qdsp_cg_destroy(qdsp_cg_t s) {
  return guarded([&] { delete s; });
}

int qdsp_csr_spmv_config(qdsp_csr_t m, int compress) {
  return guarded([&] {
    NOT_NULL(m);
    spmv_set_compress(*m->impl, compress, nullptr);
  });
}
int qdsp_csr_spmv_info(qdsp_csr_t m, void *stream, int *index_bytes_per_nnz, int *path, int *hot_columns) {
  return guarded([&] {
    NOT_NULL(m);
    spmv_prepare(*m->impl, S(stream));
    if (index_bytes_per_nnz) *index_bytes_per_nnz = m->impl->spmv_ntab > 0 ? 1 : 4;
    if (path) *path = m->impl->spmv_mode;
    if (hot_columns) *hot_columns = m->impl->spmv_mode == 2 ? m->impl->spmv_plan_hot_n : 0;
  });
}

// ---- SpMV plan (pure host) ------------------------------------------------------------------------------------------------
int qdsp_spmv_plan_host(const int32_t *indptr, int rows, int32_t *blocks_out, int64_t block_capacity, int64_t *n_blocks,
                        int *n_long_rows, int *n_segments) {
  return guarded([&] {
    NOT_NULL(indptr); NOT_NULL(n_blocks);
    QDSP_REQUIRE(rows >= 0, QDSP_ERR_INVALID, "spmv plan: negative row count");
    std::vector<int4> blocks;
    std::vector<int2> long_rows;
    int nseg = 0;
    spmv_plan_host(indptr, rows, blocks, long_rows, nseg);
    *n_blocks = static_cast<int64_t>(blocks.size());
    if (n_long_rows) *n_long_rows = static_cast<int>(long_rows.size()) - 1;
    if (n_segments) *n_segments = nseg;
    if (blocks_out) {
      QDSP_REQUIRE(block_capacity >= *n_blocks, QDSP_ERR_INVALID, "spmv plan: blocks_out too small");
      std::memcpy(blocks_out, blocks.data(), sizeof(int4) * blocks.size());
    }
  });
}

// ---- multi-GPU ------------------------------------------------------------------------------------------------------------
int qdsp_dist_plan_host(int rank, int world, const int64_t *row_offsets, int32_t *indices_inout, int64_t nnz,
                        int ghost_base, int32_t *ghost_ids_out, int64_t ghost_capacity, int64_t *n_ghost,
                        int32_t *owner_offsets_out) {
  return guarded([&] {
    NOT_NULL(n_ghost); NOT_NULL(owner_offsets_out);
    const DistPlan p = dist_plan_host(rank, world, row_offsets, indices_inout, nnz, ghost_base);
    *n_ghost = static_cast<int64_t>(p.ghost_ids.size());
    for (int q = 0; q <= world; ++q) owner_offsets_out[q] = p.owner_offsets[q];
    if (ghost_ids_out) {
      QDSP_REQUIRE(ghost_capacity >= *n_ghost, QDSP_ERR_INVALID, "dist plan: ghost_ids_out too small");
      std::memcpy(ghost_ids_out, p.ghost_ids.data(), sizeof(int32_t) * p.ghost_ids.size());
    }
  });
}
int qdsp_dist_create(int rank, int world, qdsp_dist_t *out) {
  return guarded([&] {
    NOT_NULL(out);
    auto *h = new qdsp_dist_s();
    try { h->impl = std::make_unique<DistContext>(rank, world); } catch (...) { delete h; throw; }
    *out = h;
  });
}
int qdsp_dist_destroy(qdsp_dist_t d) {
  return guarded([&] { delete d; });
}
int qdsp_dist_localize(qdsp_dist_t d, qdsp_csr_t m, const int64_t *row_offsets, void *stream, int *n_local,
                       int *ghost_base, int *n_ghost, int32_t *owner_offsets_out) {
  return guarded([&] {
    NOT_NULL(d); NOT_NULL(m);
    d->impl->localize(*m->impl, row_offsets, S(stream));
    if (n_local) *n_local = d->impl->n_local();
    if (ghost_base) *ghost_base = d->impl->ghost_base();
    if (n_ghost) *n_ghost = d->impl->n_ghost();
    if (owner_offsets_out)
      for (int q = 0; q <= d->impl->world(); ++q) owner_offsets_out[q] = d->impl->plan().owner_offsets[q];
  });
}
int qdsp_dist_ghost_ids(qdsp_dist_t d, const int32_t **host_ids) {
  return guarded([&] { NOT_NULL(d); NOT_NULL(host_ids); *host_ids = d->impl->plan().ghost_ids.data(); });
}
int qdsp_dist_set_send_list(qdsp_dist_t d, int peer, const int32_t *global_ids, int count, int dst_offset,
                            int dst_offset_z) {
  return guarded([&] { NOT_NULL(d); d->impl->set_send_list(peer, global_ids, count, dst_offset, dst_offset_z); });
}
int qdsp_dist_comm_handle(qdsp_dist_t d, void *out64) {
  return guarded([&] { NOT_NULL(d); NOT_NULL(out64); d->impl->comm_handle(out64); });
}
int qdsp_dist_attach(qdsp_dist_t d, const void *handles) {
  return guarded([&] { NOT_NULL(d); NOT_NULL(handles); d->impl->attach(handles); });
}
int qdsp_dist_nccl_unique_id(void *out128) {
  return guarded([&] { NOT_NULL(out128); DistContext::nccl_unique_id(out128); });
}
int qdsp_dist_nccl_init(qdsp_dist_t d, const void *id128) {
  return guarded([&] { NOT_NULL(d); NOT_NULL(id128); d->impl->nccl_init(id128); });
}
int qdsp_dist_spmv(qdsp_dist_t d, qdsp_csr_t m, const void *x_local, void *y_local, void *stream) {
  return guarded([&] { NOT_NULL(d); NOT_NULL(m); NOT_NULL(x_local); NOT_NULL(y_local); d->impl->spmv(*m->impl, x_local, y_local, S(stream)); });
}
int qdsp_cg_create_dist(qdsp_csr_t m, qdsp_dist_t d, int max_iters, double tol, int mode, int verbose, qdsp_cg_t *out) {
  return guarded([&] {
    NOT_NULL(m); NOT_NULL(d); NOT_NULL(out);
    auto *h = new qdsp_cg_s();
    try {
      h->impl = std::make_unique<CgSolver>(*m->impl, max_iters, tol, mode, verbose != 0, d->impl.get());
    } catch (...) { delete h; throw; }
    h->matrix = m;
    *out = h;
  });
}

// ---- algorithms -----------------------------------------------------------------------------------------------------------
size_t qdsp_sort_tmp_bytes(int64_t n, int key_kind, int value_bytes) {
  if (n < 0) return 0;
  return sort_tmp_bytes(n, key_kind >= QDSP_KEY_U64 ? 8 : 4, value_bytes);
}
int qdsp_sort_pairs(void *keys, void *values, int64_t n, int key_kind, int value_bytes, void *tmp, size_t tmp_bytes,
                    void *stream) {
  return guarded([&] { sort_pairs(keys, values, n, key_kind, value_bytes, tmp, tmp_bytes, S(stream)); });
}
size_t qdsp_scan_tmp_bytes(int64_t n) { return n < 0 ? 0 : scan_tmp_bytes(n); }
int qdsp_inclusive_scan_i32(int32_t *inout, int64_t n, void *tmp, size_t tmp_bytes, void *stream) {
  return guarded([&] { inclusive_scan_i32(inout, n, tmp, tmp_bytes, S(stream)); });
}
int64_t qdsp_scan_static_max_n(void) { return scan_static_max_n(); }
int qdsp_inclusive_scan_i32_static(int32_t *inout, int64_t n, void *stream) {
  return guarded([&] { inclusive_scan_i32_static(inout, n, S(stream)); });
}

// ---- device-wide algorithms ---------------------------------------------------------------------------------
size_t qdsp_algo_scratch_bytes(int algo, int64_t capacity, int elem_bytes) {
  return algo_scratch_bytes(algo, capacity, elem_bytes);
}
int qdsp_reduce(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op, void *scratch,
                size_t scratch_bytes, void *stream) {
  return guarded([&] { device_reduce(in, out, capacity, n_dev, elem, op, scratch, scratch_bytes, S(stream)); });
}
int qdsp_exclusive_scan(const void *in, void *out, int64_t capacity, const int32_t *n_dev, int elem, int op,
                        void *scratch, size_t scratch_bytes, void *stream) {
  return guarded([&] { device_exclusive_scan(in, out, capacity, n_dev, elem, op, scratch, scratch_bytes, S(stream)); });
}
int qdsp_select(const void *in, const int32_t *flags, void *out, int32_t *num_out, int64_t capacity,
                const int32_t *n_dev, int item_bytes, void *scratch, size_t scratch_bytes, void *stream) {
  return guarded([&] {
    device_select(in, flags, out, num_out, capacity, n_dev, item_bytes, scratch, scratch_bytes, S(stream));
  });
}
int qdsp_reduce_by_key_add(const void *keys_in, const void *values_in, void *keys_out, void *values_out,
                           int32_t *num_runs, int64_t capacity, const int32_t *n_dev, int key_elem, int value_elem,
                           void *scratch, size_t scratch_bytes, void *stream) {
  return guarded([&] {
    device_reduce_by_key_add(keys_in, values_in, keys_out, values_out, num_runs, capacity, n_dev, key_elem, value_elem,
                             scratch, scratch_bytes, S(stream));
  });
}
int qdsp_sort(void *keys, void *tmp_keys, void *values, void *tmp_values, int64_t capacity, const int32_t *n_dev,
              int key_elem, int value_bytes, int end_bit, void *scratch, size_t scratch_bytes, void *stream) {
  return guarded([&] {
    sort_modern(keys, tmp_keys, values, tmp_values, capacity, n_dev, key_elem, value_bytes, end_bit, scratch,
                scratch_bytes, S(stream));
  });
}

// ---- triplet emission ---------------------------------------------------------------------------------------
int qdsp_builder_append(qdsp_builder_t b, const int32_t *row, const int32_t *col, const void *val, int64_t count,
                        int negate, void *stream) {
  return guarded([&] { NOT_NULL(b); b->impl->append(row, col, val, count, negate != 0, S(stream)); });
}
int qdsp_builder_append_elements(qdsp_builder_t b, const int32_t *nodes, const void *mats, int64_t n_elems, int k,
                                 void *stream) {
  return guarded([&] { NOT_NULL(b); b->impl->append_elements(nodes, mats, n_elems, k, S(stream)); });
}

// ---- SparseSolver (solver.cu) ---------------------------------------------------------------------------------------------------
int qdsp_solver_create(int dtype, int solver_type, int ordering, qdsp_solver_t *out) {
  return guarded([&] {
    NOT_NULL(out);
    auto *h = new qdsp_solver_s();
    try {
      h->impl = std::make_unique<DirectSolver>(dtype, solver_type, ordering);
    } catch (...) {
      delete h;
      throw;
    }
    *out = h;
  });
}
int qdsp_solver_analyze_pattern(qdsp_solver_t s, qdsp_csr_t m, void *stream) {
  return guarded([&] { NOT_NULL(s); NOT_NULL(m); s->impl->analyze_pattern(*m->impl, S(stream)); });
}
int qdsp_solver_factorize(qdsp_solver_t s, qdsp_csr_t m, void *stream, int *success, int *zero_pivot) {
  return guarded([&] {
    NOT_NULL(s); NOT_NULL(m);
    s->impl->factorize(*m->impl, S(stream));
    if (success) *success = s->impl->info() ? 1 : 0;
    if (zero_pivot) *zero_pivot = s->impl->zero_pivot();
  });
}
int qdsp_solver_solve(qdsp_solver_t s, const void *b, void *x, void *stream) {
  return guarded([&] { NOT_NULL(s); s->impl->solve(b, x, S(stream)); });
}
int qdsp_solver_solve_host(qdsp_solver_t s, const void *b_host, void *x_host, void *stream) {
  return guarded([&] { NOT_NULL(s); s->impl->solve_host(b_host, x_host, S(stream)); });
}
int qdsp_solver_info(qdsp_solver_t s, int *success) {
  return guarded([&] { NOT_NULL(s); NOT_NULL(success); *success = s->impl->info() ? 1 : 0; });
}
int qdsp_solver_stats(qdsp_solver_t s, int64_t *factor_nnz, int *etree_height, double *flops) {
  return guarded([&] {
    NOT_NULL(s);
    const SymbolicFactor &f = s->impl->symbolic();
    if (factor_nnz) *factor_nnz = f.factor_nnz;
    if (etree_height) *etree_height = f.height;
    if (flops) *flops = f.flops;
  });
}
int qdsp_solver_destroy(qdsp_solver_t s) {
  return guarded([&] { delete s; });
}
int qdsp_solver_lu_row_matching_host(int n, const int32_t *indptr, const int32_t *indices, const double *abs_values,
                                     int32_t *rowof_out, int *zero_free_diagonal) {
  return guarded([&] {
    NOT_NULL(indptr);
    NOT_NULL(rowof_out);
    QDSP_REQUIRE(n >= 0, QDSP_ERR_INVALID, "solver row matching: negative size");
    const int64_t nnz = indptr[n];
    QDSP_REQUIRE(nnz == 0 || (indices != nullptr && abs_values != nullptr), QDSP_ERR_INVALID, "solver row matching: null array");
    for (int64_t q = 0; q < nnz; ++q)
      QDSP_REQUIRE(indices[q] >= 0 && indices[q] < n, QDSP_ERR_INVALID, "solver row matching: column index out of range");
    std::vector<double> w(abs_values, abs_values + nnz);
    std::vector<int32_t> rowof;
    const bool ok = solver_lu_row_matching_host(n, indptr, indices, w, rowof);
    std::copy(rowof.begin(), rowof.end(), rowof_out);
    if (zero_free_diagonal) *zero_free_diagonal = ok ? 1 : 0;
  });
}
int qdsp_solver_analyze_host(int n, const int32_t *indptr, const int32_t *indices, int natural_order, int32_t *perm_out,
                             int32_t *parent_out, int32_t *colptr_out, int32_t *rowidx_out, int64_t rowidx_capacity,
                             int64_t *factor_nnz, int *etree_height) {
  return guarded([&] {
    NOT_NULL(indptr);
    QDSP_REQUIRE(n >= 0, QDSP_ERR_INVALID, "solver analyse: negative size");
    QDSP_REQUIRE(indices != nullptr || indptr[n] == 0, QDSP_ERR_INVALID, "solver analyse: null indices");
    std::vector<int32_t> perm(n);
    const auto t0 = std::chrono::steady_clock::now();
    if (natural_order) for (int i = 0; i < n; ++i) perm[i] = i;
    else solver_order_host(n, indptr, indices, perm);
    const auto t1 = std::chrono::steady_clock::now();
    SymbolicFactor f;
    solver_symbolic_host(n, indptr, indices, perm, f);
    if (std::getenv("QDSP_SOLVER_TIMING"))
      std::fprintf(stderr, "[qdsp solver] ordering %.1f ms, symbolic %.1f ms\n",
                   std::chrono::duration<double, std::milli>(t1 - t0).count(),
                   std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count());
    if (perm_out) std::memcpy(perm_out, f.perm.data(), sizeof(int32_t) * n);
    if (parent_out) std::memcpy(parent_out, f.parent.data(), sizeof(int32_t) * n);
    if (colptr_out) std::memcpy(colptr_out, f.colptr.data(), sizeof(int32_t) * (n + 1));
    if (rowidx_out) {
      QDSP_REQUIRE(rowidx_capacity >= f.factor_nnz, QDSP_ERR_INVALID, "solver analyse: rowidx_capacity too small");
      std::memcpy(rowidx_out, f.rowidx.data(), sizeof(int32_t) * f.factor_nnz);
    }
    if (factor_nnz) *factor_nnz = f.factor_nnz;
    if (etree_height) *etree_height = f.height;
  });
}

// ---- dense vector kernels (krylov.cu) --------------------------------------------------------------------------------------------
int qdsp_vec_dot2(const void *a, const void *b, const void *c, const void *d, int64_t n, int dtype, void *stream,
                  double *out) {
  return guarded([&] { vec_dot2(a, b, c, d, n, dtype, S(stream), out); });
}
int qdsp_vec_lincomb(void *out, int64_t n, int dtype, int k, const void *const *xs, const double *coefs, void *stream) {
  return guarded([&] { vec_lincomb(out, n, dtype, k, xs, coefs, S(stream)); });
}
int qdsp_vec_cg_step(void *x, void *r, const void *p, const void *ap, double alpha, int64_t n, int dtype, void *stream,
                     double *rr) {
  return guarded([&] { vec_cg_step(x, r, p, ap, alpha, n, dtype, S(stream), rr); });
}

}  // extern "C"
