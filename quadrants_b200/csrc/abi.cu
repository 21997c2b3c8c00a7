// abi.cu -- the extern "C" surface declared in include/qdsp.h.  Thin: argument checks, handle casts, and the
// exception -> (status, thread-local message) translation.  No exception crosses this file.
#include <cstring>
#include <new>

#include "host.hpp"

using namespace qdsp;

struct qdsp_builder_s { std::unique_ptr<Builder> impl; };
struct qdsp_csr_s { std::unique_ptr<CsrMatrix> impl; };
struct qdsp_cg_s { std::unique_ptr<CgSolver> impl; qdsp_csr_s *matrix; };

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
  return guarded([&] {
    NOT_NULL(s); NOT_NULL(x_host); NOT_NULL(b_host);
    const CsrMatrix &A = *s->matrix->impl;
    const size_t bytes = A.value_bytes() * static_cast<size_t>(A.rows);
    DeviceBuffer x(bytes), b(bytes);
    QDSP_CUDA(cudaMemcpyAsync(b.get(), b_host, bytes, cudaMemcpyHostToDevice, S(stream)));
    QDSP_CUDA(cudaMemcpyAsync(x.get(), x_host, bytes, cudaMemcpyHostToDevice, S(stream)));  // CUCG starts from x
    s->impl->solve(x.get(), b.get(), S(stream), iterations, error, success);
    QDSP_CUDA(cudaMemcpyAsync(x_host, x.get(), bytes, cudaMemcpyDeviceToHost, S(stream)));
    QDSP_CUDA(cudaStreamSynchronize(S(stream)));
  });
}
int qdsp_cg_run_fixed(qdsp_cg_t s, void *x, const void *b, int iters, void *stream) {
  return guarded([&] { NOT_NULL(s); s->impl->run_fixed(x, b, iters, S(stream)); });
}
int qdsp_cg_destroy(qdsp_cg_t s) {
  return guarded([&] { delete s; });
}

// ---- algorithms -----------------------------------------------------------------------------------------------------------
size_t qdsp_sort_tmp_bytes(int64_t n, int key_kind, int value_bytes) {
  if (n < 0) return 0;
  return sort_tmp_bytes(n, key_kind == QDSP_KEY_U64 ? 8 : 4, value_bytes);
}
int qdsp_sort_pairs(void *keys, void *values, int64_t n, int key_kind, int value_bytes, void *tmp, size_t tmp_bytes,
                    void *stream) {
  return guarded([&] { sort_pairs(keys, values, n, key_kind, value_bytes, tmp, tmp_bytes, S(stream)); });
}
size_t qdsp_scan_tmp_bytes(int64_t n) { return n < 0 ? 0 : scan_tmp_bytes(n); }
int qdsp_inclusive_scan_i32(int32_t *inout, int64_t n, void *tmp, size_t tmp_bytes, void *stream) {
  return guarded([&] { inclusive_scan_i32(inout, n, tmp, tmp_bytes, S(stream)); });
}

}  // extern "C"
