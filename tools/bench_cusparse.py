#!/usr/bin/env python
"""BENCH-ONLY harness: the reference's own CUDA path for SpMV and CG, called exactly as the reference calls it, on the same
B200 and the same matrix, so that "the kernel to beat"  is a number.

  * CuSparseMatrix::spmv (quadrants/program/sparse_matrix.cpp:654-686): per call cusparseCreate + cusparseCreateDnVec x 2 +
    cusparseSpMV_bufferSize + cuMemAlloc + cusparseSpMV(NON_TRANSPOSE, alpha = 1, beta = 0, CUDA_R_32F, CSR_ALG1) +
    destroy + cuMemFree.  Timed twice: AS SHIPPED (everything per call) and HOISTED (handle, descriptors and buffer made
    once; only cusparseSpMV in the loop) -- the second is cuSPARSE's kernel, the first is what a Quadrants user gets.
  * CUCG::solve (quadrants/program/conjugate_gradient.cpp:19-83): r = b - A x, p = r, then per iteration SpMV, Sdot (host
    result), Saxpy x, Saxpy r, Sdot (host result), Sscal + Saxpy for p.  Timed with the hoisted SpMV (best case for the
    reference) and with the as-shipped SpMV.

The libraries are the pip packages already in the image (nvidia-cusparse-cu12, nvidia-cublas-cu12), loaded with ctypes.
NOTHING under quadrants_b200/ imports this file or these libraries (tests/test_abi.py checks); bench.py calls it only to
print roofline.kernel_to_beat_ms beside its own numbers.

    python tools/bench_cusparse.py            # C2 matrix (256^3), one JSON line
"""
import ctypes as C
import glob
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CUSPARSE_INDEX_32I = 2
CUSPARSE_INDEX_BASE_ZERO = 0
CUDA_R_32F = 0
CUSPARSE_OPERATION_NON_TRANSPOSE = 0
CUSPARSE_SPMV_CSR_ALG1 = 2


def _find(name):
    import site

    roots = list(site.getsitepackages()) + [os.path.dirname(os.path.dirname(os.__file__))]
    for r in roots:
        hits = sorted(glob.glob(os.path.join(r, "nvidia", "*", "lib", name + ".so*")))
        if hits:
            return hits[-1]
    for cand in (f"/usr/local/cuda/lib64/{name}.so", f"{name}.so.12", f"{name}.so"):
        try:
            C.CDLL(cand)
            return cand
        except OSError:
            continue
    raise OSError(f"{name} not found")


class Libs:
    def __init__(self):
        self.sp = C.CDLL(_find("libcusparse"))
        self.bl = C.CDLL(_find("libcublas"))
        self.rt = C.CDLL(_find("libcudart")) if False else None


def ck(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} -> status {rc}")


class CuSparseSpmv:
    """y = A x with cusparseSpMV(CSR_ALG1, f32); `shipped` re-does everything per call like sparse_matrix.cpp:654-686."""

    def __init__(self, libs, rows, cols, nnz, indptr, indices, values):
        import torch

        self.L, self.torch = libs, torch
        self.rows, self.cols, self.nnz = rows, cols, nnz
        self.indptr, self.indices, self.values = indptr, indices, values
        sp = libs.sp
        self.mat = C.c_void_p()
        # the reference creates the CSR descriptor once at build time (sparse_matrix.cpp:375-377)
        ck(sp.cusparseCreateCsr(C.byref(self.mat), C.c_int64(rows), C.c_int64(cols), C.c_int64(nnz),
                                C.c_void_p(indptr.data_ptr()), C.c_void_p(indices.data_ptr()),
                                C.c_void_p(values.data_ptr()), CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I,
                                CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F), "cusparseCreateCsr")
        self.one, self.zero = C.c_float(1.0), C.c_float(0.0)
        self.h = None

    def _descr(self, x, y):
        sp = self.L.sp
        vx, vy = C.c_void_p(), C.c_void_p()
        ck(sp.cusparseCreateDnVec(C.byref(vx), C.c_int64(self.cols), C.c_void_p(x.data_ptr()), CUDA_R_32F), "CreateDnVec")
        ck(sp.cusparseCreateDnVec(C.byref(vy), C.c_int64(self.rows), C.c_void_p(y.data_ptr()), CUDA_R_32F), "CreateDnVec")
        return vx, vy

    def shipped(self, x, y):
        sp, torch = self.L.sp, self.torch
        h = C.c_void_p()
        ck(sp.cusparseCreate(C.byref(h)), "cusparseCreate")
        vx, vy = self._descr(x, y)
        size = C.c_size_t()
        ck(sp.cusparseSpMV_bufferSize(h, CUSPARSE_OPERATION_NON_TRANSPOSE, C.byref(self.one), self.mat, vx,
                                      C.byref(self.zero), vy, CUDA_R_32F, CUSPARSE_SPMV_CSR_ALG1, C.byref(size)), "bufferSize")
        buf = torch.empty(max(size.value, 4), dtype=torch.uint8, device="cuda")  # cuMemAlloc in the reference
        ck(sp.cusparseSpMV(h, CUSPARSE_OPERATION_NON_TRANSPOSE, C.byref(self.one), self.mat, vx, C.byref(self.zero), vy,
                           CUDA_R_32F, CUSPARSE_SPMV_CSR_ALG1, C.c_void_p(buf.data_ptr())), "cusparseSpMV")
        sp.cusparseDestroyDnVec(vx)
        sp.cusparseDestroyDnVec(vy)
        sp.cusparseDestroy(h)
        del buf

    def hoist(self, x, y):
        sp, torch = self.L.sp, self.torch
        self.h = C.c_void_p()
        ck(sp.cusparseCreate(C.byref(self.h)), "cusparseCreate")
        ck(sp.cusparseSetStream(self.h, C.c_void_p(torch.cuda.current_stream().cuda_stream)), "SetStream")
        self.vx, self.vy = self._descr(x, y)
        size = C.c_size_t()
        ck(sp.cusparseSpMV_bufferSize(self.h, CUSPARSE_OPERATION_NON_TRANSPOSE, C.byref(self.one), self.mat, self.vx,
                                      C.byref(self.zero), self.vy, CUDA_R_32F, CUSPARSE_SPMV_CSR_ALG1, C.byref(size)), "bufferSize")
        self.buf = torch.empty(max(size.value, 4), dtype=torch.uint8, device="cuda")

    def hoisted(self):
        ck(self.L.sp.cusparseSpMV(self.h, CUSPARSE_OPERATION_NON_TRANSPOSE, C.byref(self.one), self.mat, self.vx,
                                  C.byref(self.zero), self.vy, CUDA_R_32F, CUSPARSE_SPMV_CSR_ALG1, C.c_void_p(self.buf.data_ptr())),
           "cusparseSpMV")


def time_ms(fn, reps, torch):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def cucg_iterations(libs, spmv_fn_factory, n, b, iters, torch):
    """CUCG::solve's loop (conjugate_gradient.cpp:43-78) for a fixed number of iterations; returns seconds."""
    bl = libs.bl
    h = C.c_void_p()
    ck(bl.cublasCreate_v2(C.byref(h)), "cublasCreate")
    ck(bl.cublasSetStream_v2(h, C.c_void_p(torch.cuda.current_stream().cuda_stream)), "cublasSetStream")
    x = torch.zeros(n, dtype=torch.float32, device="cuda")
    r = b.clone()
    p = r.clone()
    ax = torch.zeros(n, dtype=torch.float32, device="cuda")
    spmv = spmv_fn_factory(p, ax)
    r1 = C.c_float(0.0)
    dot = C.c_float(0.0)
    ck(bl.cublasSdot_v2(h, n, C.c_void_p(r.data_ptr()), 1, C.c_void_p(r.data_ptr()), 1, C.byref(r1)), "Sdot")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        spmv()                                                                                 # Ax = A p
        ck(bl.cublasSdot_v2(h, n, C.c_void_p(p.data_ptr()), 1, C.c_void_p(ax.data_ptr()), 1, C.byref(dot)), "Sdot")
        a = C.c_float(r1.value / dot.value if dot.value != 0 else 0.0)
        ck(bl.cublasSaxpy_v2(h, n, C.byref(a), C.c_void_p(p.data_ptr()), 1, C.c_void_p(x.data_ptr()), 1), "Saxpy")
        na = C.c_float(-a.value)
        ck(bl.cublasSaxpy_v2(h, n, C.byref(na), C.c_void_p(ax.data_ptr()), 1, C.c_void_p(r.data_ptr()), 1), "Saxpy")
        r0 = r1.value
        ck(bl.cublasSdot_v2(h, n, C.c_void_p(r.data_ptr()), 1, C.c_void_p(r.data_ptr()), 1, C.byref(r1)), "Sdot")
        beta = C.c_float(r1.value / r0 if r0 != 0 else 0.0)
        one = C.c_float(1.0)
        ck(bl.cublasSscal_v2(h, n, C.byref(beta), C.c_void_p(p.data_ptr()), 1), "Sscal")                # p = r + beta p
        ck(bl.cublasSaxpy_v2(h, n, C.byref(one), C.c_void_p(r.data_ptr()), 1, C.c_void_p(p.data_ptr()), 1), "Saxpy")
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    bl.cublasDestroy_v2(h)
    return dt


def measure(indptr, indices, values, rows, cols, spmv_reps=50, cg_iters=100):
    """indptr/indices/values: CUDA torch tensors (i32, i32, f32).  -> dict of the reference-path numbers."""
    import torch

    libs = Libs()
    nnz = int(indices.numel())
    sp = CuSparseSpmv(libs, rows, cols, nnz, indptr, indices, values)
    x = torch.empty(cols, dtype=torch.float32, device="cuda").uniform_(-1, 1)
    y = torch.zeros(rows, dtype=torch.float32, device="cuda")
    sp.hoist(x, y)
    for _ in range(5):
        sp.hoisted()
    hoisted_ms = time_ms(sp.hoisted, spmv_reps, torch)
    for _ in range(3):
        sp.shipped(x, y)
    shipped_ms = time_ms(lambda: sp.shipped(x, y), max(5, spmv_reps // 5), torch)
    alg = nnz * 8 + (rows + 1) * 4 + cols * 4 + rows * 4
    b = torch.ones(rows, dtype=torch.float32, device="cuda")

    def hoisted_factory(p, ax):
        s2 = CuSparseSpmv(libs, rows, cols, nnz, indptr, indices, values)
        s2.hoist(p, ax)
        return s2.hoisted

    def shipped_factory(p, ax):
        return lambda: sp.shipped(p, ax)

    cucg_iterations(libs, hoisted_factory, rows, b, 5, torch)
    t_h = cucg_iterations(libs, hoisted_factory, rows, b, cg_iters, torch)
    t_s = cucg_iterations(libs, shipped_factory, rows, b, max(10, cg_iters // 4), torch)
    return {"library": "cuSPARSE / cuBLAS from the image's pip packages (nvidia-cusparse-cu12, nvidia-cublas-cu12)",
            "spmv_csr_alg1_hoisted_ms": hoisted_ms, "spmv_csr_alg1_hoisted_gbs": alg / (hoisted_ms * 1e-3) / 1e9,
            "spmv_as_shipped_ms": shipped_ms, "spmv_as_shipped_gbs": alg / (shipped_ms * 1e-3) / 1e9,
            "cucg_hoisted_iters_per_s": cg_iters / t_h, "cucg_as_shipped_iters_per_s": max(10, cg_iters // 4) / t_s,
            "reference": "sparse_matrix.cpp:654-686 (SpMV), conjugate_gradient.cpp:43-78 (CUCG loop)"}


def main():
    import torch

    import quadrants_b200 as qd
    from bench import poisson3d_triplets_device

    qd.init(arch=qd.cuda, device=0)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    dev = torch.device("cuda", 0)
    r, c, v = poisson3d_triplets_device(n, 0, n, dev)
    K = qd.linalg.SparseMatrixBuilder(n ** 3, n ** 3, max_num_triplets=int(r.shape[0]), dtype=qd.f32)
    K.ptr.insert_triplets(r, c, v)
    del r, c, v
    A = K.build()
    del K
    ip, ix, vals = A.matrix.csr_tensors()
    out = measure(ip, ix, vals, n ** 3, n ** 3)
    # ours, same run
    x = qd.ndarray(qd.f32, n ** 3)
    x.tensor.uniform_(-1, 1)
    y = qd.ndarray(qd.f32, n ** 3)
    for _ in range(5):
        A.matrix.spmv(None, x, y)
    out["qdsp_spmv_ms"] = time_ms(lambda: A.matrix.spmv(None, x, y), 50, torch)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
