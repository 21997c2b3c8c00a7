"""ctypes front-end of the CPU checker (oracle/qd_oracle.c and, when built, oracle/_ref).

TEST INFRASTRUCTURE ONLY.  May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  The product package (quadrants_b200) never imports it.

Numpy arrays in, numpy arrays out.  ``Oracle`` = the plain-C restatement (always available after
``make -C oracle``); ``EigenRef`` = the reference's vendored Eigen compiled with the reference's flags (available when
oracle/_ref/libeigen_ref.so was built, i.e. in a container that has /root/reference, or prebuilt and shipped).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_I32P = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_VOIDP = C.c_void_p


def build(force: bool = False) -> None:
    """Compile the checker (gcc; plus the Eigen _ref when the reference tree is present)."""
    if force:
        subprocess.run(["make", "-C", _HERE, "clean"], check=False, capture_output=True)
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if r.returncode != 0 or not os.path.exists(os.path.join(_HERE, "libqd_oracle.so")):
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)


def _load(path):
    return C.CDLL(path) if os.path.exists(path) else None


def make_triplet_buffer(rows, cols, vals, dtype, capacity=None):
    """The builder buffer of SURVEY 8(a) A1: [count | (row, col, value-bits)...] with 4-byte (f32) or 8-byte (f64) slots
    (quadrants/runtime/llvm/runtime_module/internal_functions.h:12-21)."""
    dtype = np.dtype(dtype)
    slot = np.int32 if dtype == np.float32 else np.int64
    rows = np.asarray(rows)
    n = rows.shape[0]
    cap = n if capacity is None else capacity
    buf = np.zeros(3 * cap + 1, dtype=slot)
    buf[0] = n
    body = buf[1 : 1 + 3 * n].reshape(n, 3)
    body[:, 0] = rows
    body[:, 1] = np.asarray(cols)
    body[:, 2] = np.ascontiguousarray(vals, dtype=dtype).view(slot)
    return buf


class Oracle:
    """Plain-C restatement (oracle/qd_oracle.c)."""

    def __init__(self):
        path = os.path.join(_HERE, "libqd_oracle.so")
        if not os.path.exists(path):
            build()
        self.lib = C.CDLL(path)
        L = self.lib
        for name, vt in (("f32", C.c_float), ("f64", C.c_double)):
            getattr(L, f"qdo_build_{name}").restype = C.c_int64
            getattr(L, f"qdo_build_{name}").argtypes = [_VOIDP, C.c_int, C.c_int, C.c_int, _VOIDP, _VOIDP, _VOIDP]
            for fn in (f"qdo_spmv_{name}", f"qdo_spmv_nofma_{name}"):
                getattr(L, fn).restype = None
                getattr(L, fn).argtypes = [C.c_int, C.c_int, C.c_int, _VOIDP, _VOIDP, _VOIDP, _VOIDP, _VOIDP]
            getattr(L, f"qdo_cg_{name}").restype = C.c_int
            getattr(L, f"qdo_cg_{name}").argtypes = [C.c_int, _VOIDP, _VOIDP, _VOIDP, _VOIDP, _VOIDP, C.c_int, vt,
                                                     C.POINTER(C.c_int), C.POINTER(vt)]
        L.qdo_cucg_f32.restype = C.c_int
        L.qdo_cucg_f32.argtypes = [C.c_int, _VOIDP, _VOIDP, _VOIDP, _VOIDP, _VOIDP, C.c_int, C.c_float,
                                   C.POINTER(C.c_float)]
        for fn in ("qdo_network_sort_i32", "qdo_network_sort_f32"):
            getattr(L, fn).restype = C.c_int
            getattr(L, fn).argtypes = [_VOIDP, _VOIDP, C.c_int, C.c_int64]
        L.qdo_stable_sort_u32.restype = None
        L.qdo_stable_sort_u32.argtypes = [_VOIDP, _VOIDP, C.c_int, C.c_int64, C.c_int]
        L.qdo_stable_sort_u64.restype = None
        L.qdo_stable_sort_u64.argtypes = [_VOIDP, _VOIDP, C.c_int64]
        L.qdo_inclusive_scan_i32.restype = None
        L.qdo_inclusive_scan_i32.argtypes = [_VOIDP, C.c_int64]
        for name in ("f32", "f64"):
            getattr(L, f"qdo_direct_{name}").restype = C.c_int
            getattr(L, f"qdo_direct_{name}").argtypes = [C.c_int, _VOIDP, _VOIDP, _VOIDP, C.c_int, _VOIDP, _VOIDP]
            getattr(L, f"qdo_lu_{name}").restype = C.c_int
            getattr(L, f"qdo_lu_{name}").argtypes = [C.c_int, _VOIDP, _VOIDP, _VOIDP, _VOIDP, _VOIDP]

    # -- A16 ------------------------------------------------------------------------------------------------------
    def direct_solve(self, indptr, indices, vals, b, ldlt=False):
        """Simplicial LLT / LDLT of the LOWER triangle of a CSR matrix + solve (small systems: dense workspace).
        Returns (x, ok)."""
        indptr, indices = (np.ascontiguousarray(a, dtype=np.int32) for a in (indptr, indices))
        vals = np.ascontiguousarray(vals)
        name = "f64" if vals.dtype == np.float64 else "f32"
        b = np.ascontiguousarray(b, dtype=vals.dtype)
        x = np.zeros_like(b)
        rc = getattr(self.lib, f"qdo_direct_{name}")(indptr.shape[0] - 1, indptr.ctypes.data, indices.ctypes.data,
                                                     vals.ctypes.data, int(ldlt), b.ctypes.data, x.ctypes.data)
        return x, rc == 0

    def lu_solve(self, indptr, indices, vals, b):
        """LU with partial pivoting of a CSR matrix (dense workspace: small systems) + solve.  Returns (x, ok)."""
        indptr, indices = (np.ascontiguousarray(a, dtype=np.int32) for a in (indptr, indices))
        vals = np.ascontiguousarray(vals)
        name = "f64" if vals.dtype == np.float64 else "f32"
        b = np.ascontiguousarray(b, dtype=vals.dtype)
        x = np.zeros_like(b)
        rc = getattr(self.lib, f"qdo_lu_{name}")(indptr.shape[0] - 1, indptr.ctypes.data, indices.ctypes.data,
                                                 vals.ctypes.data, b.ctypes.data, x.ctypes.data)
        return x, rc == 0

    # -- A3 -------------------------------------------------------------------------------------------------------
    def build_matrix(self, buf: np.ndarray, rows: int, cols: int, row_major: bool = True):
        """-> (outer[int32], inner[int32], values).  row_major=True gives exactly CSR (indptr, indices, values)."""
        f64 = buf.dtype == np.int64
        assert buf.dtype in (np.int32, np.int64)
        count = int(buf[0])
        out_n = rows if row_major else cols
        outer = np.zeros(out_n + 1, np.int32)
        inner = np.zeros(max(count, 1), np.int32)
        vals = np.zeros(max(count, 1), np.float64 if f64 else np.float32)
        fn = self.lib.qdo_build_f64 if f64 else self.lib.qdo_build_f32
        nnz = fn(buf.ctypes.data, rows, cols, int(row_major), outer.ctypes.data, inner.ctypes.data, vals.ctypes.data)
        if nnz < 0:
            raise ValueError("triplet index out of range")
        return outer, inner[:nnz].copy(), vals[:nnz].copy()

    # -- A7 -------------------------------------------------------------------------------------------------------
    def spmv(self, outer, inner, vals, x, rows, cols, row_major=True, fma=True):
        f64 = vals.dtype == np.float64
        x = np.ascontiguousarray(x, dtype=vals.dtype)
        y = np.zeros(rows, vals.dtype)
        name = f"qdo_spmv_{'' if fma else 'nofma_'}{'f64' if f64 else 'f32'}"
        getattr(self.lib, name)(int(row_major), rows, cols, outer.ctypes.data, inner.ctypes.data, vals.ctypes.data,
                                x.ctypes.data, y.ctypes.data)
        return y

    # -- A9 / A10 ---------------------------------------------------------------------------------------------------
    def cg(self, outer, inner, vals, b, max_iters, tol):
        """Eigen Jacobi-PCG semantics -> (x, success, iterations, error)."""
        f64 = vals.dtype == np.float64
        n = outer.shape[0] - 1
        b = np.ascontiguousarray(b, dtype=vals.dtype)
        x = np.zeros(n, vals.dtype)
        it = C.c_int(0)
        if f64:
            err = C.c_double(0)
            info = self.lib.qdo_cg_f64(n, outer.ctypes.data, inner.ctypes.data, vals.ctypes.data, b.ctypes.data,
                                       x.ctypes.data, max_iters, float(np.float32(tol)), C.byref(it), C.byref(err))
        else:
            err = C.c_float(0)
            info = self.lib.qdo_cg_f32(n, outer.ctypes.data, inner.ctypes.data, vals.ctypes.data, b.ctypes.data,
                                       x.ctypes.data, max_iters, tol, C.byref(it), C.byref(err))
        return x, info == 0, it.value, err.value

    def cucg(self, indptr, indices, vals, b, max_iters, tol):
        """Reference CUDA-arch CG semantics (plain CG, absolute r.r <= tol^2) -> (x, iterations, r1)."""
        n = indptr.shape[0] - 1
        b = np.ascontiguousarray(b, dtype=np.float32)
        x = np.zeros(n, np.float32)
        r1 = C.c_float(0)
        k = self.lib.qdo_cucg_f32(n, indptr.ctypes.data, indices.ctypes.data, vals.ctypes.data, b.ctypes.data,
                                  x.ctypes.data, max_iters, tol, C.byref(r1))
        return x, k, r1.value

    # -- A13 / A14 --------------------------------------------------------------------------------------------------
    def network_sort(self, keys, values=None):
        """Exact replay of the reference's odd-even merge network (small N)."""
        keys = np.ascontiguousarray(keys).copy()
        fn = {np.dtype(np.int32): self.lib.qdo_network_sort_i32, np.dtype(np.float32): self.lib.qdo_network_sort_f32}[
            keys.dtype]
        if values is not None:
            values = np.ascontiguousarray(values).copy()
            assert values.dtype.itemsize == 4
            fn(keys.ctypes.data, values.ctypes.data, 1, keys.shape[0])
        else:
            fn(keys.ctypes.data, None, 0, keys.shape[0])
        return keys, values

    def stable_sort(self, keys, values=None):
        """Stable ascending key(/value) sort; keys u32/i32/f32/u64, values 4- or 8-byte items."""
        keys = np.ascontiguousarray(keys).copy()
        if values is not None:
            values = np.ascontiguousarray(values).copy()
        if keys.dtype == np.uint64:
            assert values is None or values.dtype.itemsize == 4
            self.lib.qdo_stable_sort_u64(keys.ctypes.data, None if values is None else values.ctypes.data,
                                         keys.shape[0])
            return keys, values
        kind = {np.dtype(np.uint32): 0, np.dtype(np.int32): 1, np.dtype(np.float32): 2}[keys.dtype]
        words = 1 if values is None else values.dtype.itemsize // 4
        self.lib.qdo_stable_sort_u32(keys.ctypes.data, None if values is None else values.ctypes.data, words,
                                     keys.shape[0], kind)
        return keys, values

    def inclusive_scan(self, arr):
        out = np.ascontiguousarray(arr, dtype=np.int32).copy()
        self.lib.qdo_inclusive_scan_i32(out.ctypes.data, out.shape[0])
        return out


# -- N1: the modern device-wide ops, restated in numpy -----------------------------------------------
# Semantics follow docs/source/user_guide/algorithms.md ("Semantics") and the host references the reference's own tests
# compare against (tests/python/test_algorithms.py:267-695): np.sum / min / max, np.cumsum-style prefix, boolean-mask
# compaction, run-length collapse, np.argsort(kind="stable").  Test infrastructure only.
def _identity(op, dtype):
    """_reduce.py:137-203: 0 for add; largest value (+inf) for min; smallest value (-inf, 0 for unsigned) for max."""
    dtype = np.dtype(dtype)
    if op == "add":
        return dtype.type(0)
    if dtype.kind == "f":
        return dtype.type(np.inf if op == "min" else -np.inf)
    info = np.iinfo(dtype)
    return dtype.type(info.max if op == "min" else info.min)


def algo_reduce(arr, op):
    """out[0] of reduce_{add,min,max}: integer add wraps at the dtype's width; float add is the f64-accumulated sum (the
    device uses a tree order, so tests compare with a tolerance like test_algorithms.py:316-321)."""
    arr = np.asarray(arr)
    if arr.shape[0] == 0:
        return _identity(op, arr.dtype)
    if op == "add":
        if arr.dtype.kind == "f":
            return arr.dtype.type(np.sum(arr.astype(np.float64)))
        with np.errstate(over="ignore"):
            return np.add.reduce(arr, dtype=arr.dtype)
    return arr.min() if op == "min" else arr.max()


def algo_exclusive_scan(arr, op):
    """out[i] = op(arr[0:i]), out[0] = identity (test_algorithms.py:404-451)."""
    arr = np.asarray(arr)
    out = np.empty_like(arr)
    if arr.shape[0] == 0:
        return out
    out[0] = _identity(op, arr.dtype)
    if arr.shape[0] > 1:
        head = arr[:-1]
        if op == "add":
            if arr.dtype.kind == "f":
                out[1:] = np.cumsum(head.astype(np.float64)).astype(arr.dtype)
            else:
                with np.errstate(over="ignore"):
                    out[1:] = np.cumsum(head, dtype=arr.dtype)
        elif op == "min":
            out[1:] = np.minimum.accumulate(head)
        else:
            out[1:] = np.maximum.accumulate(head)
    return out


def algo_select(arr, flags):
    """(compacted prefix, count) of select (test_algorithms.py:454-535)."""
    arr, flags = np.asarray(arr), np.asarray(flags)
    sel = arr[flags == 1]
    return sel, int(sel.shape[0])


def algo_reduce_by_key_add(keys, values):
    """(run keys, run sums, number of runs): a run ends where key[i] != key[i-1] under the key dtype's own `!=` (NaN opens a
    run of its own, +0 == -0).  Integer sums wrap; float sums are accumulated in f64 (compare with a tolerance)."""
    keys, values = np.asarray(keys), np.asarray(values)
    n = keys.shape[0]
    if n == 0:
        return keys[:0], values[:0], 0
    head = np.ones(n, dtype=bool)
    head[1:] = keys[1:] != keys[:-1]
    starts = np.flatnonzero(head)
    if values.dtype.kind == "f":
        sums = np.add.reduceat(values.astype(np.float64), starts).astype(values.dtype)
    else:
        with np.errstate(over="ignore"):
            sums = np.add.reduceat(values, starts, dtype=values.dtype)
    return keys[starts], sums, int(starts.shape[0])


def algo_sort_order(keys, end_bit=None):
    """The permutation a stable ascending radix sort of `keys` produces: order of the sign/IEEE-twiddled bit patterns
    (two's complement -> offset binary; floats -> total order with -0 < +0, -NaN first, +NaN last), restricted to the low
    `end_bit` bits when given (_radix_sort.py:94-137 twiddle, :277-344 passes)."""
    keys = np.ascontiguousarray(keys)
    w = keys.dtype.itemsize * 8
    ut = np.uint32 if w == 32 else np.uint64
    b = keys.view(ut).copy()
    sign = ut(1) << ut(w - 1)
    if keys.dtype.kind == "i":
        b ^= sign
    elif keys.dtype.kind == "f":
        neg = (b & sign) != 0
        b = np.where(neg, ~b, b | sign)
    if end_bit is not None and end_bit < w:
        b &= (ut(1) << ut(end_bit)) - ut(1)
    return np.argsort(b, kind="stable")


class EigenRef:
    """The reference's vendored Eigen behind the reference's call sequence (oracle/eigen_ref.cpp)."""

    def __init__(self, omp: bool = False):
        path = os.path.join(_HERE, "_ref", "libeigen_ref_omp.so" if omp else "libeigen_ref.so")
        self.lib = _load(path)
        if self.lib is None:
            raise FileNotFoundError(path)
        self.omp = omp
        L = self.lib
        L.ref_matrix_build.restype = _VOIDP
        L.ref_matrix_build.argtypes = [_VOIDP, C.c_int, C.c_int, C.c_int, C.c_int]
        L.ref_matrix_free.argtypes = [_VOIDP]
        L.ref_matrix_from_compressed.restype = _VOIDP
        L.ref_matrix_from_compressed.argtypes = [C.c_int, C.c_int, C.c_int64, _VOIDP, _VOIDP, _VOIDP, C.c_int, C.c_int]
        L.ref_matrix_nnz.restype = C.c_int64
        L.ref_matrix_nnz.argtypes = [_VOIDP]
        L.ref_matrix_is_compressed.argtypes = [_VOIDP]
        L.ref_matrix_export.argtypes = [_VOIDP, _VOIDP, _VOIDP, _VOIDP]
        L.ref_matrix_spmv.argtypes = [_VOIDP, _VOIDP, _VOIDP]
        L.ref_cg_solve.restype = C.c_int
        L.ref_cg_solve.argtypes = [_VOIDP, _VOIDP, _VOIDP, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int),
                                   C.POINTER(C.c_double)]
        if hasattr(L, "ref_direct_solve"):
            L.ref_direct_solve.restype = C.c_int
            L.ref_direct_solve.argtypes = [_VOIDP, C.c_int, C.c_int, _VOIDP, _VOIDP]
        L.ref_num_threads.restype = C.c_int
        L.ref_set_num_threads.argtypes = [C.c_int]
        L.ref_eigen_version.restype = C.c_char_p

    @staticmethod
    def available(omp: bool = False) -> bool:
        return os.path.exists(os.path.join(_HERE, "_ref", "libeigen_ref_omp.so" if omp else "libeigen_ref.so"))

    def version(self):
        return self.lib.ref_eigen_version().decode()

    def threads(self):
        return self.lib.ref_num_threads()

    def set_threads(self, n):
        self.lib.ref_set_num_threads(int(n))

    class Matrix:
        def __init__(self, ref, buf, rows, cols, row_major, compressed=None):
            self.ref, self.rows, self.cols, self.row_major = ref, rows, cols, row_major
            if compressed is not None:
                outer, inner, vals = (np.ascontiguousarray(a) for a in compressed)
                assert outer.dtype == np.int32 and inner.dtype == np.int32
                self.f64 = vals.dtype == np.float64
                self.dtype = vals.dtype.type
                self.h = ref.lib.ref_matrix_from_compressed(rows, cols, inner.shape[0], outer.ctypes.data,
                                                            inner.ctypes.data, vals.ctypes.data, int(self.f64),
                                                            int(row_major))
                return
            self.f64 = buf.dtype == np.int64
            self.dtype = np.float64 if self.f64 else np.float32
            self.h = ref.lib.ref_matrix_build(buf.ctypes.data, rows, cols, int(self.f64), int(row_major))

        def __del__(self):
            if getattr(self, "h", None):
                self.ref.lib.ref_matrix_free(self.h)
                self.h = None

        def export(self):
            nnz = self.ref.lib.ref_matrix_nnz(self.h)
            outer = np.zeros((self.rows if self.row_major else self.cols) + 1, np.int32)
            inner = np.zeros(max(nnz, 1), np.int32)
            vals = np.zeros(max(nnz, 1), self.dtype)
            self.ref.lib.ref_matrix_export(self.h, outer.ctypes.data, inner.ctypes.data, vals.ctypes.data)
            return outer, inner[:nnz].copy(), vals[:nnz].copy()

        def spmv(self, x):
            x = np.ascontiguousarray(x, dtype=self.dtype)
            y = np.zeros(self.rows, self.dtype)
            self.ref.lib.ref_matrix_spmv(self.h, x.ctypes.data, y.ctypes.data)
            return y

        def cg(self, b, max_iters, tol, allow_row_major=False):
            b = np.ascontiguousarray(b, dtype=self.dtype)
            x = np.zeros(self.cols, self.dtype)
            it, err = C.c_int(0), C.c_double(0)
            info = self.ref.lib.ref_cg_solve(self.h, b.ctypes.data, x.ctypes.data, max_iters, tol,
                                             int(allow_row_major), C.byref(it), C.byref(err))
            if info < 0:
                raise RuntimeError("SparseCG accepts col_major matrices only (conjugate_gradient.h:40)")
            return x, info == 0, it.value, err.value

        def direct_solve(self, b, solver_type="LLT", ordering="AMD"):
            """EigenSparseSolver analyze_pattern + factorize + solve (sparse_solver.cpp:86-107).  Returns (x, ok)."""
            b = np.ascontiguousarray(b, dtype=self.dtype)
            x = np.zeros(self.cols, self.dtype)
            info = self.ref.lib.ref_direct_solve(self.h, {"LLT": 0, "LDLT": 1, "LU": 2}[solver_type],
                                                 {"AMD": 0, "COLAMD": 1}[ordering], b.ctypes.data, x.ctypes.data)
            if info == -1:
                raise RuntimeError("the direct solvers are declared on col_major matrices (sparse_solver.h:9-20)")
            return x, info == 0

    def matrix(self, buf, rows, cols, row_major=True):
        return EigenRef.Matrix(self, buf, rows, cols, row_major)

    def matrix_from_compressed(self, outer, inner, vals, rows, cols, row_major=True):
        """Eigen matrix over existing compressed arrays (copied); same arithmetic afterwards."""
        return EigenRef.Matrix(self, None, rows, cols, row_major, compressed=(outer, inner, vals))
