"""Objects carrying the `_qd_core` (nanobind) method names of the sparse path
(quadrants/python/export_lang.cpp:518-531, 1166-1295), implemented over the C ABI.  python/quadrants/linalg/*.py can run
against these unmodified; quadrants_b200.linalg is the same thin layer re-created."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .lang import Ndarray, QuadrantsRuntimeError, f32, f64, to_dtype, _require_prog


def _dt(dtype) -> int:
    d = to_dtype(dtype)
    if d is f32:
        return _lib.F32
    if d is f64:
        return _lib.F64
    raise _lib.QdspError(_lib.QDSP_OK + 1, "Unsupported sparse matrix data type!")


def _stream(prog=None) -> int:
    return (prog or _require_prog()).stream()


class _DevView:
    """Exposes raw device memory of the library to torch through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr, n, typestr, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3,
                                         "strides": None}


def device_tensor(ptr: int, n: int, np_dtype, owner):
    import torch

    if n == 0:
        return torch.empty(0, dtype=getattr(torch, np.dtype(np_dtype).name), device="cuda")
    return torch.as_tensor(_DevView(ptr, n, np.dtype(np_dtype).str, owner), device="cuda")


class SparseMatrixBuilder:
    """export_lang.cpp:1167-1177 over quadrants/program/sparse_matrix.cpp:70-228."""

    def __init__(self, rows, cols, max_num_triplets, dt=f32, storage_format="col_major"):
        self.rows, self.cols, self.max_num_triplets = int(rows), int(cols), int(max_num_triplets)
        self.dtype = to_dtype(dt)
        self.storage_format = storage_format
        self._h = C.c_void_p()
        self._lib = _lib.load()
        _require_prog()
        _lib.check(self._lib.qdsp_builder_create(self.rows, self.cols, self.max_num_triplets, _dt(dt),
                                                 storage_format.encode(), C.byref(self._h)))

    # the reference allocates in create_ndarray(prog); here the ctor already did (slot 0 zeroed)
    def create_ndarray(self, prog):
        return None

    def delete_ndarray(self, prog):
        self._destroy()

    def _destroy(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.qdsp_builder_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    def get_ndarray_data_ptr(self) -> int:
        p = C.c_void_p()
        _lib.check(self._lib.qdsp_builder_buffer_ptr(self._h, C.byref(p)))
        return p.value

    def get_addr(self) -> int:
        return self._h.value

    def num_triplets(self) -> int:
        n = C.c_int64()
        _lib.check(self._lib.qdsp_builder_count(self._h, _stream(), C.byref(n)))
        return n.value

    def buffer_tensor(self):
        """The triplet buffer [count | (row, col, bits)...] as a torch tensor of slot-width integers (zero copy)."""
        slot = np.int32 if self.dtype is f32 else np.int64
        return device_tensor(self.get_ndarray_data_ptr(), 3 * self.max_num_triplets + 1, slot, self)

    def _dev(self, x, torch_dtype):
        import torch

        if not torch.is_tensor(x):
            x = torch.as_tensor(np.asarray(x))
        dev = f"cuda:{_require_prog().config().device}"
        return x.to(device=dev, dtype=torch_dtype).reshape(-1).contiguous()

    def _reserve(self, k):
        """Host-side bound on slot 0, so an overflow raises at the call like the old synchronous path did without
        reading the count back every time; only when the bound is hit is the true device count fetched."""
        bound = getattr(self, "_appended", 0) + k
        if bound > self.max_num_triplets:
            bound = self.num_triplets() + k
            if bound > self.max_num_triplets:
                raise QuadrantsRuntimeError(f"builder overflow: {bound} triplets > max_num_triplets "
                                            f"{self.max_num_triplets}")
        self._appended = bound

    def insert_triplets(self, rows, cols, vals, negate=False):
        """What a JIT fill kernel does through insert_triplet_f32/f64 (internal_functions.h:12-39), in bulk and without
        its per-triplet atomic: qdsp_builder_append writes (row, col, bits(value)) behind the triplets already there,
        in array order, and bumps slot 0 on the stream.  `A[i, j] -= v` inserts -v."""
        import torch

        rows, cols = self._dev(rows, torch.int32), self._dev(cols, torch.int32)
        vals = self._dev(vals, self.dtype.torch_dtype)
        k = rows.shape[0]
        if cols.shape[0] != k or vals.shape[0] != k:
            raise QuadrantsRuntimeError("insert_triplets: rows, cols and vals differ in length")
        if k == 0:
            return
        self._reserve(k)
        _lib.check(self._lib.qdsp_builder_append(self._h, rows.data_ptr(), cols.data_ptr(), vals.data_ptr(), k,
                                                 1 if negate else 0, _stream()))

    def insert_elements(self, nodes, mats):
        """Element-wise assembly (the FEM fill pattern, tests/python/test_implicit_fem.py:179-204): nodes [E, k] int,
        mats [E, k, k]; element e adds mats[e, a, b] at (nodes[e, a], nodes[e, b])."""
        import torch

        nodes_t = nodes if torch.is_tensor(nodes) else torch.as_tensor(np.asarray(nodes))
        if nodes_t.dim() != 2:
            raise QuadrantsRuntimeError("insert_elements: nodes must be [n_elems, k]")
        e, k = int(nodes_t.shape[0]), int(nodes_t.shape[1])
        nodes_d = self._dev(nodes_t, torch.int32)
        mats_d = self._dev(mats, self.dtype.torch_dtype)
        if mats_d.shape[0] != e * k * k:
            raise QuadrantsRuntimeError("insert_elements: mats must be [n_elems, k, k]")
        if e == 0:
            return
        self._reserve(e * k * k)
        _lib.check(self._lib.qdsp_builder_append_elements(self._h, nodes_d.data_ptr(), mats_d.data_ptr(), e, k,
                                                          _stream()))

    def fill_from_host(self, triplet_buffer: np.ndarray):
        """Copies a host buffer already in the builder layout (slot 0 = count)."""
        slot = np.int32 if self.dtype is f32 else np.int64
        b = np.ascontiguousarray(triplet_buffer, dtype=slot)
        n = int(b[0])
        _lib.check(self._lib.qdsp_builder_fill_from_host(self._h, b[1:].ctypes.data, n, _stream()))
        self._appended = n

    def print_triplets_cuda(self):
        _lib.check(self._lib.qdsp_builder_print_triplets(self._h))

    print_triplets_eigen = print_triplets_cuda

    def build_cuda(self):
        out = C.c_void_p()
        _lib.check(self._lib.qdsp_builder_build(self._h, _stream(), C.byref(out)))
        self._appended = 0  # build() clears the buffer
        return CuSparseMatrix._wrap(out)

    build = build_cuda


class CuSparseMatrix:
    """export_lang.cpp:1179-1189,1217-1228 over quadrants/program/sparse_matrix.cpp:344-780."""

    def __init__(self, rows=0, cols=0, dt=f32):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self._rows, self._cols, self._dtype = int(rows), int(cols), to_dtype(dt)

    @classmethod
    def _wrap(cls, handle):
        self = cls.__new__(cls)
        self._lib = _lib.load()
        self._h = handle
        r, c, nnz, dt = C.c_int(), C.c_int(), C.c_int64(), C.c_int()
        _lib.check(self._lib.qdsp_csr_info(handle, C.byref(r), C.byref(c), C.byref(nnz), C.byref(dt)))
        self._rows, self._cols, self._dtype = r.value, c.value, (f64 if dt.value == _lib.F64 else f32)
        return self

    def __del__(self):
        try:
            if self._h is not None and self._h.value:
                self._lib.qdsp_csr_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _need(self):
        if not self._h.value:
            raise _lib.QdspError(3, "sparse matrix has not been built")
        return self._h

    def num_rows(self):
        return self._rows

    def num_cols(self):
        return self._cols

    def get_data_type(self):
        return self._dtype

    def nnz(self) -> int:
        n = C.c_int64()
        _lib.check(self._lib.qdsp_csr_info(self._need(), None, None, C.byref(n), None))
        return n.value

    def build_csr_from_coo(self, coo_row, coo_col, coo_val, nnz=None):
        """CuSparseMatrix::build_csr_from_coo: device int32 row/col + value arrays (torch tensors or Ndarrays)."""
        t = [a.tensor if isinstance(a, Ndarray) else a for a in (coo_row, coo_col, coo_val)]
        n = int(t[0].shape[0]) if nnz is None else int(nnz)
        out = C.c_void_p()
        _lib.check(self._lib.qdsp_csr_from_device_coo(self._rows, self._cols, n, t[0].data_ptr(), t[1].data_ptr(),
                                                      t[2].data_ptr(), _dt(self._dtype), _stream(), C.byref(out)))
        if self._h.value:
            self._lib.qdsp_csr_destroy(self._h)
        self._h = out

    def csr_tensors(self):
        """(indptr, indices, values) as zero-copy torch views of csr_row_ptr_/csr_col_ind_/csr_val_."""
        ip, ix, v = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _lib.check(self._lib.qdsp_csr_export(self._need(), C.byref(ip), C.byref(ix), C.byref(v)))
        n = self.nnz()
        return (device_tensor(ip.value, self._rows + 1, np.int32, self), device_tensor(ix.value, n, np.int32, self),
                device_tensor(v.value, n, self._dtype.np_dtype, self))

    def to_scipy_arrays(self):
        ip, ix, v = self.csr_tensors()
        return ip.cpu().numpy(), ix.cpu().numpy(), v.cpu().numpy()

    def spmv_config(self, compress):
        """qdsp_csr_spmv_config: 1 / 0 / -1 (default) -- one-byte diagonal codes instead of column indices in the SpMV."""
        _lib.check(self._lib.qdsp_csr_spmv_config(self._need(), int(compress)))

    def spmv_info(self):
        ib, path, hot = C.c_int(), C.c_int(), C.c_int()
        _lib.check(self._lib.qdsp_csr_spmv_info(self._need(), _stream(), C.byref(ib), C.byref(path), C.byref(hot)))
        return {"index_bytes_per_nnz": ib.value, "path": {1: "uniform", 2: "stream"}.get(path.value, "none"),
                "hot_columns": hot.value}

    def spmv(self, prog, x, y):
        """nd_spmv(Program*, const Ndarray& x, const Ndarray& y)"""
        _lib.check(self._lib.qdsp_csr_spmv(self._need(), x.data_ptr(), y.data_ptr(), _stream(prog)))

    def mat_vec_mul_host(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=self._dtype.np_dtype)
        y = np.empty(self._rows, self._dtype.np_dtype)
        _lib.check(self._lib.qdsp_csr_spmv_host(self._need(), x.ctypes.data, y.ctypes.data, _stream()))
        return y

    def _binary(self, fn, *args):
        out = C.c_void_p()
        _lib.check(fn(*args, _stream(), C.byref(out)))
        return CuSparseMatrix._wrap(out)

    def __add__(self, other):
        return self._binary(self._lib.qdsp_csr_add, self._need(), other._need(), 1.0, 1.0)

    def __sub__(self, other):
        return self._binary(self._lib.qdsp_csr_add, self._need(), other._need(), 1.0, -1.0)

    def __mul__(self, s):
        return self._binary(self._lib.qdsp_csr_scale, self._need(), float(s))

    __rmul__ = __mul__

    def matmul(self, other):
        return self._binary(self._lib.qdsp_csr_matmul, self._need(), other._need())

    def transpose(self):
        return self._binary(self._lib.qdsp_csr_transpose, self._need())

    def get_element(self, r, c):
        v = C.c_double()
        _lib.check(self._lib.qdsp_csr_get_element(self._need(), int(r), int(c), C.byref(v)))
        return v.value

    def set_element(self, r, c, v):
        raise _lib.QdspError(4, "set_element is not implemented for CUDA sparse matrices (QD_NOT_IMPLEMENTED)")

    def to_string(self):
        ip, ix, v = self.to_scipy_arrays()
        dense = np.zeros((self._rows, self._cols), np.float32)
        for r in range(self._rows):
            dense[r, ix[ip[r]:ip[r + 1]]] = v[ip[r]:ip[r + 1]]
        return "\n".join("[" + ", ".join(f"{x:.4g}" for x in row) + "]" for row in dense)

    def mmwrite(self, filename):
        _lib.check(self._lib.qdsp_csr_mmwrite(self._need(), str(filename).encode()))


SparseMatrix = CuSparseMatrix


class _CgBase:
    def __init__(self, A, max_iters, tol, verbose, mode):
        self._lib = _lib.load()
        self._A = A  # keep alive: the C++ solver holds a reference like CG::A_ (conjugate_gradient.h:60)
        self._h = C.c_void_p()
        self.iterations, self.error, self._success = 0, 0.0, False
        _lib.check(self._lib.qdsp_cg_create(A._need(), int(max_iters), float(tol), mode, int(bool(verbose)),
                                            C.byref(self._h)))

    def __del__(self):
        try:
            if self._h.value:
                self._lib.qdsp_cg_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _solve(self, x, b, prog=None):
        it, err, ok = C.c_int(), C.c_double(), C.c_int()
        _lib.check(self._lib.qdsp_cg_solve(self._h, x.data_ptr(), b.data_ptr(), _stream(prog), C.byref(it),
                                           C.byref(err), C.byref(ok)))
        self.iterations, self.error, self._success = it.value, err.value, bool(ok.value)

    def run_fixed(self, x, b, iters):
        _lib.check(self._lib.qdsp_cg_run_fixed(self._h, x.data_ptr(), b.data_ptr(), int(iters), _stream()))


class CUCG(_CgBase):
    """CUCG (export_lang.cpp:1294): plain CG, absolute tolerance, x is the zero-filled ndarray Python allocated."""

    def __init__(self, A, max_iters, tol, verbose):
        super().__init__(A, max_iters, np.float32(tol), verbose, _lib.CG_CUCG)

    def solve(self, prog, x, b):
        self._solve(x, b, prog)


def make_cucg_solver(A, max_iters, tol, verbose):
    return CUCG(A, max_iters, tol, verbose)


class _EigenCg(_CgBase):
    """CGf / CGd (export_lang.cpp:1269-1292): Eigen's Jacobi-PCG semantics, computed on the GPU."""
    _dtype = f32

    def __init__(self, A, max_iters, tol, verbose):
        # tol travels through a float ctor argument (conjugate_gradient.h:12,64)
        super().__init__(A, max_iters, float(np.float32(tol)), verbose, _lib.CG_EIGEN_JACOBI)
        n = A.num_rows()
        self._x = Ndarray(self._dtype, n)
        self._b = Ndarray(self._dtype, n)

    def set_b(self, b):
        self._b.from_numpy(np.asarray(b))

    def set_x(self, x):
        self._x.from_numpy(np.asarray(x))  # ignored by solve(), like cg.solve(b_) in the reference

    def set_b_ndarray(self, prog, b):
        self._b = b

    def set_x_ndarray(self, prog, x):
        pass  # the guess is ignored (conjugate_gradient.h:44: x_ = cg.solve(b_))

    def solve(self):
        self._solve(self._x, self._b)

    def get_x(self):
        return self._x.to_numpy()

    def get_x_ndarray(self):
        return self._x

    def is_success(self):
        return self._success


class CGf(_EigenCg):
    _dtype = f32


class CGd(_EigenCg):
    _dtype = f64


def make_float_cg_solver(A, max_iters, tol, verbose):
    return CGf(A, max_iters, tol, verbose)


def make_double_cg_solver(A, max_iters, tol, verbose):
    return CGd(A, max_iters, tol, verbose)


def make_cu_sparse_matrix(rows, cols, dt):
    return CuSparseMatrix(rows, cols, dt)


class CuSparseSolver:
    """CuSparseSolver (export_lang.cpp make_cusparse_solver; quadrants/program/sparse_solver.h:81-136) over solver.cu:
    host minimum-degree ordering + symbolic analysis, numeric factorisation and both triangular solves on the device."""

    def __init__(self, dt, solver_type, ordering):
        if solver_type not in _lib.SOLVER_TYPES:
            raise QuadrantsRuntimeError(f"Not supported sparse solver type: {solver_type}")
        self._lib = _lib.load()
        self._dt = to_dtype(dt)
        self._h = C.c_void_p()
        self._A = None
        _lib.check(self._lib.qdsp_solver_create(_dt(dt), _lib.SOLVER_TYPES[solver_type],
                                                _lib.SOLVER_ORDERINGS.get(ordering, 0), C.byref(self._h)))

    def __del__(self):
        try:
            if self._h.value:
                self._lib.qdsp_solver_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def analyze_pattern(self, sm):
        self._A = sm
        _lib.check(self._lib.qdsp_solver_analyze_pattern(self._h, sm._need(), _stream()))

    def factorize(self, sm):
        ok, piv = C.c_int(), C.c_int()
        _lib.check(self._lib.qdsp_solver_factorize(self._h, sm._need(), _stream(), C.byref(ok), C.byref(piv)))
        if not ok.value:
            # the reference: QD_ASSERT(singularity == -1) after csrcholZeroPivot / csrluZeroPivotHost
            # (sparse_solver.cpp:326-327, 359-360)
            raise QuadrantsRuntimeError(f"SparseSolver factorisation failed: zero pivot at (permuted) index {piv.value}")

    def compute(self, sm):
        self.analyze_pattern(sm)
        self.factorize(sm)
        return self.info()

    def solve_rf(self, prog, sm, b, x):
        n = sm.num_rows()
        if b.tensor.shape[0] != n or x.tensor.shape[0] != n:
            raise QuadrantsRuntimeError(f"SparseSolver.solve: vectors of length {b.tensor.shape[0]} / {x.tensor.shape[0]} "
                                        f"for a system of {n} rows")
        _lib.check(self._lib.qdsp_solver_solve(self._h, b.data_ptr(), x.data_ptr(), _stream(prog)))

    def solve(self, b: np.ndarray) -> np.ndarray:
        """EigenSparseSolver::solve(b) with a numpy right-hand side (export_lang.cpp solve overloads)."""
        npdt = np.float64 if self._dt is f64 else np.float32
        bb = np.ascontiguousarray(b, dtype=npdt).reshape(-1)
        if self._A is None or bb.shape[0] != self._A.num_rows():
            raise QuadrantsRuntimeError(
                f"SparseSolver.solve: right-hand side of length {bb.shape[0]} for a system of "
                f"{None if self._A is None else self._A.num_rows()} rows")
        x = np.empty_like(bb)
        _lib.check(self._lib.qdsp_solver_solve_host(self._h, bb.ctypes.data, x.ctypes.data, _stream()))
        return x

    def info(self) -> bool:
        ok = C.c_int()
        _lib.check(self._lib.qdsp_solver_info(self._h, C.byref(ok)))
        return bool(ok.value)

    def stats(self):
        nnz, h, fl = C.c_int64(), C.c_int(), C.c_double()
        _lib.check(self._lib.qdsp_solver_stats(self._h, C.byref(nnz), C.byref(h), C.byref(fl)))
        return {"factor_nnz": nnz.value, "etree_height": h.value, "flops": fl.value}


def make_cusparse_solver(dt, solver_type, ordering):
    return CuSparseSolver(dt, solver_type, ordering)


make_sparse_solver = make_cusparse_solver  # the arch=x64 factory name; there is one (device) implementation here


def solver_analyze_host(indptr, indices, natural_order=False):
    """Pure host: ordering + symbolic factorisation of a square CSR pattern (no GPU needed).  Returns a dict with perm
    (new -> old), parent (elimination tree), colptr / rowidx (factor pattern by columns) and etree_height."""
    lib = _lib.load()
    ip = np.ascontiguousarray(indptr, dtype=np.int32)
    ix = np.ascontiguousarray(indices, dtype=np.int32)
    n = ip.shape[0] - 1
    perm, parent, colptr = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n + 1, np.int32)
    nnz, h = C.c_int64(), C.c_int()
    _lib.check(lib.qdsp_solver_analyze_host(n, ip.ctypes.data, ix.ctypes.data, int(natural_order), perm.ctypes.data,
                                            parent.ctypes.data, colptr.ctypes.data, None, 0, C.byref(nnz), C.byref(h)))
    rowidx = np.empty(nnz.value, np.int32)
    _lib.check(lib.qdsp_solver_analyze_host(n, ip.ctypes.data, ix.ctypes.data, int(natural_order), perm.ctypes.data,
                                            parent.ctypes.data, colptr.ctypes.data, rowidx.ctypes.data, nnz.value,
                                            C.byref(nnz), C.byref(h)))
    return {"perm": perm, "parent": parent, "colptr": colptr, "rowidx": rowidx, "factor_nnz": nnz.value,
            "etree_height": h.value}
