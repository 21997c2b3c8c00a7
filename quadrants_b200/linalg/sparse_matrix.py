"""qd.linalg.SparseMatrix / SparseMatrixBuilder with the reference's signatures, defaults and error behaviour
(python/quadrants/linalg/sparse_matrix.py:15-303), on arch=qd.cuda only."""
from __future__ import annotations

import numpy as np

from .. import _core
from ..lang import Arch, Ndarray, QuadrantsRuntimeError, ScalarNdarray, f32, get_runtime, to_dtype


def _same_shape(a, b):
    assert a.n == b.n and a.m == b.m, (
        f"Dimension mismatch between sparse matrices ({a.n}, {a.m}) and ({b.n}, {b.m})")


class SparseMatrix:
    """Quadrants's Sparse Matrix class (reference: sparse_matrix.py:15-238)."""

    def __init__(self, n=None, m=None, sm=None, dtype=f32, storage_format="col_major"):
        self.dtype = to_dtype(dtype)
        if sm is None:
            self.n = n
            self.m = m if m else n
            self.matrix = _core.make_cu_sparse_matrix(self.n, self.m, self.dtype)
        else:
            self.n, self.m, self.matrix = sm.num_rows(), sm.num_cols(), sm

    # -- algebra (CuSparseMatrix binds + - *float float* matmul transpose; no in-place and no Hadamard on CUDA) --------
    def __add__(self, other):
        _same_shape(self, other)
        return SparseMatrix(sm=self.matrix + other.matrix, dtype=self.dtype)

    def __iadd__(self, other):
        _same_shape(self, other)
        self.matrix = self.matrix + other.matrix
        return self

    def __sub__(self, other):
        _same_shape(self, other)
        return SparseMatrix(sm=self.matrix - other.matrix, dtype=self.dtype)

    def __isub__(self, other):
        _same_shape(self, other)
        self.matrix = self.matrix - other.matrix
        return self

    def __mul__(self, other):
        if isinstance(other, float):
            return SparseMatrix(sm=other * self.matrix, dtype=self.dtype)
        if isinstance(other, SparseMatrix):
            raise QuadrantsRuntimeError("element-wise sparse product is not bound for CUDA sparse matrices")
        return None

    def __rmul__(self, other):
        if isinstance(other, float):
            return SparseMatrix(sm=self.matrix * other, dtype=self.dtype)
        return None

    def transpose(self):
        return SparseMatrix(sm=self.matrix.transpose(), dtype=self.dtype)

    def __matmul__(self, other):
        if isinstance(other, SparseMatrix):
            assert self.m == other.n, (
                f"Dimension mismatch between sparse matrices ({self.n}, {self.m}) and ({other.n}, {other.m})")
            return SparseMatrix(sm=self.matrix.matmul(other.matrix), dtype=self.dtype)
        if isinstance(other, Ndarray):
            if self.m != other.shape[0]:
                raise QuadrantsRuntimeError(
                    f"Dimension mismatch between sparse matrix ({self.n}, {self.m}) and vector ({other.shape})")
            res = ScalarNdarray(dtype=other.dtype, arr_shape=(self.n,))
            self.matrix.spmv(get_runtime().prog, other.arr, res.arr)
            return res
        if isinstance(other, np.ndarray):
            # host operand: the reference supports this on the CPU arch only (mat_vec_mul); here it is the
            # host-buffer entry point of the C ABI (H2D, SpMV, D2H)
            assert self.m == other.shape[0], (
                f"Dimension mismatch between sparse matrix ({self.n}, {self.m}) and vector ({other.shape})")
            return self.matrix.mat_vec_mul_host(other)
        raise QuadrantsRuntimeError(
            f"Sparse matrix-matrix/vector multiplication does not support {type(other)} for now. Supported types are "
            "SparseMatrix, qd.field, and numpy ndarray.")

    def __getitem__(self, indices):
        return self.matrix.get_element(indices[0], indices[1])

    def __setitem__(self, indices, value):
        self.matrix.set_element(indices[0], indices[1], value)

    def __str__(self):
        return self.matrix.to_string()

    __repr__ = __str__

    @property
    def shape(self):
        return (self.n, self.m)

    def build_from_ndarray(self, ndarray):
        """N x 3 float ndarray of (row, col, value) (reference: sparse_matrix.py:182-216, sparse_matrix.cpp:320-342)."""
        if not isinstance(ndarray, Ndarray):
            raise QuadrantsRuntimeError(
                "Sparse matrix only supports building from [qd.ndarray, qd.Vector.ndarray, qd.Matrix.ndarray]")
        t = ndarray.tensor.reshape(-1)
        if t.shape[0] % 3 != 0:
            raise QuadrantsRuntimeError("The number of ndarray elements must have a length that is divisible by 3.")
        import torch

        trip = t.view(-1, 3)
        self.matrix.build_csr_from_coo(trip[:, 0].to(torch.int32).contiguous(), trip[:, 1].to(torch.int32).contiguous(),
                                       trip[:, 2].to(self.dtype.torch_dtype).contiguous())

    def mmwrite(self, filename):
        self.matrix.mmwrite(filename)


class _Entry:
    """`A[i, j] += v` in Python scope (the reference does this inside a JIT kernel via SparseMatrixEntry._augassign,
    python/quadrants/lang/kernel_arguments.py:35-49: += inserts v, -= inserts -v, any other operator asserts)."""

    def __init__(self, builder, i, j):
        self.builder, self.i, self.j = builder, i, j

    def __iadd__(self, v):
        self.builder.ptr.insert_triplets([self.i], [self.j], [v])
        return self

    def __isub__(self, v):
        self.builder.ptr.insert_triplets([self.i], [self.j], [v], negate=True)
        return self


class SparseMatrixBuilder:
    """reference: sparse_matrix.py:219-300"""

    def __init__(self, num_rows=None, num_cols=None, max_num_triplets=0, dtype=f32, storage_format="col_major"):
        self.num_rows = num_rows
        self.num_cols = num_cols if num_cols else num_rows
        self.dtype = to_dtype(dtype)
        self.ptr = None
        if num_rows is not None:
            prog = get_runtime().prog
            if prog is None or prog.config().arch != Arch.cuda:
                raise QuadrantsRuntimeError("SparseMatrix only supports CPU and CUDA for now.")
            self.ptr = _core.SparseMatrixBuilder(num_rows, self.num_cols, max_num_triplets, self.dtype, storage_format)
            self.ptr.create_ndarray(prog)

    def _get_addr(self):
        return self.ptr.get_addr()

    def _get_ndarray_addr(self):
        return self.ptr.get_ndarray_data_ptr()

    def __getitem__(self, ij):
        return _Entry(self, ij[0], ij[1])

    def __setitem__(self, ij, entry):
        if not isinstance(entry, _Entry):
            raise QuadrantsRuntimeError("Only operations '+=' and '-=' are supported on sparse matrices.")

    def insert_triplets(self, rows, cols, vals):
        """Vectorised `A[rows[k], cols[k]] += vals[k]` (device-side append into the builder buffer)."""
        self.ptr.insert_triplets(rows, cols, vals)

    def insert_elements(self, nodes, mats):
        """Element-wise assembly: `A[nodes[e, a], nodes[e, b]] += mats[e, a, b]` for every element e (the FEM fill loop
        of the reference's tests/python/test_implicit_fem.py:179-204 as one device-side append)."""
        self.ptr.insert_elements(nodes, mats)

    def print_triplets(self):
        self.ptr.print_triplets_cuda()

    def build(self, dtype=f32, _format="CSR"):
        """Create a sparse matrix using the triplets.  (The reference refuses f64 on CUDA -- "CUDA sparse matrix only
        supports f32", sparse_matrix.py:292-293; f64 is implemented here.)"""
        return SparseMatrix(sm=self.ptr.build_cuda(), dtype=self.dtype)

    def __del__(self):
        rt = get_runtime()
        if self.ptr is not None and rt is not None and rt.prog is not None:
            self.ptr.delete_ndarray(rt.prog)


__all__ = ["SparseMatrix", "SparseMatrixBuilder"]
