"""qd.linalg.SparseCG (python/quadrants/linalg/sparse_cg.py:12-59) on arch=qd.cuda."""
from __future__ import annotations

import numpy as np

from .. import _core
from ..lang import Arch, Ndarray, QuadrantsRuntimeError, ScalarNdarray, f32, f64, get_runtime


class SparseCG:
    """Conjugate-gradient solver built for SparseMatrix: solves A x = b.

    Args follow the reference: A, b, x0 (ignored by the reference on every arch: Eigen's solve(b) starts from zero,
    conjugate_gradient.h:44; the CUDA path allocates a fresh zero ndarray, sparse_cg.py:53), max_iter=50, atol=1e-6.

    semantics:
      "eigen" (default) -- the reference's Eigen path: Jacobi-preconditioned, `atol` is RELATIVE to |b|, returns
                           (x, success) with success = (|r|/|b| <= atol).  This is the contract BASELINE.json checks.
      "cucg"            -- the reference's arch=cuda behaviour bit for bit in meaning: unpreconditioned, stop when
                           r.r <= atol^2 (absolute), f32 only, returns (x, True) unconditionally.
    """

    def __init__(self, A, b, x0=None, max_iter=50, atol=1e-6, semantics="eigen"):
        self.dtype = A.dtype
        prog = get_runtime().prog
        if prog is None or prog.config().arch != Arch.cuda:
            raise QuadrantsRuntimeError(f"Unsupported CG arch: {None if prog is None else prog.config().arch}")
        self.qd_arch = prog.config().arch
        self.matrix = A
        self.b = b
        self.semantics = semantics
        if semantics == "cucg":
            if self.dtype is not f32:
                raise QuadrantsRuntimeError(f"Unsupported CG dtype: {self.dtype}")
            self.cg_solver = _core.make_cucg_solver(A.matrix, max_iter, atol, False)
        elif semantics == "eigen":
            if self.dtype is f32:
                self.cg_solver = _core.make_float_cg_solver(A.matrix, max_iter, atol, False)
            elif self.dtype is f64:
                self.cg_solver = _core.make_double_cg_solver(A.matrix, max_iter, atol, False)
            else:
                raise QuadrantsRuntimeError(f"Unsupported CG dtype: {self.dtype}")
            if isinstance(b, Ndarray):
                self.cg_solver.set_b_ndarray(prog, b.arr)
            elif isinstance(b, np.ndarray):
                self.cg_solver.set_b(b)
            else:
                raise QuadrantsRuntimeError(f"Unsupported CG RHS type: {type(b)}")
        else:
            raise QuadrantsRuntimeError(f"Unknown CG semantics: {semantics}")

    def solve(self):
        if self.semantics == "cucg":
            if isinstance(self.b, Ndarray):
                x = ScalarNdarray(self.b.dtype, [self.matrix.m])
                self.cg_solver.solve(get_runtime().prog, x.arr, self.b.arr)
                return x, True
            raise QuadrantsRuntimeError(f"Unsupported CG RHS type: {type(self.b)}")
        self.cg_solver.solve()
        x = self.cg_solver.get_x_ndarray() if isinstance(self.b, Ndarray) else self.cg_solver.get_x()
        return x, self.cg_solver.is_success()

    @property
    def iterations(self):
        return self.cg_solver.iterations

    @property
    def error(self):
        return self.cg_solver.error
