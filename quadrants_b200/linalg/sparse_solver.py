"""qd.linalg.SparseSolver (python/quadrants/linalg/sparse_solver.py:14-123): the interface of the reference's direct
solvers (Eigen SimplicialLLT / LDLT / SparseLU on x64, cuSOLVER on cuda; quadrants/program/sparse_solver.cpp:74-558).

SURVEY 8(a) row A16: the surface is retained, the factorisations are NOT implemented -- they are a different problem
class (fill-in and latency bound, not HBM bound) and in no BASELINE config.
Constructing a solver validates its arguments exactly like the reference; compute / analyze_pattern / factorize /
solve raise the error the reference raises for an unimplemented path (QD_NOT_IMPLEMENTED -> RuntimeError).  There is
deliberately no fallback to an iterative or host solver behind this name: use qd.linalg.SparseCG for SPD systems.
"""
from __future__ import annotations

from ..lang import Arch, QuadrantsRuntimeError, f32, get_runtime
from .sparse_matrix import SparseMatrix

_NOT_IMPLEMENTED = ("SparseSolver direct factorisations (LLT / LDLT / LU) are not implemented in quadrants_b200 "
                    "(SURVEY 8(a) A16: interface retained, implementation deferred); use qd.linalg.SparseCG")


class SparseSolver:
    def __init__(self, dtype=f32, solver_type="LLT", ordering="AMD"):
        self.matrix = None
        self.dtype = dtype
        solver_type_list = ["LLT", "LDLT", "LU"]
        solver_ordering = ["AMD", "COLAMD"]
        if not (solver_type in solver_type_list and ordering in solver_ordering):
            raise QuadrantsRuntimeError(
                f"The solver type {solver_type} with {ordering} is not supported for now. Only {solver_type_list} with "
                f"{solver_ordering} are supported.")
        prog = get_runtime().prog
        assert prog is not None and prog.config().arch == Arch.cuda, "SparseSolver only supports CPU and CUDA for now."
        self.solver_type, self.ordering = solver_type, ordering

    @staticmethod
    def _type_assert(sparse_matrix):
        raise QuadrantsRuntimeError(
            f"The parameter type: {type(sparse_matrix)} is not supported in linear solvers for now.")

    def _check(self, sparse_matrix):
        if not isinstance(sparse_matrix, SparseMatrix):
            self._type_assert(sparse_matrix)
        if sparse_matrix.dtype != self.dtype:
            raise QuadrantsRuntimeError(
                f"The SparseSolver's dtype {self.dtype} is not consistent with the SparseMatrix's dtype "
                f"{sparse_matrix.dtype}.")
        self.matrix = sparse_matrix

    def compute(self, sparse_matrix):
        self._check(sparse_matrix)
        raise QuadrantsRuntimeError(_NOT_IMPLEMENTED)

    def analyze_pattern(self, sparse_matrix):
        self._check(sparse_matrix)
        raise QuadrantsRuntimeError(_NOT_IMPLEMENTED)

    def factorize(self, sparse_matrix):
        self._check(sparse_matrix)
        raise QuadrantsRuntimeError(_NOT_IMPLEMENTED)

    def solve(self, b):
        if self.matrix is None:
            raise QuadrantsRuntimeError("Please call compute() before calling solve().")
        raise QuadrantsRuntimeError(_NOT_IMPLEMENTED)

    def info(self):
        return False


__all__ = ["SparseSolver"]
