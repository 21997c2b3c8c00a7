"""qd.linalg.SparseSolver (python/quadrants/linalg/sparse_solver.py:14-123) on arch=qd.cuda.

The reference builds `_qd_core.make_cusparse_solver(dtype, solver_type, ordering)` on CUDA (cuSOLVER low-level Cholesky
on the device; LU entirely on the host; f32 only -- quadrants/program/sparse_solver.cpp:149-476).  Here the same calls
reach quadrants_b200/csrc/solver.cu: minimum-degree ordering and symbolic analysis on the host (as the reference's
reorder() does), numeric LLT / LDLT / LU and the triangular solves on the device, f32 and f64.  Method names, argument
checks and error strings follow the reference file line by line; there is no iterative or host fallback behind this name.
"""
from __future__ import annotations

import numpy as np

from .. import _core
from ..lang import Arch, Field, Ndarray, QuadrantsRuntimeError, ScalarNdarray, f32, get_runtime
from .sparse_matrix import SparseMatrix


class SparseSolver:
    """Sparse linear system solver.

    Args:
        solver_type (str): The factorization type ("LLT", "LDLT" or "LU").
        ordering (str): The method for matrices re-ordering ("AMD" or "COLAMD").
    """

    def __init__(self, dtype=f32, solver_type="LLT", ordering="AMD"):
        self.matrix = None
        self.dtype = dtype
        solver_type_list = ["LLT", "LDLT", "LU"]
        solver_ordering = ["AMD", "COLAMD"]
        if solver_type in solver_type_list and ordering in solver_ordering:
            prog = get_runtime().prog
            assert prog is not None and prog.config().arch == Arch.cuda, "SparseSolver only supports CPU and CUDA for now."
            self.solver = _core.make_cusparse_solver(dtype, solver_type, ordering)
        else:
            raise QuadrantsRuntimeError(
                f"The solver type {solver_type} with {ordering} is not supported for now. Only {solver_type_list} with "
                f"{solver_ordering} are supported.")

    @staticmethod
    def _type_assert(sparse_matrix):
        raise QuadrantsRuntimeError(
            f"The parameter type: {type(sparse_matrix)} is not supported in linear solvers for now.")

    def compute(self, sparse_matrix):
        """Equivalent to calling both `analyze_pattern` and then `factorize`."""
        if isinstance(sparse_matrix, SparseMatrix):
            self.matrix = sparse_matrix
            self.analyze_pattern(self.matrix)
            self.factorize(self.matrix)
        else:
            self._type_assert(sparse_matrix)

    def analyze_pattern(self, sparse_matrix):
        """Reorders the matrix so that the factorization creates less fill-in, and analyses the factor's pattern."""
        if isinstance(sparse_matrix, SparseMatrix):
            self.matrix = sparse_matrix
            if self.matrix.dtype != self.dtype:
                raise QuadrantsRuntimeError(
                    f"The SparseSolver's dtype {self.dtype} is not consistent with the SparseMatrix's dtype "
                    f"{self.matrix.dtype}.")
            self.solver.analyze_pattern(sparse_matrix.matrix)
        else:
            self._type_assert(sparse_matrix)

    def factorize(self, sparse_matrix):
        """Does the numeric factorization (on the device)."""
        if isinstance(sparse_matrix, SparseMatrix):
            self.matrix = sparse_matrix
            self.solver.factorize(sparse_matrix.matrix)
        else:
            self._type_assert(sparse_matrix)

    def solve(self, b):  # pylint: disable=R1710
        """Computes the solution of the linear system.

        Args:
            b (numpy.array, Field or Ndarray): the right-hand side.
        Returns:
            numpy.array for a numpy / Field right-hand side, Ndarray (on the device) for an Ndarray one.
        """
        if self.matrix is None:
            raise QuadrantsRuntimeError("Please call compute() before calling solve().")
        if isinstance(b, Field):
            return self.solver.solve(b.to_numpy())
        if isinstance(b, np.ndarray):
            return self.solver.solve(b)
        if isinstance(b, Ndarray):
            x = ScalarNdarray(b.dtype, [self.matrix.m])
            self.solver.solve_rf(get_runtime().prog, self.matrix.matrix, b.arr, x.arr)
            return x
        raise QuadrantsRuntimeError(f"The parameter type: {type(b)} is not supported in linear solvers for now.")

    def info(self):
        """True if the factorization succeeded."""
        return self.solver.info()


__all__ = ["SparseSolver"]
