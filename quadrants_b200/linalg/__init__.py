"""qd.linalg: SparseMatrixBuilder / SparseMatrix / SparseCG / SparseSolver and the matrix-free Krylov solvers
(python/quadrants/linalg/__init__.py:1-10)."""
from .matrixfree_cg import LinearOperator, MatrixFreeBICGSTAB, MatrixFreeCG
from .sparse_cg import SparseCG
from .sparse_matrix import SparseMatrix, SparseMatrixBuilder
from .sparse_solver import SparseSolver

__all__ = ["SparseCG", "SparseMatrix", "SparseMatrixBuilder", "SparseSolver", "LinearOperator", "MatrixFreeCG",
           "MatrixFreeBICGSTAB"]
