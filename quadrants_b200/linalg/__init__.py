"""qd.linalg: SparseMatrixBuilder / SparseMatrix / SparseCG / SparseSolver (python/quadrants/linalg/__init__.py:1-10)."""
from .sparse_cg import SparseCG
from .sparse_matrix import SparseMatrix, SparseMatrixBuilder
from .sparse_solver import SparseSolver

__all__ = ["SparseCG", "SparseMatrix", "SparseMatrixBuilder", "SparseSolver"]
