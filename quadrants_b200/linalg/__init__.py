"""qd.linalg: SparseMatrixBuilder / SparseMatrix / SparseCG (python/quadrants/linalg/__init__.py:1-10)."""
from .sparse_cg import SparseCG
from .sparse_matrix import SparseMatrix, SparseMatrixBuilder

__all__ = ["SparseCG", "SparseMatrix", "SparseMatrixBuilder"]
