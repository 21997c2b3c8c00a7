"""qd.algorithms: parallel_sort / PrefixSumExecutor (python/quadrants/algorithms/_algorithms.py)."""
from ._algorithms import PrefixSumExecutor, parallel_sort

__all__ = ["parallel_sort", "PrefixSumExecutor"]
