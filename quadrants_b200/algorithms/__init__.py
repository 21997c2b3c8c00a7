"""qd.algorithms: the deprecated pair parallel_sort / PrefixSumExecutor (python/quadrants/algorithms/_algorithms.py)
and the modern device-wide ops that replace them (python/quadrants/algorithms/__init__.py:3-43)."""
from ._algorithms import PrefixSumExecutor, parallel_sort
from ._device import (exclusive_scan_add, exclusive_scan_max, exclusive_scan_min, exclusive_scan_scratch_slots,
                      reduce_add, reduce_by_key_add, reduce_by_key_scratch_slots, reduce_max, reduce_min,
                      reduce_scratch_slots, select, select_scratch_slots, sort, sort_scratch_slots)

__all__ = [
    "PrefixSumExecutor",
    "exclusive_scan_add",
    "exclusive_scan_max",
    "exclusive_scan_min",
    "exclusive_scan_scratch_slots",
    "parallel_sort",
    "reduce_add",
    "reduce_by_key_add",
    "reduce_by_key_scratch_slots",
    "reduce_max",
    "reduce_min",
    "reduce_scratch_slots",
    "select",
    "select_scratch_slots",
    "sort",
    "sort_scratch_slots",
]
