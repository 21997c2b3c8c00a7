"""qd.algorithms.parallel_sort and PrefixSumExecutor with the reference's call shapes
(python/quadrants/algorithms/_algorithms.py:19-163), backed by the onesweep radix sort and the single-pass
decoupled look-back scan of libqdsp."""
from __future__ import annotations

import ctypes as C

from .. import _lib
from ..lang import Ndarray, f32, i32, u32, _require_prog

_KEY_KIND = {"i32": _lib.KEY_I32, "f32": _lib.KEY_F32, "u32": _lib.KEY_U32}


def _scratch(nbytes: int):
    import torch

    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device="cuda")


def parallel_sort(keys, values=None):
    """In-place ascending sort of `keys`; `values` (optional, 4- or 8-byte items) is permuted alike.

    The reference is a Batcher odd-even merge network (unstable, O(N log^2 N), one launch + sync() per stage); this is a
    stable LSD radix sort, so keys come out identical and ties keep their input order.  Unlike the reference it does
    not emit a DeprecationWarning-worthy amount of launches: 2 + 4 kernels for 32-bit keys."""
    import warnings

    warnings.warn(  # the reference's deprecation notice (_algorithms.py:33-40)
        "qd.algorithms.parallel_sort is deprecated. Use qd.algorithms.sort (the @qd.func) instead. "
        "See docs/source/user_guide/algorithms.md for migration.", DeprecationWarning, stacklevel=2)
    if not isinstance(keys, Ndarray):
        raise RuntimeError("parallel_sort expects 1-D qd fields / ndarrays")
    prog = _require_prog()
    n = keys.shape[0]
    kind = _KEY_KIND.get(keys.dtype.name)
    if kind is None:
        raise RuntimeError(f"parallel_sort: unsupported key dtype {keys.dtype}")
    vbytes, vptr = 0, None
    if values is not None:
        if values.shape[0] != n:
            raise RuntimeError("parallel_sort: keys and values differ in length")
        vbytes = values.dtype.np_dtype.itemsize
        vptr = values.data_ptr()
    lib = _lib.load()
    tmp = _scratch(lib.qdsp_sort_tmp_bytes(n, kind, vbytes))
    _lib.check(lib.qdsp_sort_pairs(keys.data_ptr(), vptr, n, kind, vbytes, tmp.data_ptr(), tmp.numel(),
                                   prog.stream()))


class PrefixSumExecutor:
    """Inclusive in-place parallel prefix sum (i32, wrapping).  One executor per length, like the reference, so the
    scratch buffer is allocated once."""

    def __init__(self, length):
        import warnings

        warnings.warn(  # the reference's deprecation notice (_algorithms.py:78-86)
            "qd.algorithms.PrefixSumExecutor is deprecated. Use qd.algorithms.exclusive_scan_add (the @qd.func) "
            "instead. See docs/source/user_guide/algorithms.md for migration.", DeprecationWarning, stacklevel=2)
        self.sorting_length = int(length)
        self._lib = _lib.load()
        _require_prog()
        self._tmp = _scratch(self._lib.qdsp_scan_tmp_bytes(self.sorting_length))

    def run(self, input_arr):
        if input_arr.dtype is not i32:
            raise RuntimeError("Only qd.i32 type is supported for prefix sum.")
        if input_arr.shape[0] < self.sorting_length:
            raise RuntimeError("PrefixSumExecutor: array shorter than the executor's length")
        _lib.check(self._lib.qdsp_inclusive_scan_i32(input_arr.data_ptr(), self.sorting_length, self._tmp.data_ptr(),
                                                     self._tmp.numel(), _require_prog().stream()))


__all__ = ["parallel_sort", "PrefixSumExecutor"]
