"""CPU-only checks of the modern qd.algorithms layer: the numpy restatements in oracle/ against the
known answers printed in the reference's user guide (docs/source/user_guide/algorithms.md, "Semantics" examples G10-G14)
and the reference tests' host formulas; the scratch-slot helpers against the reference's published figures and against
what libqdsp really needs; argument validation that happens before any kernel is launched."""
import numpy as np
import pytest

import oracle as O


# ---- G10-G14: the user guide's toy examples (algorithms.md:176-352) -------------------------------------------------------
def test_g10_reduce_doc_example():
    a = np.array([3, 1, 4, 1, 5, 9, 2, 6], np.int32)
    assert O.algo_reduce(a, "add") == 31 and O.algo_reduce(a, "min") == 1 and O.algo_reduce(a, "max") == 9


def test_g11_exclusive_scan_doc_example():
    a = np.array([3, 1, 4, 1, 5, 9, 2, 6], np.int32)
    assert O.algo_exclusive_scan(a, "add").tolist() == [0, 3, 4, 8, 9, 14, 23, 25]


def test_g12_select_doc_example():
    sel, k = O.algo_select(np.arange(10, 18, dtype=np.int32), np.array([1, 0, 1, 1, 0, 0, 1, 0], np.int32))
    assert k == 4 and sel.tolist() == [10, 12, 13, 16]


def test_g13_sort_doc_example():
    keys = np.array([3, 1, 4, 1, 5, 9, 2, 6], np.int32)
    order = O.algo_sort_order(keys)
    assert keys[order].tolist() == [1, 1, 2, 3, 4, 5, 6, 9]
    assert order.tolist() == [1, 3, 6, 0, 2, 4, 7, 5]  # stable for the tied 1s


def test_g14_reduce_by_key_doc_example():
    k, v, r = O.algo_reduce_by_key_add(np.array([1, 1, 1, 2, 2, 3, 3, 3], np.int32),
                                       np.array([5, 2, 1, 4, 4, 6, 1, 1], np.int32))
    assert r == 3 and k.tolist() == [1, 2, 3] and v.tolist() == [8, 8, 8]


def test_oracle_identities_and_wrap():
    # identities of the reference (_reduce.py:137-203)
    assert O.algo_reduce(np.zeros(0, np.int32), "min") == np.iinfo(np.int32).max
    assert O.algo_reduce(np.zeros(0, np.uint64), "max") == 0
    assert O.algo_reduce(np.zeros(0, np.float32), "max") == -np.inf
    assert O.algo_exclusive_scan(np.array([5.0, 1.0], np.float64), "min").tolist() == [np.inf, 5.0]
    # integer add wraps at the dtype's width
    a = np.array([2**31 - 1, 1, 5], np.int32)
    assert O.algo_reduce(a, "add") == np.int32(-2**31 + 5)
    assert O.algo_exclusive_scan(a, "add").tolist() == [0, 2**31 - 1, -2**31]
    # stable total order of floats: -0 before +0, by bit pattern
    f = np.array([0.0, -0.0, 1.0, -np.inf, -0.0], np.float32)
    assert O.algo_sort_order(f).tolist() == [3, 1, 4, 0, 2]
    # NaN keys open their own runs, +0 == -0 does not
    k, v, r = O.algo_reduce_by_key_add(np.array([np.nan, np.nan, 0.0, -0.0], np.float32), np.ones(4, np.int32))
    assert r == 3 and v.tolist() == [1, 1, 2]


# ---- scratch sizing ------------------------------------------------------------------------------------------------------
def test_scratch_slots_reference_figures():
    import quadrants_b200.algorithms as A

    # algorithms.md: "4112 slots at N = 1M" for the scan; reduce ~ ceil(N/256) + the staircase above it
    assert A.exclusive_scan_scratch_slots(1 << 20) == 4112
    assert A.reduce_scratch_slots(1 << 20) == 4112
    # select ~ N write indices + partials, reduce_by_key ~ 1.004 N, sort ~ ceil(N/256)*256 + staircase
    n = 1 << 20
    assert A.select_scratch_slots(n) == n + 4096 + 16
    assert A.reduce_by_key_scratch_slots(n) == n + 4096 + 16
    assert A.sort_scratch_slots(n, 3) == n + 4096 + 16


@pytest.mark.parametrize("n", [0, 1, 2, 256])
def test_scratch_slots_are_at_least_one(n):
    """tests/python/test_algorithms.py:697-718"""
    import quadrants_b200.algorithms as A

    for f in (A.reduce_scratch_slots, A.exclusive_scan_scratch_slots, A.select_scratch_slots,
              A.reduce_by_key_scratch_slots, A.sort_scratch_slots):
        assert f(n) >= 1


@pytest.mark.parametrize("n", [0, 1, 256, 100_000])
def test_scratch_slots_reject_depth_above_max(n):
    """tests/python/test_algorithms.py:721-745"""
    import quadrants_b200.algorithms as A

    for f in (A.reduce_scratch_slots, A.exclusive_scan_scratch_slots, A.sort_scratch_slots):
        with pytest.raises(ValueError):
            f(n, 5)
        assert f(n, 4) >= 1


def test_scratch_slots_cover_the_library_need():
    """A buffer of `slots` u32 (4-byte dtypes) or u64 (8-byte dtypes) elements is always enough for the kernels."""
    import quadrants_b200.algorithms as A
    from quadrants_b200 import _lib

    L = _lib.load()
    for n in [0, 1, 5, 255, 256, 257, 2047, 2048, 2049, 4095, 4096, 4097, 100_001, 1 << 20, (1 << 24) + 3, 1 << 28,
              (1 << 30) - 1]:
        for eb in (4, 8):
            assert A.reduce_scratch_slots(n) * eb >= L.qdsp_algo_scratch_bytes(_lib.ALGO_REDUCE, n, eb)
            assert A.exclusive_scan_scratch_slots(n) * eb >= L.qdsp_algo_scratch_bytes(_lib.ALGO_SCAN, n, eb)
            assert A.sort_scratch_slots(n) * 4 >= L.qdsp_algo_scratch_bytes(_lib.ALGO_SORT, n, eb)
            for d in range(1, 5):
                if 256 ** d >= n:
                    assert A.reduce_scratch_slots(n, d) * eb >= L.qdsp_algo_scratch_bytes(_lib.ALGO_REDUCE, n, eb)
                    assert A.sort_scratch_slots(n, d) * 4 >= L.qdsp_algo_scratch_bytes(_lib.ALGO_SORT, n, eb)
        assert A.select_scratch_slots(n) * 4 >= L.qdsp_algo_scratch_bytes(_lib.ALGO_SELECT, n, 4)
        assert A.reduce_by_key_scratch_slots(n) * 4 >= L.qdsp_algo_scratch_bytes(_lib.ALGO_REDUCE_BY_KEY, n, 4)


def test_compute_entries_fail_loudly_without_device():
    """No CPU fallback: with no CUDA device the new entry points return an error status, never a result."""
    import ctypes as C

    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is visible")
    from quadrants_b200 import _lib

    L = _lib.load()
    buf = (C.c_int32 * 64)()
    out = (C.c_int32 * 64)()
    scratch = (C.c_uint64 * 4096)()
    st = L.qdsp_reduce(C.addressof(buf), C.addressof(out), 64, None, _lib.ELEM_I32, _lib.OP_ADD, C.addressof(scratch),
                       C.sizeof(scratch), None)
    assert st != 0 and L.qdsp_last_error()
    st = L.qdsp_exclusive_scan(C.addressof(buf), C.addressof(out), 64, None, _lib.ELEM_I32, _lib.OP_ADD,
                               C.addressof(scratch), C.sizeof(scratch), None)
    assert st != 0
    st = L.qdsp_sort(C.addressof(buf), C.addressof(out), None, None, 64, None, _lib.ELEM_I32, 0, 32,
                     C.addressof(scratch), C.sizeof(scratch), None)
    assert st != 0


def test_abi_argument_validation_needs_no_device():
    import ctypes as C

    from quadrants_b200 import _lib

    L = _lib.load()
    buf = (C.c_int32 * 64)()
    out = (C.c_int32 * 64)()
    scratch = (C.c_uint64 * 4096)()
    a, o, s = C.addressof(buf), C.addressof(out), C.addressof(scratch)
    # in-place scan, odd pass count, unsupported key dtype of reduce_by_key, tiny scratch: refused before any launch
    assert L.qdsp_exclusive_scan(a, a, 64, None, _lib.ELEM_I32, _lib.OP_ADD, s, C.sizeof(scratch), None) == 1
    assert b"distinct" in L.qdsp_last_error()
    assert L.qdsp_sort(a, o, None, None, 64, None, _lib.ELEM_I32, 0, 24, s, C.sizeof(scratch), None) == 1
    assert L.qdsp_sort(a, a, None, None, 64, None, _lib.ELEM_I32, 0, 32, s, C.sizeof(scratch), None) == 1
    assert L.qdsp_reduce_by_key_add(a, a, o, o, o, 64, None, _lib.ELEM_I64, _lib.ELEM_I32, s, C.sizeof(scratch),
                                    None) == 4
    assert L.qdsp_reduce(a, o, 64, None, _lib.ELEM_I32, _lib.OP_ADD, s, 8, None) == 1
    assert b"scratch" in L.qdsp_last_error()
    assert L.qdsp_reduce(a, o, 64, None, 9, _lib.OP_ADD, s, C.sizeof(scratch), None) == 1
    assert L.qdsp_select(a, a, o, o, 64, None, 6, s, C.sizeof(scratch), None) == 1


def test_legacy_entries_emit_the_reference_deprecation_warnings():
    """tests/python/test_algorithms.py:748-759; the warning fires before anything touches the device."""
    import quadrants_b200 as qd

    with pytest.warns(DeprecationWarning, match="exclusive_scan_add"):
        try:
            qd.algorithms.PrefixSumExecutor(64)
        except Exception:  # no CUDA device here: qd.init was never called
            pass
    with pytest.warns(DeprecationWarning, match="sort"):
        with pytest.raises(RuntimeError):
            qd.algorithms.parallel_sort("not a field")
