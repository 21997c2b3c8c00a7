"""CPU-only: the host-side work partition of the irregular-matrix SpMV (qdsp_spmv_plan_host) covers every row and every
non-zero exactly once and respects the stage capacity the kernel assumes."""
import ctypes as C

import numpy as np
import pytest

CAP, STREAM_CAP, MAX_ROWS = 2048, 2044, 256


def plan(indptr):
    from quadrants_b200 import _lib

    lib = _lib.load()
    ip = np.ascontiguousarray(indptr, dtype=np.int32)
    rows = ip.shape[0] - 1
    cap = rows + int(ip[-1]) // 1024 + 16
    out = np.zeros((cap, 4), np.int32)
    nb, nl, ns = C.c_int64(), C.c_int(), C.c_int()
    _lib.check(lib.qdsp_spmv_plan_host(ip.ctypes.data, rows, out.ctypes.data, cap, C.byref(nb), C.byref(nl), C.byref(ns)))
    return out[: nb.value], nl.value, ns.value


def check(lens):
    indptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    blocks, n_long, n_seg = plan(indptr)
    rows = len(lens)
    next_row, next_nz, seg_seen = 0, 0, 0
    long_rows = 0
    for row0, nr, nz0, nz1 in blocks:
        assert nz0 == next_nz and nz1 >= nz0
        assert nz1 - (nz0 & ~3) <= CAP
        if nr > 0:
            assert row0 == next_row and nr <= MAX_ROWS
            assert nz0 == indptr[row0] and nz1 == indptr[row0 + nr]
            assert all(l <= STREAM_CAP for l in lens[row0:row0 + nr])
            next_row = row0 + nr
        else:
            assert -nr - 1 == seg_seen and lens[row0] > STREAM_CAP and row0 == next_row
            assert indptr[row0] <= nz0 < nz1 <= indptr[row0 + 1]
            seg_seen += 1
            if nz0 == indptr[row0]:
                long_rows += 1
            if nz1 == indptr[row0 + 1]:
                next_row = row0 + 1
        next_nz = nz1
    assert next_row == rows and next_nz == indptr[-1]
    assert n_long == long_rows == int(np.sum(np.asarray(lens) > STREAM_CAP)) and n_seg == seg_seen


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_plan_power_law(seed):
    rng = np.random.default_rng(seed)
    lens = np.minimum((rng.pareto(1.1, 5000) * 8).astype(np.int64), 20000)
    lens[rng.integers(0, 5000, 3)] = [2044, 2045, 9000]
    check(lens.tolist())


def test_plan_edges():
    check([])
    check([0])
    check([0, 0, 0, 5, 0])
    check([7] * 1000)                      # regular: blocks of 256 rows
    check([1] * 3000)                      # row-count cap binds
    check([2044, 2044, 3, 2045, 1, 100000, 0, 2])
    check([2043, 1, 1, 1, 1, 1, 1])
