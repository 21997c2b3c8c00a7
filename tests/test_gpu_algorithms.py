"""GPU parity of the modern qd.algorithms ops and of the bulk triplet emission (N2) against the numpy
restatements in oracle/ -- the same host references the reference's own tests use
(tests/python/test_algorithms.py:267-695), through the public API (-> ctypes -> C ABI -> kernels).  Needs a B200.

Bars: integer results, min / max, select, sort (keys, values, stability), run keys and run counts: bit-exact.
Float add (reduce, scan, reduce_by_key): the reference's own tolerances (test_algorithms.py:134-152), and -- stronger
than the reference, whose float results are not reproducible -- bit-identical from run to run.
"""
import math

import numpy as np
import pytest

import oracle as O
from oracle import make_triplet_buffer

pytestmark = pytest.mark.gpu

SIZES = [1, 255, 256, 257, 1024, 4095, 4096, 4097, 65537, (1 << 20) + 7]
F32_RTOL, F32_ATOL = 1e-4, 1e-4  # _F32_REDUCE_RTOL / _ATOL
F64_RTOL, F64_ATOL = 1e-12, 1e-9  # _F64_RTOL / _ATOL


@pytest.fixture(scope="module")
def qd():
    import quadrants_b200 as qd

    qd.init(arch=qd.cuda)
    return qd


def all_dtypes(qd):
    return [qd.i32, qd.u32, qd.f32, qd.i64, qd.u64, qd.f64]


def depth(n):
    d, cap = 1, 256
    while cap < n:
        cap *= 256
        d += 1
    return d


def nd(qd, dt, a):
    x = qd.ndarray(dt, a.shape[0])
    x.from_numpy(a)
    return x


def count_nd(qd, n):
    c = qd.ndarray(qd.i32, 1)
    c.from_numpy(np.array([n], np.int32))
    return c


def scratch_for(qd, dt, slots):
    return qd.ndarray(qd.u32 if dt.np_dtype.itemsize == 4 else qd.u64, slots)


def rand_vals(rng, dt, n, op):
    npd = dt.np_dtype
    if npd.kind == "f":
        return rng.uniform(-1, 1, n).astype(npd) if op == "add" else rng.uniform(-10, 10, n).astype(npd)
    if op == "add":  # full range: integer add must wrap exactly like numpy's fixed-width sum
        info = np.iinfo(npd)
        return rng.integers(info.min, info.max, n, dtype=npd, endpoint=True)
    bound = 10000
    return rng.integers(0 if npd.kind == "u" else -bound, bound, n, dtype=npd)


def tol(dt, n=1):
    return (F32_RTOL, F32_ATOL) if dt.np_dtype.itemsize == 4 else (F64_RTOL, F64_ATOL)


# ---------------------------------------------------------------------------------------------------------------------
# reduce_{add,min,max}
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("op", ["add", "min", "max"])
@pytest.mark.parametrize("n", SIZES)
def test_reduce(qd, op, n):
    rng = np.random.default_rng(7)
    fn = getattr(qd.algorithms, f"reduce_{op}")
    for dt in all_dtypes(qd):
        host = rand_vals(rng, dt, n, op)
        arr, out = nd(qd, dt, host), qd.ndarray(dt, 1)
        D = depth(n)
        scratch = scratch_for(qd, dt, qd.algorithms.reduce_scratch_slots(n, D))
        fn(arr, out, scratch, count_nd(qd, n), dt, D)  # device-resident count, like count[0] in the reference
        got = out.to_numpy()[0]
        fn(arr, out, scratch, n, dt, D)  # host count
        got2 = out.to_numpy()[0]
        exp = O.algo_reduce(host, op)
        if op == "add" and dt.np_dtype.kind == "f":
            rt, at = tol(dt)
            assert math.isclose(float(got), float(exp), rel_tol=rt, abs_tol=at), (dt, n, got, exp)
            assert got.tobytes() == got2.tobytes(), "float reduce must be reproducible"
        else:
            assert got == exp and got2 == exp, (dt, op, n, got, exp)


def test_reduce_empty_and_partial_count(qd):
    rng = np.random.default_rng(8)
    for dt in all_dtypes(qd):
        host = rand_vals(rng, dt, 5000, "min")
        arr, out = nd(qd, dt, host), qd.ndarray(dt, 1)
        scratch = scratch_for(qd, dt, qd.algorithms.reduce_scratch_slots(5000, 2))
        for op in ("add", "min", "max"):
            fn = getattr(qd.algorithms, f"reduce_{op}")
            fn(arr, out, scratch, count_nd(qd, 0), dt, 2)
            assert out.to_numpy()[0] == O.algo_reduce(host[:0], op), (dt, op)
            fn(arr, out, scratch, count_nd(qd, 4097), dt, 2)  # live count below the capacity
            exp = O.algo_reduce(host[:4097], op)
            got = out.to_numpy()[0]
            if op == "add" and dt.np_dtype.kind == "f":
                assert math.isclose(float(got), float(exp), rel_tol=1e-4, abs_tol=1e-3)
            else:
                assert got == exp, (dt, op)


# ---------------------------------------------------------------------------------------------------------------------
# exclusive_scan_{add,min,max}
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("op", ["add", "min", "max"])
@pytest.mark.parametrize("n", SIZES)
def test_exclusive_scan(qd, op, n):
    rng = np.random.default_rng(11)
    fn = getattr(qd.algorithms, f"exclusive_scan_{op}")
    for dt in all_dtypes(qd):
        host = rand_vals(rng, dt, n, op)
        arr, out = nd(qd, dt, host), qd.ndarray(dt, n)
        D = depth(n)
        scratch = scratch_for(qd, dt, qd.algorithms.exclusive_scan_scratch_slots(n, D))
        fn(arr, out, scratch, count_nd(qd, n), dt, D)
        got = out.to_numpy()
        exp = O.algo_exclusive_scan(host, op)
        if op == "add" and dt.np_dtype.kind == "f":
            rtol = 2e-5 * math.sqrt(n) if dt is qd.f32 else 1e-12  # _f32_scan_tol
            np.testing.assert_allclose(got, exp, rtol=rtol, atol=1e-3 if dt is qd.f32 else 1e-9)
            out2 = qd.ndarray(dt, n)
            fn(arr, out2, scratch, n, dt, D)
            assert out2.to_numpy().tobytes() == got.tobytes(), "float scan must be reproducible"
        else:
            assert np.array_equal(got, exp), (dt, op, n)


def test_exclusive_scan_partial_count_leaves_tail(qd):
    rng = np.random.default_rng(12)
    n, live = 20000, 12345
    host = rng.integers(-2**31, 2**31 - 1, n, dtype=np.int32)
    arr, out = nd(qd, qd.i32, host), qd.ndarray(qd.i32, n)
    out.fill(-7)
    scratch = qd.ndarray(qd.u32, qd.algorithms.exclusive_scan_scratch_slots(n, 2))
    qd.algorithms.exclusive_scan_add(arr, out, scratch, count_nd(qd, live), qd.i32, 2)
    got = out.to_numpy()
    assert np.array_equal(got[:live], O.algo_exclusive_scan(host[:live], "add"))
    assert np.all(got[live:] == -7)


def test_exclusive_scan_misaligned_views(qd):
    """Buffers that are only element-aligned take the scalar path: same result."""
    import torch

    rng = np.random.default_rng(13)
    n = 9001
    host = rng.integers(-1000, 1000, n, dtype=np.int32)
    big_in, big_out = qd.ndarray(qd.i32, n + 1), qd.ndarray(qd.i32, n + 1)
    big_in.tensor[1:].copy_(torch.from_numpy(host))
    arr = qd.Ndarray(qd.i32, n, _tensor=big_in.tensor[1:])
    out = qd.Ndarray(qd.i32, n, _tensor=big_out.tensor[1:])
    scratch = qd.ndarray(qd.u32, qd.algorithms.exclusive_scan_scratch_slots(n, 2))
    qd.algorithms.exclusive_scan_add(arr, out, scratch, n, qd.i32, 2)
    assert np.array_equal(out.to_numpy(), O.algo_exclusive_scan(host, "add"))


# ---------------------------------------------------------------------------------------------------------------------
# select
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", SIZES)
def test_select(qd, n):
    rng = np.random.default_rng(17)
    for dt in (qd.i32, qd.f32, qd.u64, qd.f64):
        host = rand_vals(rng, dt, n, "min")
        for density in (0.5, 0.0, 1.0, 0.01):
            flags = (rng.random(n) < density).astype(np.int32)
            arr, fl, out = nd(qd, dt, host), nd(qd, qd.i32, flags), qd.ndarray(dt, n)
            sentinel = dt.np_dtype.type(77)
            out.fill(sentinel)
            num = qd.ndarray(qd.i32, 1)
            scratch = qd.ndarray(qd.u32, qd.algorithms.select_scratch_slots(n))
            qd.algorithms.select(arr, fl, out, num, scratch, count_nd(qd, n), depth(n))
            exp, k = O.algo_select(host, flags)
            assert int(num.to_numpy()[0]) == k, (dt, n, density)
            got = out.to_numpy()
            assert got[:k].tobytes() == exp.tobytes()
            assert np.all(got[k:] == sentinel), "the tail past num_out must be left untouched"


# ---------------------------------------------------------------------------------------------------------------------
# reduce_by_key_add
# ---------------------------------------------------------------------------------------------------------------------
def run_rbk(qd, kdt, vdt, keys, vals, live=None):
    n = keys.shape[0]
    ki, vi = nd(qd, kdt, keys), nd(qd, vdt, vals)
    ko, vo, nr = qd.ndarray(kdt, max(n, 1)), qd.ndarray(vdt, max(n, 1)), qd.ndarray(qd.i32, 1)
    ko.fill(kdt.np_dtype.type(55))
    vo.fill(vdt.np_dtype.type(55))
    scratch = qd.ndarray(qd.u32, qd.algorithms.reduce_by_key_scratch_slots(n))
    cnt = count_nd(qd, n if live is None else live)
    qd.algorithms.reduce_by_key_add(ki, vi, ko, vo, nr, scratch, cnt, vdt, depth(max(n, 1)))
    return ko.to_numpy(), vo.to_numpy(), int(nr.to_numpy()[0])


def check_rbk(qd, kdt, vdt, keys, vals, live=None):
    m = keys.shape[0] if live is None else live
    gk, gv, r = run_rbk(qd, kdt, vdt, keys, vals, live)
    ek, ev, er = O.algo_reduce_by_key_add(keys[:m], vals[:m])
    assert r == er
    assert gk[:r].tobytes() == ek.tobytes()
    if vdt.np_dtype.kind == "f":
        # run sums of up to ~1e5 values in [-1, 1]: _F32_LARGE_N_RTOL / _ATOL of the reference's RBK test
        np.testing.assert_allclose(gv[:r], ev, rtol=1e-3, atol=1e-3)
    else:
        assert np.array_equal(gv[:r], ev)
    assert np.all(gk[r:] == kdt.np_dtype.type(55)) and np.all(gv[r:] == vdt.np_dtype.type(55)), "tails untouched"


@pytest.mark.parametrize("n", SIZES)
def test_reduce_by_key_add(qd, n):
    rng = np.random.default_rng(19)
    for kdt in (qd.i32, qd.f32, qd.u32):
        for vdt in (qd.i32, qd.f32, qd.u32):
            # runs of random length: mean 1 (mostly unique), 8, and 3000 (runs that cross tiles)
            for mean_run in (1, 8, 3000):
                nruns = max(n // mean_run, 1)
                cuts = np.sort(rng.integers(0, n, nruns))
                run_id = np.searchsorted(cuts, np.arange(n), side="right")
                labels = rng.integers(-50, 50, nruns + 1)
                keys = labels[run_id].astype(kdt.np_dtype) if kdt is not qd.u32 else (labels[run_id] + 50).astype(np.uint32)
                vals = rand_vals(rng, vdt, n, "add")
                check_rbk(qd, kdt, vdt, keys, vals)


def test_reduce_by_key_single_run_nan_and_zero_keys(qd):
    n = 65537
    rng = np.random.default_rng(23)
    vals = rng.integers(-5, 5, n, dtype=np.int32)
    check_rbk(qd, qd.i32, qd.i32, np.full(n, 3, np.int32), vals)  # one run spanning every tile
    fk = np.zeros(n, np.float32)
    fk[1::2] = -0.0  # +0 == -0: still one run
    check_rbk(qd, qd.f32, qd.i32, fk, vals)
    fk = rng.integers(0, 3, n).astype(np.float32)
    fk[rng.integers(0, n, 500)] = np.nan  # every NaN is a run of its own
    fk[100:104] = np.nan
    gk, gv, r = run_rbk(qd, qd.f32, qd.i32, fk, vals)
    ek, ev, er = O.algo_reduce_by_key_add(fk, vals)
    assert r == er and gk[:r].tobytes() == ek.tobytes() and np.array_equal(gv[:r], ev)
    check_rbk(qd, qd.i32, qd.i32, rng.integers(0, 4, n, dtype=np.int32), vals, live=40000)  # live count below capacity
    check_rbk(qd, qd.i32, qd.i32, rng.integers(0, 4, n, dtype=np.int32), vals, live=0)


# ---------------------------------------------------------------------------------------------------------------------
# sort
# ---------------------------------------------------------------------------------------------------------------------
def special_floats(npd):
    fi = np.finfo(npd)
    return np.array([0.0, -0.0, np.inf, -np.inf, fi.tiny, -fi.tiny, fi.smallest_subnormal, -fi.smallest_subnormal,
                     fi.max, -fi.max, 1.0, -1.0], dtype=npd)


@pytest.mark.parametrize("n", [1, 257, 1024, 65536, (1 << 20) + 3])
def test_sort_key_value(qd, n):
    rng = np.random.default_rng(1357)
    for kdt in all_dtypes(qd):
        npd = kdt.np_dtype
        if npd.kind == "f":
            keys = rng.standard_normal(n).astype(npd)
            sp = special_floats(npd)
            if n > 4 * sp.shape[0]:
                keys[rng.choice(n, 4 * sp.shape[0], replace=False)] = np.tile(sp, 4)
        else:
            info = np.iinfo(npd)
            keys = rng.integers(info.min, info.max, n, dtype=npd, endpoint=True)
            keys[rng.integers(0, n, n // 3)] = keys[0]  # ties: stability is observable through the values
        for vdt in (qd.i32, qd.u64):
            vals = np.arange(n).astype(vdt.np_dtype)
            k, tk = nd(qd, kdt, keys), qd.ndarray(kdt, n)
            v, tv = nd(qd, vdt, vals), qd.ndarray(vdt, n)
            D = depth(n)
            scratch = qd.ndarray(qd.u32, qd.algorithms.sort_scratch_slots(n, D))
            cnt = qd.ndarray(qd.i32, ())
            cnt.fill(n)
            qd.algorithms.sort(k, tk, v, tv, scratch, cnt, kdt, True, npd.itemsize * 8, D)
            order = O.algo_sort_order(keys)
            assert k.to_numpy().tobytes() == keys[order].tobytes(), (kdt, vdt, n)
            assert np.array_equal(v.to_numpy(), vals[order]), (kdt, vdt, n)


def test_sort_keys_only_partial_bits_and_partial_count(qd):
    rng = np.random.default_rng(29)
    n, live = 100001, 70001
    keys = rng.integers(0, 1 << 16, n, dtype=np.uint32)
    k, tk = nd(qd, qd.u32, keys), qd.ndarray(qd.u32, n)
    scratch = qd.ndarray(qd.u32, qd.algorithms.sort_scratch_slots(n, 3))
    qd.algorithms.sort(k, tk, k, tk, scratch, count_nd(qd, live), qd.u32, False, 16, 3)  # placeholders for the values
    got = k.to_numpy()
    assert np.array_equal(got[:live], np.sort(keys[:live], kind="stable"))
    assert np.array_equal(got[live:], keys[live:]), "elements past the live count must not move"
    # the low 16 bits only: the high half rides along, ties on the low half keep their input order
    keys = rng.integers(0, 2**32 - 1, n, dtype=np.uint32)
    k.from_numpy(keys)
    qd.algorithms.sort(k, tk, k, tk, scratch, n, qd.u32, False, 16, 3)
    assert np.array_equal(k.to_numpy(), keys[O.algo_sort_order(keys, end_bit=16)])
    # 64-bit keys only
    keys = rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64)
    k64, tk64 = nd(qd, qd.i64, keys), qd.ndarray(qd.i64, n)
    qd.algorithms.sort(k64, tk64, k64, tk64, scratch, n, qd.i64, False, 64, 3)
    assert np.array_equal(k64.to_numpy(), np.sort(keys, kind="stable"))


def test_algorithm_errors(qd):
    A = qd.algorithms
    a, o = qd.ndarray(qd.i32, 100), qd.ndarray(qd.i32, 100)
    s = qd.ndarray(qd.u32, A.exclusive_scan_scratch_slots(100, 1))
    with pytest.raises(ValueError):
        A.exclusive_scan_add(a, a, s, 100, qd.i32, 1)  # in place
    with pytest.raises(ValueError):
        A.reduce_add(a, o, s, 100, qd.i32, 5)  # depth above 4
    with pytest.raises(ValueError):
        A.reduce_add(a, o, s, 300, qd.i32, 2)  # n beyond the array
    with pytest.raises(ValueError):
        A.sort(a, o, a, o, qd.ndarray(qd.u32, A.sort_scratch_slots(100)), 100, qd.i32, False, 24, 1)  # odd passes
    with pytest.raises(RuntimeError):
        A.reduce_add(a, o, qd.ndarray(qd.u32, 1), 100, qd.i32, 1)  # scratch too small
    with pytest.raises(NotImplementedError):
        A.reduce_by_key_add(qd.ndarray(qd.i64, 4), a, qd.ndarray(qd.i64, 4), o, qd.ndarray(qd.i32, 1), s, 4, qd.i32, 1)


# ---------------------------------------------------------------------------------------------------------------------
# N2: bulk triplet emission
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_builder_append_and_elements_match_oracle(qd, oracle, dtype):
    rng = np.random.default_rng(31)
    n_nodes, n_elems, k = 500, 3000, 4
    nodes = np.stack([rng.choice(n_nodes, k, replace=False) for _ in range(n_elems)]).astype(np.int32)
    mats = rng.uniform(-1, 1, (n_elems, k, k)).astype(dtype)
    extra_r = rng.integers(0, n_nodes, 1000).astype(np.int32)
    extra_c = rng.integers(0, n_nodes, 1000).astype(np.int32)
    extra_v = rng.uniform(-1, 1, 1000).astype(dtype)
    dt = qd.f32 if dtype == np.float32 else qd.f64
    K = qd.linalg.SparseMatrixBuilder(n_nodes, n_nodes, max_num_triplets=n_elems * k * k + 2000, dtype=dt,
                                      storage_format="row_major")
    K.insert_triplets(extra_r, extra_c, extra_v)
    K.insert_elements(nodes, mats)
    K.ptr.insert_triplets(extra_r, extra_c, extra_v, negate=True)  # A[i, j] -= v
    assert K.ptr.num_triplets() == n_elems * k * k + 2000
    A = K.build()
    # the same triplet stream on the host, element-major then (a, b) row-major, handed to the oracle's setFromTriplets
    er = np.repeat(nodes, k, axis=1).reshape(-1)
    ec = np.tile(nodes, (1, k)).reshape(-1)
    r = np.concatenate([extra_r, er, extra_r])
    c = np.concatenate([extra_c, ec, extra_c])
    v = np.concatenate([extra_v, mats.reshape(-1), -extra_v])
    buf = make_triplet_buffer(r, c, v, dtype)
    oo, oi, ov = oracle.build_matrix(buf, n_nodes, n_nodes, True)
    ip, ix, vals = A.matrix.to_scipy_arrays()
    assert np.array_equal(ip, oo) and np.array_equal(ix, oi)
    assert vals.tobytes() == ov.tobytes(), "duplicate sums must follow the append order bit for bit"


def test_builder_append_overflow(qd):
    K = qd.linalg.SparseMatrixBuilder(8, 8, max_num_triplets=10)
    K.insert_triplets([0] * 6, [1] * 6, np.ones(6, np.float32))
    with pytest.raises(RuntimeError):
        K.insert_triplets([0] * 6, [1] * 6, np.ones(6, np.float32))
    assert K.ptr.num_triplets() == 6  # the refused append wrote nothing
    K.insert_triplets([2] * 4, [3] * 4, np.ones(4, np.float32))
    A = K.build()
    assert A[0, 1] == 6.0 and A[2, 3] == 4.0
