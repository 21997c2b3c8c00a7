"""Pins the CPU oracle (oracle/qd_oracle.c) against the reference's own golden vectors.
CPU only."""
import numpy as np
import pytest

from oracle import make_triplet_buffer
from tests import workloads as W


def _dense(outer, inner, vals, rows, cols, row_major):
    a = np.zeros((rows, cols), vals.dtype)
    for o in range(outer.shape[0] - 1):
        for k in range(outer[o], outer[o + 1]):
            if row_major:
                a[o, inner[k]] = vals[k]
            else:
                a[inner[k], o] = vals[k]
    return a


@pytest.mark.parametrize("row_major", [False, True])
def test_g1_gpu_sparse_matrix_known_answer(oracle, row_major):
    # tests/python/test_sparse_matrix.py:414-451
    r = [0, 0, 0, 1, 2, 2, 2, 3, 3]
    c = [0, 2, 3, 1, 0, 2, 3, 1, 3]
    v = np.arange(1, 10, dtype=np.float32)
    buf = make_triplet_buffer(r, c, v, np.float32, capacity=50)
    outer, inner, vals = oracle.build_matrix(buf, 4, 4, row_major)
    y = oracle.spmv(outer, inner, vals, np.array([1, 2, 3, 4], np.float32), 4, 4, row_major)
    assert y.tolist() == [19.0, 8.0, 51.0, 52.0]
    if row_major:
        assert outer.tolist() == [0, 3, 4, 7, 9]
        assert inner.tolist() == c


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("row_major", [False, True])
@pytest.mark.parametrize("case", ["i", "i-j", "i+j"])
def test_g2_spmv_8x8(oracle, dtype, row_major, case):
    # tests/python/test_spmv.py:18-103
    n = 8
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    val = {"i": i, "i-j": i - j, "i+j": i + j}[case].astype(dtype)
    buf = make_triplet_buffer(i.ravel(), j.ravel(), val.ravel(), dtype, capacity=100)
    outer, inner, vals = oracle.build_matrix(buf, n, n, row_major)
    y = oracle.spmv(outer, inner, vals, np.ones(n, dtype), n, n, row_major)
    want = {"i": [8 * k for k in range(n)], "i-j": [-28, -20, -12, -4, 4, 12, 20, 28],
            "i+j": [28, 36, 44, 52, 60, 68, 76, 84]}[case]
    assert y.tolist() == [float(w) for w in want]


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("row_major", [False, True])
def test_g3_builder_elements(oracle, dtype, row_major):
    # tests/python/test_sparse_matrix.py:44-57 (A[i,j] += i+j, n=8) and :395-411 (n=2 -> [1,3])
    n = 8
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    buf = make_triplet_buffer(i.ravel(), j.ravel(), (i + j).ravel().astype(dtype), dtype, capacity=100)
    outer, inner, vals = oracle.build_matrix(buf, n, n, row_major)
    assert np.array_equal(_dense(outer, inner, vals, n, n, row_major), (i + j).astype(dtype))
    n = 2
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    buf = make_triplet_buffer(i.ravel(), j.ravel(), (i + j).ravel().astype(dtype), dtype, capacity=100)
    outer, inner, vals = oracle.build_matrix(buf, n, n, row_major)
    assert oracle.spmv(outer, inner, vals, np.ones(n, dtype), n, n, row_major).tolist() == [1.0, 3.0]


def test_g4_scipy_seeded_build(oracle):
    # tests/python/test_sparse_matrix.py:454-523: seed-2 random 5x4 matrices; the build must reproduce scipy's entries.
    from numpy.random import default_rng
    from scipy import stats
    from scipy.sparse import random

    rng = default_rng(2)
    rvs = stats.poisson(3, loc=1).rvs
    s1 = random(5, 4, density=0.5, random_state=rng, data_rvs=rvs).astype(np.float32).tocoo()
    buf = make_triplet_buffer(s1.row, s1.col, s1.data, np.float32)
    outer, inner, vals = oracle.build_matrix(buf, 5, 4, True)
    csr = s1.tocsr()
    csr.sort_indices()
    assert np.array_equal(outer, csr.indptr) and np.array_equal(inner, csr.indices) and np.array_equal(vals, csr.data)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_g5_cg_10x10(oracle, dtype):
    # tests/python/test_sparse_cg.py:9-39: SparseCG(max_iter=50, atol=1e-6) on A A^T, b = 1..10
    i, j, v, a, b = W.spd_dense_10(0, dtype)
    buf = make_triplet_buffer(i, j, v, dtype, capacity=500)
    outer, inner, vals = oracle.build_matrix(buf, 10, 10, False)
    x, ok, it, err = oracle.cg(outer, inner, vals, b, 50, 1e-6)
    want = np.linalg.solve(a.astype(np.float64), b.astype(np.float64))
    if dtype == np.float64:
        assert ok
    # the reference's own tolerance is rel=1.0 (very loose); hold the oracle to the true-residual instead
    res = np.linalg.norm(a.astype(np.float64) @ x - b) / np.linalg.norm(b)
    assert res < (5e-3 if dtype == np.float32 else 1e-5)
    assert np.allclose(x, want, rtol=1.0, atol=1e-2 * np.abs(want).max())


def test_g5_cucg_10x10(oracle):
    # tests/python/test_sparse_cg.py:42-70 (reference CUDA semantics: plain CG, absolute tolerance)
    i, j, v, a, b = W.spd_dense_10(0, np.float32)
    buf = make_triplet_buffer(i, j, v, np.float32, capacity=500)
    outer, inner, vals = oracle.build_matrix(buf, 10, 10, True)
    x, k, r1 = oracle.cucg(outer, inner, vals, b, 50, 1e-6)
    assert k <= 50
    res = np.linalg.norm(a.astype(np.float64) @ x - b) / np.linalg.norm(b)
    assert res < 5e-2


@pytest.mark.parametrize("dtype", [np.int32, np.float32])
@pytest.mark.parametrize("n", [1, 256, 100001])
def test_g6_sort(oracle, dtype, n):
    # tests/python/test_sort.py:9-62: keys = floor(rand*N), values = keys; ascending and keys == values
    rng = np.random.default_rng(n)
    keys = (rng.random(n) * n).astype(dtype)
    k, v = oracle.stable_sort(keys, keys.view(np.uint32))
    assert np.all(k[:-1] <= k[1:]) and np.array_equal(k.view(np.uint32), v)
    assert np.array_equal(k, np.sort(keys, kind="stable"))
    if n <= 256:
        # the exact comparator network of the reference gives the same keys (and the same values since values==keys)
        nk, nv = oracle.network_sort(keys, keys.view(np.int32))
        assert np.array_equal(nk, k) and np.array_equal(nv.view(np.uint32), v)


def test_network_vs_stable_small(oracle):
    """The reference network is unstable in general; its KEY output always equals the stable sort's, and per-key value
    multisets agree.  N need not be a power of two."""
    rng = np.random.default_rng(11)
    for n in (2, 3, 17, 100, 1000, 4097):
        keys = rng.integers(0, 8, n).astype(np.int32)
        vals = np.arange(n, dtype=np.int32)
        nk, nv = oracle.network_sort(keys, vals)
        sk, sv = oracle.stable_sort(keys, vals.view(np.uint32))
        assert np.array_equal(nk, sk)
        for key in range(8):
            assert sorted(nv[nk == key].tolist()) == sv.view(np.int32)[sk == key].tolist()


@pytest.mark.parametrize("n", [512, 1024, 4096])
def test_g7_scan(oracle, n):
    # tests/python/test_scan.py:9-63
    rng = np.random.default_rng(n)
    a = (rng.random(n) * n).astype(np.int32)
    out = oracle.inclusive_scan(a)
    cur = 0
    for i in range(n):
        cur += int(a[i])
        assert out[i] == cur
    big = rng.integers(-2**31, 2**31, 100000, dtype=np.int64).astype(np.int32)
    assert np.array_equal(oracle.inclusive_scan(big), np.cumsum(big, dtype=np.int32))


@pytest.mark.parametrize("dtype", [np.uint32, np.int32, np.float32])
@pytest.mark.parametrize("n", [257, 1024, 65536])
def test_g8_stable_sort_vs_numpy(oracle, dtype, n):
    # tests/python/test_algorithms.py:541-590: keys + u32 payload == np.argsort(kind="stable"), seed 1357,
    # floats include +-0, +-inf, denormals
    rng = np.random.default_rng(1357)
    if dtype == np.float32:
        keys = rng.standard_normal(n).astype(np.float32)
        keys[:8] = [0.0, -0.0, np.inf, -np.inf, 1e-45, -1e-45, 1.0, -1.0]
        rng.shuffle(keys)
    else:
        info = np.iinfo(dtype)
        keys = rng.integers(info.min, info.max, n, dtype=np.int64, endpoint=True).astype(dtype)
        keys[: n // 4] = keys[n // 4 : 2 * (n // 4)]  # ties
    payload = np.arange(n, dtype=np.uint32)
    k, v = oracle.stable_sort(keys, payload)
    if dtype == np.float32:
        # total order on the bit pattern: -0.0 sorts before +0.0 (numpy treats them as equal), so compare to a
        # stable argsort on the order-preserving integer image of the key
        u = keys.view(np.uint32)
        img = np.where(u >> 31 == 1, ~u, u | np.uint32(0x80000000))
        order = np.argsort(img, kind="stable")
    else:
        order = np.argsort(keys, kind="stable")
    assert np.array_equal(k.view(np.uint32), keys[order].view(np.uint32))
    assert np.array_equal(v, payload[order])


def test_g9_spmv_linear_solve_script(oracle):
    # misc/spMv_linear_solve.py:5-39: A = 2 I_8, b = [1,0,0,1,0,0,1,0] => A@b = 2b
    n = 8
    buf = make_triplet_buffer(np.arange(n), np.arange(n), np.full(n, 2.0, np.float32), np.float32, capacity=100)
    outer, inner, vals = oracle.build_matrix(buf, n, n, False)
    b = np.array([1, 0, 0, 1, 0, 0, 1, 0], np.float32)
    assert np.array_equal(oracle.spmv(outer, inner, vals, b, n, n, False), 2 * b)
    x, ok, it, err = oracle.cg(outer, inner, vals, b, 50, 1e-6)
    assert ok and np.allclose(x, b / 2)


def test_duplicates_and_explicit_zeros(oracle):
    """SURVEY 8(c) probe: 4x4 tridiagonal with duplicated diagonal triplets; explicit zeros are kept."""
    r = [0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 1]
    c = [0, 0, 1, 0, 1, 1, 2, 1, 2, 2, 3, 2, 3, 3, 3]
    v = [2, 2, -1, -1, 2, 2, -1, -1, 2, 2, -1, -1, 2, 2, 0]
    buf = make_triplet_buffer(r, c, np.array(v, np.float32), np.float32)
    outer, inner, vals = oracle.build_matrix(buf, 4, 4, True)
    assert outer.tolist() == [0, 2, 6, 9, 11]
    assert inner.tolist() == [0, 1, 0, 1, 2, 3, 1, 2, 3, 2, 3]
    assert vals.tolist() == [4, -1, -1, 4, -1, 0, -1, 4, -1, -1, 4]
