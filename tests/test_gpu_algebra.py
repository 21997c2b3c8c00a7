"""N3: the device matrix algebra (A +- B, alpha A, A^T, A B; quadrants/program/sparse_matrix.cpp:413-652) beyond the
reference's 5x4 toy test (tests/python/test_sparse_matrix.py:454-523), against scipy on the same inputs.

Bars: indptr / indices bit-exact for every op (canonical CSR: rows ascending, columns ascending, structural zeros from
cancellation kept like cusparse's csrgeam2 / SpGEMM keep them -- compared after scipy's pattern is taken WITHOUT
eliminate_zeros); values bit-exact for +, -, scaling and transpose (one rounding each, no reassociation); A B within
1e-5 (fp32) / 1e-12 (fp64) of sum |a_ik b_kj| (the summation order differs from scipy's).  Needs a B200.
"""
import os

import numpy as np
import pytest
import scipy.sparse as sp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def qd():
    import quadrants_b200 as qd

    qd.init(arch=qd.cuda)
    return qd


def rand_csr(n, m, nnz, seed, dt):
    rng = np.random.default_rng(seed)
    r = rng.integers(0, n, nnz)
    c = rng.integers(0, m, nnz)
    v = rng.uniform(-1, 1, nnz).astype(dt)
    a = sp.coo_matrix((v, (r, c)), shape=(n, m)).tocsr()  # duplicates summed (few)
    a.sort_indices()
    return a


def to_qd(qd, a):
    dt = qd.f32 if a.dtype == np.float32 else qd.f64
    coo = a.tocoo()
    K = qd.linalg.SparseMatrixBuilder(a.shape[0], a.shape[1], max_num_triplets=max(coo.nnz, 1), dtype=dt)
    K.insert_triplets(coo.row.astype(np.int32), coo.col.astype(np.int32), coo.data)
    return K.build()


def arrays(Q):
    ip, ix, v = Q.matrix.to_scipy_arrays()
    return ip, ix, v


def same_pattern(Q, S):
    S = S.tocsr()
    S.sort_indices()
    ip, ix, v = arrays(Q)
    assert Q.shape == S.shape
    assert np.array_equal(ip, S.indptr) and np.array_equal(ix, S.indices)
    return v, S.data


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32 if a.dtype.itemsize == 4 else np.uint64)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_add_sub_scale_transpose_200k(qd, dt):
    n, m = 200_000, 150_000
    S1, S2 = rand_csr(n, m, 2_000_000, 1, dt), rand_csr(n, m, 1_500_000, 2, dt)
    A, B = to_qd(qd, S1), to_qd(qd, S2)
    v, ref = same_pattern(A, S1)
    assert np.array_equal(bits(v), bits(ref))
    for Q, S in ((A + B, S1 + S2), (A - B, S1 - S2), (A * 2.5, S1 * dt(2.5)), (2.5 * A, S1 * dt(2.5)),
                 (A.transpose(), S1.T)):
        v, ref = same_pattern(Q, S)
        assert np.array_equal(bits(v), bits(ref.astype(dt)))
    # A - A: every entry cancels; the pattern stays (csrgeam2 keeps structural zeros)
    Z = A - A
    ip, ix, v = arrays(Z)
    assert np.array_equal(ip, S1.indptr) and np.array_equal(ix, S1.indices) and not v.any()


def check_product(Q, S1, S2, dt):
    P = (S1 @ S2).tocsr()
    P.sort_indices()
    scale = (abs(S1) @ abs(S2)).tocsr()
    scale.sort_indices()
    ip, ix, v = arrays(Q)
    # scipy drops nothing either: the pattern of |A| |B| is the structural pattern
    assert np.array_equal(ip, scale.indptr) and np.array_equal(ix, scale.indices)
    Pd = np.zeros(scale.nnz, np.float64)
    # align P (which may have dropped exact-zero sums on some scipy versions) onto the structural pattern
    key_s = np.repeat(np.arange(scale.shape[0]), np.diff(scale.indptr)).astype(np.int64) * scale.shape[1] + scale.indices
    key_p = np.repeat(np.arange(P.shape[0]), np.diff(P.indptr)).astype(np.int64) * P.shape[1] + P.indices
    Pd[np.searchsorted(key_s, key_p)] = P.data
    eps = 1e-5 if dt == np.float32 else 1e-12
    assert np.all(np.abs(v.astype(np.float64) - Pd) <= eps * scale.data + 1e-300)
    return ip, ix, v


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_matmul_200k_and_chunking_is_invisible(qd, dt):
    n = 200_000
    S1, S2 = rand_csr(n, n, 1_200_000, 3, dt), rand_csr(n, n, 1_400_000, 4, dt)
    A, B = to_qd(qd, S1), to_qd(qd, S2)
    ip, ix, v = check_product(A @ B, S1, S2, dt)
    os.environ["QDSP_SPGEMM_CHUNK"] = "1000000"  # ~9 chunks
    try:
        ip2, ix2, v2 = arrays(A @ B)
    finally:
        del os.environ["QDSP_SPGEMM_CHUNK"]
    assert np.array_equal(ip, ip2) and np.array_equal(ix, ix2) and np.array_equal(bits(v), bits(v2))
    # (A B)^T == B^T A^T structurally and to rounding
    T1 = (A @ B).transpose()
    T2 = B.transpose() @ A.transpose()
    a1, a2 = arrays(T1), arrays(T2)
    assert np.array_equal(a1[0], a2[0]) and np.array_equal(a1[1], a2[1])
    assert np.allclose(a1[2], a2[2], rtol=1e-4 if dt == np.float32 else 1e-11, atol=1e-6 if dt == np.float32 else 1e-13)


def test_matmul_poisson_1m_rows(qd):
    """A A for the 1000 x 1000 5-point Laplacian (10^6 rows, 5 M nnz -> 13-point stencil, 25 M partial products)."""
    nx = 1000
    T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(nx, nx), format="csr")
    I = sp.identity(nx, format="csr")
    S = (sp.kron(I, T) + sp.kron(T, I)).tocsr()
    S.eliminate_zeros()
    S.sort_indices()
    A = to_qd(qd, S)
    ip, ix, v = check_product(A @ A, S, S, np.float64)
    assert ip[-1] == (S @ S).nnz
    # integer-valued matrix: the product is exact in any order
    P = (S @ S).tocsr()
    P.sort_indices()
    assert np.array_equal(v, P.data)


def test_matmul_errors_and_empty(qd):
    S1 = rand_csr(50, 40, 300, 5, np.float32)
    S2 = rand_csr(30, 20, 100, 6, np.float32)
    A, B = to_qd(qd, S1), to_qd(qd, S2)
    with pytest.raises(AssertionError):
        A @ B
    E = to_qd(qd, sp.csr_matrix((40, 20), dtype=np.float32))
    Q = A @ E
    assert Q.shape == (50, 20) and arrays(Q)[0][-1] == 0


def test_add_of_non_canonical_csr_takes_the_sort_path(qd):
    """A matrix adopted through qdsp_csr_from_device_csr may have unsorted rows: A + B must still be right (the merge
    kernel notices and the expand-sort-collapse path runs instead)."""
    import ctypes as C

    import torch

    from quadrants_b200 import _core, _lib

    lib = _lib.load()
    ip = np.array([0, 3, 5, 5, 6], np.int32)
    ix = np.array([2, 0, 1, 3, 1, 0], np.int32)           # rows 0 and 1 are not ascending
    v = np.array([1, 2, 3, 4, 5, 6], np.float32)
    t = [torch.from_numpy(a).cuda() for a in (ip, ix, v)]
    h = C.c_void_p()
    stream = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.qdsp_csr_from_device_csr(4, 4, 6, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), _lib.F32, 0,
                                            stream, C.byref(h)))
    A = qd.linalg.SparseMatrix(sm=_core.CuSparseMatrix._wrap(h), dtype=qd.f32)
    S = sp.csr_matrix((v, ix, ip), shape=(4, 4))
    Q = A + A
    ip2, ix2, v2 = arrays(Q)
    R = (S + S).tocsr()
    R.sort_indices()
    assert np.array_equal(ip2, R.indptr) and np.array_equal(ix2, R.indices) and np.array_equal(v2, R.data)


def test_scratch_pool_gives_way_when_the_device_fills_up(qd):
    """blocks parked in the scratch pool are invisible to every other allocator on the device (torch's in the
    host application).  The pool therefore never keeps more than a quarter of the memory that would be free without it,
    re-checked whenever a block comes back (qdsp_scratch_info reports both numbers)."""
    import ctypes as C
    import time

    import torch

    from quadrants_b200 import _lib

    lib = _lib.load()

    def info():
        cached, cap = C.c_size_t(0), C.c_size_t(0)
        _lib.check(lib.qdsp_scratch_info(C.byref(cached), C.byref(cap)))
        return cached.value, cap.value

    if os.environ.get("QDSP_SCRATCH_POOL_MB") == "0":
        pytest.skip("pool disabled")
    _lib.check(lib.qdsp_release_scratch())
    assert info()[0] == 0
    A = to_qd(qd, rand_csr(300_000, 300_000, 3_000_000, 5, np.float32))
    At = A.transpose()
    cached, cap = info()
    assert 0 < cached <= cap, (cached, cap)          # ample memory: the temporaries of the transpose stay cached
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    keep = 2 * cached                                 # leaves the pool above a quarter of (free + cached): it has to shrink
    hog = torch.empty(max(free - keep, 0), dtype=torch.uint8, device="cuda")  # the host application takes the device
    try:
        time.sleep(0.05)                              # (the pool re-reads the free memory at most every 20 ms)
        At2 = A.transpose()                           # re-uses cached blocks and hands them back: the cap is re-applied
        torch.cuda.synchronize()
        cached2, cap2 = info()
        free2, _ = torch.cuda.mem_get_info()
        assert cap2 <= (free2 + cached2) // 4 + (2 << 20), (cap2, free2, cached2)
        assert cached2 <= cap2 and cached2 < cached, (cached, cached2, cap2)
        assert np.array_equal(arrays(At2)[1], arrays(At)[1])
    finally:
        del hog
        torch.cuda.empty_cache()
