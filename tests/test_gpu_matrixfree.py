"""N4: qd.linalg.MatrixFreeCG / MatrixFreeBICGSTAB over the vector kernels of krylov.cu.  Ports of the reference's
tests/python/test_matrixfree_cg.py and test_matrixfree_bicgstab.py (same operator, same right-hand side, same check),
the vector kernels against numpy, and the solvers with a SparseMatrix as the operator against SparseCG / numpy."""
import ctypes as C
import math

import numpy as np
import pytest

from tests import workloads as W

pytestmark = pytest.mark.gpu
GRID = 32


@pytest.fixture(scope="module")
def qd():
    import quadrants_b200 as qd

    qd.init(arch=qd.cuda)
    return qd


def stencil_operator(qd):
    import torch

    def compute_Ax(v, mv):  # test_matrixfree_cg.py:32-40: mv = 20 v - l - r - t - b on a GRID x GRID grid
        a = v.tensor.view(GRID, GRID)
        out = 20 * a
        out[1:, :] -= a[:-1, :]
        out[:-1, :] -= a[1:, :]
        out[:, 1:] -= a[:, :-1]
        out[:, :-1] -= a[:, 1:]
        mv.tensor.copy_(out.reshape(-1))

    return qd.linalg.LinearOperator(compute_Ax), compute_Ax, torch


@pytest.mark.parametrize("solver", ["cg", "bicgstab"])
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_matrixfree_reference_case(qd, solver, dt):
    qdt = qd.f32 if dt == "f32" else qd.f64
    A, compute_Ax, torch = stencil_operator(qd)
    Ax = qd.field(dtype=qdt, shape=(GRID, GRID))
    x = qd.field(dtype=qdt, shape=(GRID, GRID))
    b = qd.field(dtype=qdt, shape=(GRID, GRID))
    i, j = np.meshgrid(np.arange(GRID), np.arange(GRID), indexing="ij")
    b.from_numpy(np.sin(2 * math.pi * i / (GRID - 1)) * np.sin(2 * math.pi * j / (GRID - 1)))
    fn = qd.linalg.MatrixFreeCG if solver == "cg" else qd.linalg.MatrixFreeBICGSTAB
    fn(A, b, x, maxiter=10 * GRID * GRID, tol=1e-18, quiet=True)
    compute_Ax(x, Ax)
    assert np.all(np.abs(Ax.to_numpy() - b.to_numpy()) < 1e-6)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_vector_kernels_against_numpy(qd, dt):
    from quadrants_b200 import _lib
    from quadrants_b200.linalg.matrixfree_cg import _Vec

    qdt = qd.f32 if dt == np.float32 else qd.f64
    for n in (1, 31, 1000, 1 << 20, (1 << 22) + 13):
        rng = np.random.default_rng(n)
        arrs = [rng.uniform(-1, 1, n).astype(dt) for _ in range(4)]
        nd = []
        for a in arrs:
            v = qd.ndarray(qdt, n)
            v.from_numpy(a)
            nd.append(v)
        V = _Vec(qdt, n)
        a64 = [a.astype(np.float64) for a in arrs]
        d0, d1 = V.dot2(nd[0], nd[1], nd[2], nd[3])
        assert abs(d0 - a64[0] @ a64[1]) <= 1e-12 * n and abs(d1 - a64[2] @ a64[3]) <= 1e-12 * n
        assert V.dot(nd[0], nd[0]) == V.dot(nd[0], nd[0])  # reproducible
        out = qd.ndarray(qdt, n)
        V.lincomb(out, (0.5, nd[0]), (-2.0, nd[1]), (3.0, nd[2]))
        ref = dt(0.5) * arrs[0] + dt(-2.0) * arrs[1] + dt(3.0) * arrs[2]
        assert np.allclose(out.to_numpy(), ref, rtol=1e-6 if dt == np.float32 else 1e-14, atol=1e-6 if dt == np.float32 else 1e-14)
        V.lincomb(nd[3], (1.0, nd[3]), (2.0, nd[0]))  # aliasing
        assert np.allclose(nd[3].to_numpy(), arrs[3] + dt(2.0) * arrs[0], rtol=1e-6, atol=1e-6)
        rr = V.cg_step(nd[0], nd[1], nd[2], out, 0.25)
        x1 = arrs[0] + dt(0.25) * arrs[2]
        r1 = arrs[1] - dt(0.25) * ref
        assert np.allclose(nd[0].to_numpy(), x1, rtol=1e-6, atol=1e-6) and np.allclose(nd[1].to_numpy(), r1, rtol=1e-5, atol=1e-5)
        assert abs(rr - r1.astype(np.float64) @ r1.astype(np.float64)) <= 1e-4 * max(1.0, rr)


def test_matrixfree_with_sparse_matrix_operator(qd):
    """The operator is a SparseMatrix's SpMV: MatrixFreeCG must land on the solution SparseCG / numpy give."""
    nx = 48
    n = nx * nx
    r, c, v = W.poisson2d_triplets(nx, dtype=np.float64)
    K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=r.shape[0], dtype=qd.f64)
    K.insert_triplets(r.astype(np.int32), c.astype(np.int32), v)
    Am = K.build()
    A = qd.linalg.LinearOperator(lambda vin, vout: Am.matrix.spmv(None, vin, vout))
    xt = np.random.default_rng(1).uniform(-1, 1, n)
    xd = qd.ndarray(qd.f64, n)
    xd.from_numpy(xt)
    b = Am @ xd
    x = qd.ndarray(qd.f64, n)
    assert qd.linalg.MatrixFreeCG(A, b, x, tol=1e-10, maxiter=2000)
    assert np.linalg.norm(x.to_numpy() - xt) <= 1e-8 * np.linalg.norm(xt)
    x2 = qd.ndarray(qd.f64, n)
    assert qd.linalg.MatrixFreeBICGSTAB(A, b, x2, tol=1e-10, maxiter=2000)
    assert np.linalg.norm(x2.to_numpy() - xt) <= 1e-7 * np.linalg.norm(xt)
    # a warm start from the solution does nothing and succeeds
    assert qd.linalg.MatrixFreeCG(A, b, x, tol=1e-6, maxiter=5)


def test_matrixfree_bicgstab_unsymmetric_and_failures(qd):
    import torch

    n = 500
    rng = np.random.default_rng(3)
    m = rng.uniform(-1, 1, (n, n)) * (rng.random((n, n)) < 0.02)
    m[np.arange(n), np.arange(n)] = 1.0 + np.abs(m).sum(1)
    M = torch.from_numpy(m).cuda()
    A = qd.linalg.LinearOperator(lambda vin, vout: vout.tensor.copy_(M @ vin.tensor))
    b = qd.ndarray(qd.f64, n)
    b.from_numpy(np.cos(np.arange(n) * 0.1))
    x = qd.ndarray(qd.f64, n)
    assert qd.linalg.MatrixFreeBICGSTAB(A, b, x, tol=1e-10, maxiter=500)
    assert np.linalg.norm(m @ x.to_numpy() - b.to_numpy()) <= 1e-8
    x0 = qd.ndarray(qd.f64, n)
    assert not qd.linalg.MatrixFreeCG(A, b, x0, tol=1e-30, maxiter=3)  # budget exhausted
    from quadrants_b200.linalg.matrixfree_cg import QuadrantsTypeError

    with pytest.raises(QuadrantsTypeError):
        qd.linalg.MatrixFreeCG(A, b, qd.ndarray(qd.f32, n))
    with pytest.raises(qd.QuadrantsRuntimeError):
        qd.linalg.MatrixFreeCG(A, b, qd.ndarray(qd.f64, n + 1))
    with pytest.raises(qd.QuadrantsRuntimeError):
        A.matvec(b, qd.ndarray(qd.f64, n + 1))
