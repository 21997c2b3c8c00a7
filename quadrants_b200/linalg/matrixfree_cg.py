"""qd.linalg.LinearOperator / MatrixFreeCG / MatrixFreeBICGSTAB (python/quadrants/linalg/matrixfree_cg.py:12-310).

The reference writes both solvers in its DSL: the operator is a user kernel `matvec(x, Ax)`, every dot product a JIT
reduction returning a host scalar, every vector update its own kernel.  Here the operator is any callable acting on device
arrays (e.g. `lambda v, mv: A.matrix.spmv(None, v, mv)` for a SparseMatrix, or a torch stencil on `v.tensor`), and the
vector work runs in the kernels of quadrants_b200/csrc/krylov.cu through the C ABI (qdsp_vec_dot2 / qdsp_vec_lincomb /
qdsp_vec_cg_step): the recurrences, the start from the caller's x, the absolute tolerance on sqrt(r.r), the return value
and its quirk (the final check compares r.r, not its root, with tol) are the reference's.  Dot products are accumulated
in double and in a fixed order (the reference sums in the solver dtype, in whatever order its atomics land), and every
scalar of the recurrence -- the dot results, alpha, beta, omega, rho -- is then rounded to the solver dtype, because the
reference keeps them in fields of that dtype and its reductions return it (matrixfree_cg.py:67-68,80-86,194-198): an f32
solve follows the f32 recurrence, not a double-precision one.
"""
from __future__ import annotations

import ctypes as C
from math import sqrt

import numpy as np

from .. import _lib
from ..lang import Ndarray, QuadrantsRuntimeError, f32, f64, get_runtime


class QuadrantsTypeError(TypeError):
    """python/quadrants/lang/exception.py: raised for dtype mismatches."""


class LinearOperator:
    def __init__(self, matvec_kernel):
        self._matvec = matvec_kernel

    def matvec(self, x, Ax):
        if x.shape != Ax.shape:
            raise QuadrantsRuntimeError(f"Dimension mismatch x.shape{x.shape} != Ax.shape{Ax.shape}.")
        self._matvec(x, Ax)


class _Vec:
    """The dense vector kernels over the C ABI for one (dtype, n)."""

    def __init__(self, dtype, n):
        self.lib = _lib.load()
        self.dt = _lib.F64 if dtype is f64 else _lib.F32
        self.n = int(n)
        self.stream = get_runtime().prog.stream()
        # a scalar as the reference holds it: a field / kernel return value of the solver dtype
        self.q = float if dtype is f64 else (lambda v: float(np.float32(v)))

    def dot(self, a, b):
        out = (C.c_double * 2)()
        _lib.check(self.lib.qdsp_vec_dot2(a.data_ptr(), b.data_ptr(), None, None, self.n, self.dt, self.stream, out))
        return self.q(out[0])

    def dot2(self, a, b, c, d):
        out = (C.c_double * 2)()
        _lib.check(self.lib.qdsp_vec_dot2(a.data_ptr(), b.data_ptr(), c.data_ptr(), d.data_ptr(), self.n, self.dt,
                                          self.stream, out))
        return self.q(out[0]), self.q(out[1])

    def lincomb(self, out, *terms):
        """out = sum coef * vec over (coef, vec) terms."""
        k = len(terms)
        xs = (C.c_void_p * k)(*[v.data_ptr() for _, v in terms])
        cs = (C.c_double * k)(*[float(c) for c, _ in terms])
        _lib.check(self.lib.qdsp_vec_lincomb(out.data_ptr(), self.n, self.dt, k, xs, cs, self.stream))

    def cg_step(self, x, r, p, ap, alpha):
        rr = C.c_double()
        _lib.check(self.lib.qdsp_vec_cg_step(x.data_ptr(), r.data_ptr(), p.data_ptr(), ap.data_ptr(), float(alpha),
                                             self.n, self.dt, self.stream, C.byref(rr)))
        return self.q(rr.value)


def _check_args(b, x, name):
    if b.dtype != x.dtype:
        raise QuadrantsTypeError(f"Dtype mismatch b.dtype({b.dtype}) != x.dtype({x.dtype}).")
    if b.dtype is not f32 and b.dtype is not f64:
        raise QuadrantsTypeError(f"Not supported dtype: {b.dtype}")
    if b.shape != x.shape:
        raise QuadrantsRuntimeError(f"Dimension mismatch b.shape{b.shape} != x.shape{x.shape}.")
    if len(b.shape) > 3:
        raise QuadrantsRuntimeError(f"{name} only support 1D, 2D, 3D inputs; your inputs is {len(b.shape)}-D.")
    if not isinstance(b, Ndarray) or not isinstance(x, Ndarray):
        raise QuadrantsRuntimeError(f"{name}: b and x must be fields / ndarrays on the device")


def _like(v):
    w = Ndarray(v.dtype, v.tensor.shape[0])
    w.shape = v.shape
    return w


def MatrixFreeCG(A, b, x, tol=1e-6, maxiter=5000, quiet=True):
    """Matrix-free conjugate-gradient solver: A x = b with A a LinearOperator; starts from the x passed in; stops when
    sqrt(r.r) < tol (absolute).  Returns True on success (matrixfree_cg.py:23-139)."""
    _check_args(b, x, "MatrixFreeCG")
    n = b.tensor.shape[0]
    V = _Vec(b.dtype, n)
    p, r, Ap, Ax = _like(b), _like(b), _like(b), _like(b)
    succeeded = True
    A._matvec(x, Ax)
    V.lincomb(r, (1.0, b), (-1.0, Ax))            # init(): r = b - Ax, p = 0, Ap = 0 (fresh arrays are zero)
    initial_rTr = V.dot(r, r)
    if not quiet:
        print(f">>> Initial residual = {initial_rTr:e}")
    old_rTr = new_rTr = initial_rTr
    V.lincomb(p, (1.0, r), (0.0, p))              # update_p() with beta = 0
    if sqrt(initial_rTr) >= tol:
        for i in range(maxiter):
            A._matvec(p, Ap)
            pAp = V.dot(p, Ap)
            alpha = V.q(old_rTr / pAp)
            new_rTr = V.cg_step(x, r, p, Ap, alpha)  # update_x + update_r + reduce(r, r) in one pass
            if sqrt(new_rTr) < tol:
                if not quiet:
                    print(">>> Conjugate Gradient method converged.")
                    print(f">>> #iterations {i}")
                break
            beta = V.q(new_rTr / old_rTr)
            V.lincomb(p, (1.0, r), (beta, p))
            old_rTr = new_rTr
            if not quiet:
                print(f">>> Iter = {i + 1:4}, Residual = {sqrt(new_rTr):e}")
    if new_rTr >= tol:
        if not quiet:
            print(f">>> Conjugate Gradient method failed to converge in {maxiter} iterations: "
                  f"Residual = {sqrt(new_rTr):e}")
        succeeded = False
    return succeeded


def MatrixFreeBICGSTAB(A, b, x, tol=1e-6, maxiter=5000, quiet=True):
    """Matrix-free BiCGSTAB (matrixfree_cg.py:142-310): same conventions as MatrixFreeCG; r_tld = b."""
    _check_args(b, x, "MatrixFreeBICGSTAB")
    n = b.tensor.shape[0]
    V = _Vec(b.dtype, n)
    p, r, s, t, Ap, Ax = (_like(b) for _ in range(6))
    r_tld = b
    succeeded = True
    A._matvec(x, Ax)
    V.lincomb(r, (1.0, b), (-1.0, Ax))
    rho = 0.0
    rho_1 = alpha = beta = omega = 1.0
    initial_rTr = V.dot(r, r)
    rTr = initial_rTr
    if not quiet:
        print(f">>> Initial residual = {initial_rTr:e}")
    if sqrt(initial_rTr) >= tol:
        for i in range(maxiter):
            rho = V.dot(r, r_tld)
            if rho == 0.0:
                if not quiet:
                    print(">>> BICGSTAB failed because r@r_tld = 0.")
                succeeded = False
                break
            if i == 0:
                V.lincomb(p, (1.0, r))
            else:
                beta = V.q((rho / rho_1) * (alpha / omega))
                V.lincomb(p, (1.0, r), (beta, p), (-beta * omega, Ap))   # p = r + beta (p - omega Ap)
            A._matvec(p, Ap)                                             # p_hat = p (no preconditioner)
            alpha = V.q(rho / V.dot(r_tld, Ap))
            V.lincomb(s, (1.0, r), (-alpha, Ap))
            A._matvec(s, t)                                              # s_hat = s; t = A s_hat
            omega_upper, omega_lower = V.dot2(t, s, t, t)
            omega = V.q(omega_upper / (omega_lower + 1e-16) if omega_lower == 0.0 else omega_upper / omega_lower)
            V.lincomb(x, (1.0, x), (alpha, p), (omega, s))
            V.lincomb(r, (1.0, s), (-omega, t))
            rTr = V.dot(r, r)
            if not quiet:
                print(f">>> Iter = {i + 1:4}, Residual = {sqrt(rTr):e}")
            if sqrt(rTr) < tol:
                if not quiet:
                    print(f">>> BICGSTAB method converged at #iterations {i}")
                break
            rho_1 = rho
    if rTr >= tol:
        if not quiet:
            print(f">>> BICGSTAB failed to converge in {maxiter} iterations: Residual = {sqrt(rTr):e}")
        succeeded = False
    return succeeded


__all__ = ["LinearOperator", "MatrixFreeCG", "MatrixFreeBICGSTAB"]
