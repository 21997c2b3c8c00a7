#!/usr/bin/env python
"""One small invocation of the round-2 kernels outside the CG path -- direct solver (factor, forward, backward), row-merge
A + B, the Krylov vector kernels -- for an ncu capture:
    ncu --set full --clock-control none -k regex:"solver_|geam_|vec_" -c 12 -o out python tools/profile_new_kernels.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

import quadrants_b200 as qd
from tests import workloads as W


def main():
    qd.init(arch=qd.cuda, device=0)
    nx = 256
    n = nx * nx
    r, c, v = W.poisson2d_triplets(nx, dtype=np.float64)
    K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=int(r.shape[0]), dtype=qd.f64)
    K.insert_triplets(r.astype(np.int32), c.astype(np.int32), v)
    A = K.build()
    b = qd.ndarray(qd.f64, n)
    b.fill(1.0)
    s = qd.linalg.SparseSolver(dtype=qd.f64, solver_type="LLT")
    s.compute(A)
    x = s.solve(b)
    B = A + A
    op = qd.linalg.LinearOperator(lambda vin, vout: A.matrix.spmv(None, vin, vout))
    xm = qd.ndarray(qd.f64, n)
    qd.linalg.MatrixFreeCG(op, b, xm, tol=1e-30, maxiter=3)
    print("ok", float(np.abs((A @ x).to_numpy() - 1.0).max()), B.matrix.nnz())


if __name__ == "__main__":
    main()
