#!/usr/bin/env python
"""CPU-only: where does the C2 solve (3-D Poisson 256^3, b = 1, Jacobi-PCG from x = 0) stand after 100 iterations?

  * in exact-ish arithmetic (the same recurrence in fp64, numpy/scipy),
  * in the reference's own fp32 run (Eigen 3.4.90 through oracle/_ref, float dot products).

The bench line's e2e.final_rel_residual is the same quantity from the GPU (fp32 vectors, dot products accumulated in
double): profiles/r01s_bench.json has 3.33687, r01l 3.33700 -- within 4e-5 of the fp64 trajectory (3.33693), while
Eigen's fp32 run has drifted to 3.2624 (2.2 %).  (The residual of CG is not monotone: > 1 this early is expected.)
Takes about two minutes; prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import scipy.sparse as sp

import oracle
from bench import poisson3d_csr_host


def main():
    n, iters = 256, 100
    indptr, indices, vals = poisson3d_csr_host(n)
    N = n ** 3
    A = sp.csr_matrix((vals.astype(np.float64), indices, indptr), shape=(N, N))
    b = np.ones(N)
    dinv = 1.0 / A.diagonal()
    x, r = np.zeros(N), b.copy()
    p = dinv * r
    abs_new, rhs2 = r @ p, b @ b
    for _ in range(iters):  # conjugate_gradient.h:36-49 / Eigen ConjugateGradient.h:29-85
        tmp = A @ p
        alpha = abs_new / (p @ tmp)
        x += alpha * p
        r -= alpha * tmp
        z = dinv * r
        abs_old, abs_new = abs_new, r @ z
        p = z + (abs_new / abs_old) * p
    out = {"iterations": iters, "fp64_recurrence": float(np.sqrt((r @ r) / rhs2))}
    if oracle.EigenRef.available():
        m = oracle.EigenRef().matrix_from_compressed(indptr, indices, vals, N, N, row_major=False)
        xe, ok, it, err = m.cg(np.ones(N, np.float32), iters, 1e-30)
        out["eigen_fp32_error"] = float(err)
        out["eigen_fp32_true_residual"] = float(np.linalg.norm(b - A @ xe.astype(np.float64)) / np.sqrt(rhs2))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
