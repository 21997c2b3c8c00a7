"""Generates tests/golden/solver_golden.npz: solutions of the A16 (SparseSolver) cases by the reference's own Eigen
solvers -- SimplicialLLT / SimplicialLDLT / SparseLU with AMD and COLAMD orderings, f32 and f64, exactly the twelve types
of quadrants/program/sparse_solver.h:9-20,60-71 called as sparse_solver.cpp:86-107 calls them -- through
oracle/_ref/libeigen_ref.so (compiled from /root/reference/external/eigen by oracle/Makefile).  Run in the build
container (needs /root/reference); the GPU box only reads the .npz.

    python tests/golden/make_solver_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from tests import workloads as W  # noqa: E402


def main():
    oracle.build()
    ref = oracle.EigenRef()
    out = {}
    for name, n, r, c, v, b, sym in W.solver_cases():
        for dt, tag in ((np.float32, "f32"), (np.float64, "f64")):
            buf = oracle.make_triplet_buffer(r, c, v.astype(dt), dt)
            m = ref.matrix(buf, n, n, row_major=False)
            for st in (("LLT", "LDLT", "LU") if sym else ("LU",)):
                for od in ("AMD", "COLAMD"):
                    x, ok = m.direct_solve(b.astype(dt), st, od)
                    assert ok, (name, tag, st, od)
                    out[f"{name}/{tag}/{st}/{od}"] = x
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "solver_golden.npz")
    np.savez_compressed(path, **out)
    print(path, len(out), "solutions,", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
