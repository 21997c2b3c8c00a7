"""Generates tests/golden/eigen_golden.npz from the reference's vendored Eigen (oracle/_ref/libeigen_ref.so, built from
/root/reference/external/eigen by oracle/Makefile with the reference's compile flags).  Run in a container that has
/root/reference:   python tests/golden/make_golden.py
The fixtures pin (a) the C restatement in oracle/qd_oracle.c and (b) the CUDA path, on machines where the reference
tree (and possibly oracle/_ref) is absent."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402
from tests import workloads as W  # noqa: E402


def cases():
    rng = np.random.default_rng(0)
    # C1-mini: 2-D Poisson 24x24, diagonal split in two triplets, order permuted
    for dt in (np.float32, np.float64):
        r, c, v = W.poisson2d_triplets(24, dtype=dt, split_diag=True)
        p = rng.permutation(r.shape[0])
        yield f"poisson2d_{np.dtype(dt).name}", r[p], c[p], v[p], 576, 576, dt, True
    # C4-mini: random coordinates with ~30 % repeats, non-square
    for dt in (np.float32, np.float64):
        r, c, v = W.random_triplets(300, 200, 1400, 600, seed=6, dtype=dt)
        yield f"random_{np.dtype(dt).name}", r, c, v, 300, 200, dt, False
    # C5-mini: power-law FEM stiffness
    r, c, v, n = W.fem_tet_triplets(400, 3000, seed=7, dtype=np.float64)
    yield "fem_float64", r, c, v, n, n, np.float64, True
    r, c, v, n = W.fem_tet_triplets(400, 3000, seed=7, dtype=np.float32)
    yield "fem_float32", r, c, v, n, n, np.float32, True


def main():
    ref = O.EigenRef()
    out = {"eigen_version": np.array(ref.version())}
    rng = np.random.default_rng(1)
    for name, r, c, v, rows, cols, dt, spd in cases():
        buf = O.make_triplet_buffer(r, c, v, dt)
        out[f"{name}/buf"] = buf
        out[f"{name}/shape"] = np.array([rows, cols])
        x = rng.uniform(-1, 1, cols).astype(dt)
        out[f"{name}/x"] = x
        for rm in (True, False):
            m = ref.matrix(buf, rows, cols, rm)
            outer, inner, vals = m.export()
            tag = "rm" if rm else "cm"
            out[f"{name}/{tag}/outer"], out[f"{name}/{tag}/inner"], out[f"{name}/{tag}/vals"] = outer, inner, vals
            out[f"{name}/{tag}/y"] = m.spmv(x)
        if spd:
            m = ref.matrix(buf, rows, cols, False)
            b = np.ones(rows, dt)
            tol = 1e-6 if dt == np.float32 else 1e-10
            xs, ok, it, err = m.cg(b, 2000, tol)
            out[f"{name}/cg/x"], out[f"{name}/cg/b"] = xs, b
            out[f"{name}/cg/meta"] = np.array([float(ok), float(it), err, tol, 2000.0])
            print(name, "cg", ok, it, err)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "eigen_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
