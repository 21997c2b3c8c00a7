#!/usr/bin/env python
"""Debug aid: fixed-budget CG, persistent kernel vs three-kernel iteration, on the stream-path FEM matrix (fp32 / fp64) and on
a regular matrix forced to the stream path; prints |x_fused - x_classic| / |x_classic| after k iterations."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import quadrants_b200 as qd
from oracle import make_triplet_buffer
from tests import workloads as W


def build(buf, n, dt):
    K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=int(buf[0]), dtype=dt)
    K.ptr.fill_from_host(buf)
    return K.build()


def nd(a, dt):
    x = qd.ndarray(dt, a.shape[0])
    x.from_numpy(a)
    return x


def compare(name, A, b, dt, ks=(1, 2, 3, 5, 10, 40, 200)):
    n = b.shape[0]
    res = {}
    for mode in ("1", "0"):
        os.environ["QDSP_CG_FUSED"] = mode
        cg = qd.linalg.SparseCG(A, nd(b, dt), max_iter=10, atol=1e-30)
        for k in ks:
            x = qd.ndarray(dt, n)
            cg.cg_solver.run_fixed(x, nd(b, dt), k)
            torch.cuda.synchronize()
            res[(mode, k)] = x.to_numpy().astype(np.float64)
    f, c = res[("1", ks[0])], res[("0", ks[0])]
    ratio = f / np.where(c == 0, 1.0, c)
    bad = np.flatnonzero(np.abs(ratio - 1.0) > 1e-4)
    print(f"{name} k={ks[0]}: ratio x_f/x_c min {ratio.min():.6f} max {ratio.max():.6f} median {np.median(ratio):.6f}; "
          f"{bad.shape[0]} entries off by > 1e-4, first {bad[:12].tolist()} last {bad[-6:].tolist()}", flush=True)
    for k in ks:
        f, c = res[("1", k)], res[("0", k)]
        print(f"{name} k={k:4d} rel diff {np.linalg.norm(f - c) / max(np.linalg.norm(c), 1e-300):.3e}  |x_c| {np.linalg.norm(c):.4e} "
              f"nan_f {int(np.isnan(f).sum())} nan_c {int(np.isnan(c).sum())}", flush=True)
    for mode in ("1", "0"):
        os.environ["QDSP_CG_FUSED"] = mode
        cg = qd.linalg.SparseCG(A, nd(b, dt), max_iter=5000, atol=1e-5 if dt is qd.f32 else 1e-10)
        x, ok = cg.solve()
        print(f"{name} solve fused={mode}: ok={ok} iterations={cg.iterations} error={cg.error:.3e}", flush=True)


def main():
    qd.init(arch=qd.cuda)
    for dtype, dt in ((np.float32, qd.f32), (np.float64, qd.f64)):
        r, c, v, n = W.fem_tet_triplets(30000, 200000, seed=7, dtype=dtype, compact=True)
        A = build(make_triplet_buffer(r, c, v, dtype), n, dt)
        b = np.random.default_rng(9).uniform(-1, 1, n).astype(dtype)
        compare(f"fem {np.dtype(dtype).name}", A, b, dt)
    os.environ["QDSP_SPMV_PATH"] = "stream"
    r, c, v = W.poisson2d_triplets(256, dtype=np.float32)
    n = 256 * 256
    A = build(make_triplet_buffer(r, c, v, np.float32), n, qd.f32)
    compare("poisson2d f32 forced stream", A, np.ones(n, np.float32), qd.f32)
    del os.environ["QDSP_SPMV_PATH"]
    A = build(make_triplet_buffer(r, c, v, np.float32), n, qd.f32)
    compare("poisson2d f32 uniform", A, np.ones(n, np.float32), qd.f32)


if __name__ == "__main__":
    main()
