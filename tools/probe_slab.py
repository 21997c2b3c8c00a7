#!/usr/bin/env python
"""Probe (not a bench line): the per-rank work of the multi-GPU CG on ONE GPU, without any communication.
A 256 x 256 x nz slab of the 7-point Poisson grid is what one of 256/nz ranks owns in C2; timing its fixed-budget CG
here separates kernel/launch overheads from NVLink costs.   python tools/probe_slab.py [nz ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import quadrants_b200 as qd


def slab_triplets(nz, n, dev):
    i = torch.arange(nz, device=dev, dtype=torch.int64).view(-1, 1, 1)
    j = torch.arange(n, device=dev, dtype=torch.int64).view(1, -1, 1)
    k = torch.arange(n, device=dev, dtype=torch.int64).view(1, 1, -1)
    r = ((i * n + j) * n + k).reshape(-1)
    rows, cols, vals = [], [], []
    for d, (di, dj, dk) in enumerate(((0, 0, 0), (-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1))):
        ii, jj, kk = i + di, j + dj, k + dk
        ok = ((ii >= 0) & (ii < nz) & (jj >= 0) & (jj < n) & (kk >= 0) & (kk < n)).reshape(-1)
        c = ((ii * n + jj) * n + kk).reshape(-1)
        rows.append(r[ok]); cols.append(c[ok])
        vals.append(torch.full((int(ok.sum()),), 6.0 if d == 0 else -1.0, device=dev, dtype=torch.float32))
    return torch.cat(rows), torch.cat(cols), torch.cat(vals)


def main():
    qd.init(arch=qd.cuda, device=0)
    dev = torch.device("cuda", 0)
    n = 256
    for nz in [int(a) for a in sys.argv[1:]] or [32, 64, 128, 256]:
        r, c, v = slab_triplets(nz, n, dev)
        rows = nz * n * n
        K = qd.linalg.SparseMatrixBuilder(rows, rows, max_num_triplets=int(r.shape[0]), dtype=qd.f32)
        K.ptr.insert_triplets(r, c, v)
        del r, c, v
        A = K.build()
        del K
        b = qd.ndarray(qd.f32, rows); b.fill(1.0)
        x = qd.ndarray(qd.f32, rows)
        cg = qd.linalg.SparseCG(A, b, max_iter=400, atol=1e-30)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            cg.cg_solver.run_fixed(x, b, 400)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            cg.cg_solver.run_fixed(x, b, 400)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 2000
        nnz = A.matrix.nnz()
        byts = nnz * 8 + (rows + 1) * 4 + 12 * rows * 4
        print(f"slab nz={nz}: rows={rows} nnz={nnz} graph-replayed CG {us:.2f} us/iter  ({byts / us / 1e3:.0f} GB/s of the 12n model; "
              f"roofline at 6.5 TB/s {byts / 6.5e6:.1f} us)", flush=True)
        os.environ["QDSP_CG_TRACE"] = "1"
        cg.cg_solver.run_fixed(x, b, 200)
        torch.cuda.synchronize()
        del os.environ["QDSP_CG_TRACE"]
        del cg, A, b, x


if __name__ == "__main__":
    main()
