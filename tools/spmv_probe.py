#!/usr/bin/env python
"""SpMV / CG timing probe on the C2 matrix (3-D Poisson n^3 fp32): canonical 4-byte indices against the one-byte diagonal
codes (qdsp_csr_spmv_config), same process, same matrix.  python tools/spmv_probe.py [n=256]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
import quadrants_b200 as qd


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    qd.init(arch=qd.cuda, device=0)
    dev = torch.device("cuda", 0)
    r, c, v = bench.poisson3d_triplets_device(n, 0, n, dev)
    N = n ** 3
    K = qd.linalg.SparseMatrixBuilder(N, N, max_num_triplets=int(r.shape[0]), dtype=qd.f32)
    K.ptr.insert_triplets(r, c, v)
    A = K.build()
    del K, r, c, v
    x, y, b = qd.ndarray(qd.f32, N), qd.ndarray(qd.f32, N), qd.ndarray(qd.f32, N)
    x.tensor.uniform_(-1, 1)
    b.fill(1.0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ref = None
    for compress in (0, 1, 0, 1):
        A.matrix.spmv_config(compress)
        info = A.matrix.spmv_info()
        for _ in range(5):
            A.matrix.spmv(None, x, y)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(50):
            A.matrix.spmv(None, x, y)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 50
        same = True if ref is None else bool(torch.equal(ref, y.tensor))
        ref = y.tensor.clone() if ref is None else ref
        cg = qd.linalg.SparseCG(A, b, max_iter=100, atol=1e-30)
        xs = qd.ndarray(qd.f32, N)
        for _ in range(2):
            cg.cg_solver.run_fixed(xs, b, 100)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            cg.cg_solver.run_fixed(xs, b, 100)
        e1.record()
        torch.cuda.synchronize()
        its = 500 / (e0.elapsed_time(e1) * 1e-3)
        del cg
        print(json.dumps({"compress": compress, "index_bytes": info["index_bytes_per_nnz"], "spmv_ms": round(ms, 4),
                          "cg_iters_per_s": round(its, 1), "y_bit_identical_to_first": same}), flush=True)


if __name__ == "__main__":
    main()
