#!/usr/bin/env python
"""Times SparseMatrixBuilder.build() of the C2 matrix (256^3 Poisson, 117 M triplets) three times in one process:
the first is cold (lazy kernel loading, first cudaMallocs), the others warm."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import quadrants_b200 as qd
from bench import poisson3d_triplets_device

qd.init(arch=qd.cuda, device=0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
r, c, v = poisson3d_triplets_device(n, 0, n, "cuda:0")
out = []
for it in range(3):
    K = qd.linalg.SparseMatrixBuilder(n ** 3, n ** 3, max_num_triplets=int(r.shape[0]), dtype=qd.f32)
    K.ptr.insert_triplets(r, c, v)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    A = K.build()
    e1.record()
    torch.cuda.synchronize()
    out.append(round(e0.elapsed_time(e1), 3))
    del A, K
print(json.dumps({"config": f"C2 build {n}^3", "triplets": int(r.shape[0]), "build_ms": out}))
