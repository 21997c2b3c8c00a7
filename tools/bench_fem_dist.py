#!/usr/bin/env python
"""C5 of BASELINE.json at 1/2/4/8 GPUs: irregular FEM tetrahedral stiffness matrix (power-law row lengths), fp64
Jacobi-PCG, row blocks chosen by nnz balance.  NOT the driver's bench line (that is C2, bench.py); run it as

    python tools/bench_fem_dist.py                                                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29540 tools/bench_fem_dist.py [n_vertices=400000] [n_tets=2000000] [iters=200]

Every rank generates the same triplets on the host (seeded), keeps those of its rows, assembles them on its GPU through
the builder (global column ids) and localizes the matrix.  Checks, before timing: the distributed SpMV against the
host CSR product of the rank's rows (1e-12 * sum |a_ij x_j| per row, the BASELINE bar for fp64) and a converged solve
(tol 1e-10: same iteration count on every rank, true residual).  Timing: `iters` fixed-budget iterations, CUDA events,
barrier + synchronize on both sides, max over ranks.  One JSON line on rank 0."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import scipy.sparse as sp
import torch
import torch.distributed as dist

import quadrants_b200 as qd
from quadrants_b200 import distributed as qdist
from tests import workloads as W


def main():
    nv = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
    nt = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 200
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    qd.init(arch=qd.cuda, device=local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    r, c, v, n = W.fem_tet_triplets(nv, nt, seed=7, dtype=np.float64, compact=True)
    Ag = sp.coo_matrix((v, (r, c)), shape=(n, n)).tocsr()  # host copy: partition + checks (duplicates summed)
    Ag.sort_indices()
    ro = qdist.nnz_balanced_row_offsets(Ag.indptr, world) if world > 1 else np.array([0, n], np.int64)
    lo, hi = int(ro[rank]), int(ro[rank + 1])
    mine = (r >= lo) & (r < hi)
    rl, cl, vl = r[mine] - lo, c[mine], v[mine]
    K = qd.linalg.SparseMatrixBuilder(hi - lo, n, max_num_triplets=max(int(rl.shape[0]), 1), dtype=qd.f64)
    K.ptr.insert_triplets(torch.from_numpy(rl).to(dev), torch.from_numpy(cl).to(dev), torch.from_numpy(vl).to(dev))
    A = K.build()
    nnz_local = A.matrix.nnz()
    ctx = qdist.RowBlockContext(rank, world).localize(A, ro)

    # ---- parity: SpMV of the rank's rows --------------------------------------------------------------------------------
    x = np.random.default_rng(8).uniform(-1, 1, n)
    xl, yl = qd.ndarray(qd.f64, hi - lo), qd.ndarray(qd.f64, hi - lo)
    xl.from_numpy(x[lo:hi])
    for _ in range(2):
        ctx.spmv(xl, yl)
    torch.cuda.synchronize()
    Al = Ag[lo:hi]
    y_ref = Al @ x
    bound = 1e-12 * (abs(Al) @ np.abs(x)) + 1e-300
    spmv_ok = bool(np.all(np.abs(yl.to_numpy() - y_ref) <= bound))

    # ---- parity: a converged solve ----------------------------------------------------------------------------------------
    b = np.random.default_rng(9).uniform(-1, 1, n)
    bl, xs = qd.ndarray(qd.f64, hi - lo), qd.ndarray(qd.f64, hi - lo)
    bl.from_numpy(b[lo:hi])
    solver = ctx.cg(5000, 1e-10)
    solver.solve(xs, bl)
    torch.cuda.synchronize()
    its = torch.tensor([solver.iterations], device=dev)
    if world > 1:
        allx = [torch.zeros(int(ro[g + 1] - ro[g]), dtype=torch.float64, device=dev) for g in range(world)]
        dist.all_gather(allx, xs.tensor)
        xg = torch.cat(allx).cpu().numpy()
        allits = [torch.zeros_like(its) for _ in range(world)]
        dist.all_gather(allits, its)
        same_iters = len({int(t.item()) for t in allits}) == 1
    else:
        xg, same_iters = xs.to_numpy(), True
    true_res = float(np.linalg.norm(b - Ag @ xg) / np.linalg.norm(b))

    # ---- timing ----------------------------------------------------------------------------------------------------------
    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(3):
        solver.run_fixed(xs, bl, iters)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        solver.run_fixed(xs, bl, iters)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    stats = torch.tensor([float(nnz_local), float(ctx.n_ghost)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        allstats = [torch.zeros_like(stats) for _ in range(world)]
        dist.all_gather(allstats, stats)
    else:
        allstats = [stats]
    if rank == 0:
        nnz = int(Ag.nnz)
        ms_it = float(ms.item()) / iters
        bytes_it = nnz * 12 + (n + 1) * 4 + 13 * n * 8  # SpMV + 11 n vector passes, fp64
        lens = np.diff(Ag.indptr)
        print(json.dumps({
            "config": "C5 FEM tet stiffness fp64 Jacobi-PCG", "n_gpus": world, "rows": int(n), "nnz": nnz,
            "row_len_mean": round(float(lens.mean()), 1), "row_len_max": int(lens.max()),
            "iters_per_s": round(1e3 / ms_it, 1), "us_per_iter": round(1e3 * ms_it, 2),
            "algorithmic_gbs_whole_job": round(bytes_it / ms_it / 1e6, 1),
            "nnz_per_rank": [int(s[0].item()) for s in allstats], "ghosts_per_rank": [int(s[1].item()) for s in allstats],
            "parity": {"spmv_within_1e-12": spmv_ok, "solve_iterations": int(solver.iterations),
                       "same_iterations_on_all_ranks": bool(same_iters), "solve_success": bool(solver.success),
                       "true_relative_residual": true_res}}), flush=True)
    if not spmv_ok:
        print(f"rank {rank}: distributed SpMV outside the fp64 bar", file=sys.stderr)
    del solver, ctx
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0 if spmv_ok else 1


if __name__ == "__main__":
    sys.exit(main())
