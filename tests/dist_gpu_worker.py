"""Multi-GPU parity worker: run under torchrun, one rank per GPU (tests/test_dist_gpu.py launches it).

Each rank assembles ITS row block of the 3-D Poisson matrix through the builder (global column ids), localizes it, and
checks against a host-side scipy/oracle computation of the whole problem:
  * distributed SpMV == the rank's slice of A x          (bit-exact: same per-row fma order as one GPU)
  * distributed Jacobi-PCG: same iteration count on every rank, solution within 1e-5 of the fp64 oracle solution
    (fp64) / no worse than the single-rank fp32 run (fp32), true residual at the tolerance
  * the same with QDSP_ALLREDUCE=nccl
"""
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
import quadrants_b200 as qd  # noqa: E402
from quadrants_b200 import distributed as qdist  # noqa: E402
from tests import workloads as W  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    qd.init(arch=qd.cuda, device=local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    N = n ** 3
    ro = qdist.even_row_offsets(N, world, multiple=n * n)
    lo, hi = int(ro[rank]), int(ro[rank + 1])
    for np_dtype, qdt in ((np.float32, qd.f32), (np.float64, qd.f64)):
        r, c, v = W.poisson3d_triplets(n, dtype=np_dtype, planes=(lo // (n * n), hi // (n * n)))
        K = qd.linalg.SparseMatrixBuilder(hi - lo, N, max_num_triplets=r.shape[0], dtype=qdt)
        K.ptr.insert_triplets(r - lo, c, v)
        A = K.build()
        # whole problem on the host (small): reference product and solution
        rg, cg_, vg = W.poisson3d_triplets(n, dtype=np_dtype)
        Ag = sp.coo_matrix((vg.astype(np.float64), (rg, cg_)), shape=(N, N)).tocsr()
        x = np.random.default_rng(2).uniform(-1, 1, N).astype(np_dtype)
        for mode in ("p2p", "nccl"):
            os.environ["QDSP_ALLREDUCE"] = mode
            if mode == "nccl":  # a fresh context: localize renumbers the matrix in place, so rebuild it
                K = qd.linalg.SparseMatrixBuilder(hi - lo, N, max_num_triplets=r.shape[0], dtype=qdt)
                K.ptr.insert_triplets(r - lo, c, v)
                A = K.build()
            ctx = qdist.RowBlockContext(rank, world).localize(A, ro, use_nccl=(mode == "nccl"))
            assert ctx.n_local == hi - lo
            exp_ghost = (0 if rank == 0 else n * n) + (0 if rank == world - 1 else n * n)
            assert ctx.n_ghost == exp_ghost, (ctx.n_ghost, exp_ghost)
            # ---- SpMV, twice (sequence numbers advance) -------------------------------------------------------------
            xl = qd.ndarray(qdt, hi - lo)
            xl.from_numpy(x[lo:hi])
            yl = qd.ndarray(qdt, hi - lo)
            for rep in range(3):
                ctx.spmv(xl, yl)
                torch.cuda.synchronize()
                y = yl.to_numpy()
                # one-GPU oracle order: col-major Eigen product = per-row left-to-right fma
                o = oracle.Oracle()
                ip, ix, vv = Ag.indptr.astype(np.int32), Ag.indices.astype(np.int32), Ag.data.astype(np_dtype)
                yo = o.spmv(ip, ix, vv, x, N, N, False)[lo:hi]
                assert np.array_equal(y.view(np.uint8), yo.view(np.uint8)), f"rank {rank} {mode} spmv rep {rep} differs"
            # ---- CG ------------------------------------------------------------------------------------------------
            b = np.ones(N, np_dtype)
            bl = qd.ndarray(qdt, hi - lo)
            bl.from_numpy(b[lo:hi])
            xs = qd.ndarray(qdt, hi - lo)
            tol = 1e-6 if np_dtype == np.float32 else 1e-10
            solver = ctx.cg(2000, tol)
            for rep in range(2):
                solver.solve(xs, bl)
                torch.cuda.synchronize()
                its = torch.tensor([solver.iterations], device=dev)
                allits = [torch.zeros_like(its) for _ in range(world)]
                dist.all_gather(allits, its)
                assert len({int(t.item()) for t in allits}) == 1, f"ranks disagree on iterations: {allits}"
                assert solver.success, (mode, solver.iterations, solver.error)
                xo, oko, ito, _ = o.cg(Ag.indptr.astype(np.int32), Ag.indices.astype(np.int32), Ag.data.astype(np.float64),
                                       b.astype(np.float64), 2000, tol)
                xg = xs.to_numpy().astype(np.float64)
                rel = np.linalg.norm(xg - xo[lo:hi]) / np.linalg.norm(xo[lo:hi])
                assert rel < (1e-4 if np_dtype == np.float32 else 1e-8), (mode, rel)
                assert abs(solver.iterations - ito) <= (0.1 * ito if np_dtype == np.float32 else 2), (solver.iterations, ito)
            # fixed-budget run (the bench path) must not hang or diverge
            solver.run_fixed(xs, bl, 40)
            torch.cuda.synchronize()
            assert np.isfinite(xs.to_numpy()).all()
            del solver, ctx
            dist.barrier()
    if rank == 0:
        print(f"DIST_GPU_OK world={world} n={n}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
