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


def build_rows(qdt, np_dtype, r, c, v, lo, hi, ncols):
    """This rank's rows [lo, hi) of the global triplet list, through the builder, with global column ids."""
    sel = (r >= lo) & (r < hi)
    K = qd.linalg.SparseMatrixBuilder(hi - lo, ncols, max_num_triplets=max(1, int(sel.sum())), dtype=qdt)
    K.ptr.insert_triplets(r[sel] - lo, c[sel], v[sel].astype(np_dtype))
    return K.build()


def changing_x_case(rank, world, dev, n):
    """successive stand-alone SpMVs whose x CHANGES, with no host synchronisation between the calls
    (power-iteration style).  Without the consumed-flag rendezvous a fast rank's push k+1 can overwrite ghosts a slow
    neighbour is still reading in call k.  fp32 Poisson, one thread per row => bit-exact against the host replay."""
    N = n ** 3
    ro = qdist.even_row_offsets(N, world, multiple=n * n)
    lo, hi = int(ro[rank]), int(ro[rank + 1])
    rg, cg_, vg = W.poisson3d_triplets(n, dtype=np.float32)
    A = build_rows(qd.f32, np.float32, rg, cg_, vg, lo, hi, N)
    ctx = qdist.RowBlockContext(rank, world).localize(A, ro)
    Ag = sp.coo_matrix((vg.astype(np.float64), (rg, cg_)), shape=(N, N)).tocsr()
    ip, ix, vv = Ag.indptr.astype(np.int32), Ag.indices.astype(np.int32), Ag.data.astype(np.float32)
    o = oracle.Oracle()
    x = np.random.default_rng(12).uniform(-1, 1, N).astype(np.float32)
    xl, yl = qd.ndarray(qd.f32, hi - lo), qd.ndarray(qd.f32, hi - lo)
    xl.from_numpy(x[lo:hi])
    scale = np.float32(1.0 / 12.0)
    reps = 12
    for _ in range(reps):               # no synchronize, no collective: ranks drift freely
        ctx.spmv(xl, yl)
        xl.tensor.copy_(yl.tensor * float(scale))
        if rank == 0:                  # make rank 0 the slow one: extra device work between its calls
            torch.empty(1 << 24, device=dev).normal_()
    torch.cuda.synchronize()
    for _ in range(reps):
        x = (o.spmv(ip, ix, vv, x, N, N, False) * scale).astype(np.float32)
    assert np.array_equal(xl.to_numpy().view(np.uint32), x[lo:hi].view(np.uint32)), f"rank {rank}: changing-x SpMV chain differs"
    del ctx
    dist.barrier()


def one_way_ghost_case(rank, world, dev):
    """A structurally UNSYMMETRIC partition: row i reads column i and column i - K of the previous rank, so rank 0 sends
    and receives nothing back, and the last rank receives but sends nothing (its halo wait compared the
    flags with its OWN push count, which never advanced)."""
    K = 5000
    N = K * world
    ro = np.arange(world + 1, dtype=np.int64) * K
    lo, hi = int(ro[rank]), int(ro[rank + 1])
    i = np.arange(N, dtype=np.int64)
    r = np.concatenate([i, i[K:]])
    c = np.concatenate([i, i[K:] - K])
    v = np.concatenate([np.full(N, 2.0), -1.0 - (i[K:] % 7)]).astype(np.float64)
    A = build_rows(qd.f64, np.float64, r, c, v, lo, hi, N)
    ctx = qdist.RowBlockContext(rank, world).localize(A, ro)
    assert ctx.n_ghost == (0 if rank == 0 else K)
    x = np.random.default_rng(13).uniform(-1, 1, N)
    xl, yl = qd.ndarray(qd.f64, K), qd.ndarray(qd.f64, K)
    for rep in range(3):
        xr = x * (rep + 1)
        xl.from_numpy(xr[lo:hi])
        ctx.spmv(xl, yl)
        want = 2.0 * xr[lo:hi]
        if rank > 0:
            want = want + v[N + lo - K: N + hi - K] * xr[lo - K:hi - K]
        assert np.allclose(yl.to_numpy(), want, rtol=1e-14, atol=0), f"rank {rank}: one-way ghost SpMV rep {rep} differs"
    del ctx
    dist.barrier()


def fem_case(rank, world, dev):
    """C5-shaped: power-law FEM stiffness fp64, nnz-balanced row blocks, irregular (non-contiguous) send lists, the
    stream-path SpMV waiting for the halo.  SpMV within 1e-12 per row of the fp64 host product; Jacobi-PCG to 1e-10:
    identical iteration count on every rank, within 2 % of the oracle's, solution within 1e-7."""
    r, c, v, n = W.fem_tet_triplets(30000, 200000, seed=7, dtype=np.float64, compact=True)
    o = oracle.Oracle()
    buf = oracle.make_triplet_buffer(r, c, v, np.float64)
    ip, ix, vv = o.build_matrix(buf, n, n, True)
    ro = qdist.nnz_balanced_row_offsets(ip, world)
    lo, hi = int(ro[rank]), int(ro[rank + 1])
    A = build_rows(qd.f64, np.float64, r, c, v, lo, hi, n)
    ctx = qdist.RowBlockContext(rank, world).localize(A, ro)
    x = np.random.default_rng(8).uniform(-1, 1, n)
    xl, yl = qd.ndarray(qd.f64, hi - lo), qd.ndarray(qd.f64, hi - lo)
    yref = o.spmv(ip, ix, vv, x, n, n, True)
    scale = W.row_abs_scale(ip, ix, vv, x)
    for rep in range(2):
        xl.from_numpy(x[lo:hi] * (rep + 1))
        ctx.spmv(xl, yl)
        err = np.abs(yl.to_numpy() - (rep + 1) * yref[lo:hi])
        assert np.all(err <= 1e-12 * (rep + 1) * scale[lo:hi] + 1e-300), f"rank {rank}: FEM dist SpMV rep {rep}: {err.max()}"
    b = np.random.default_rng(9).uniform(-1, 1, n)
    bl, xs = qd.ndarray(qd.f64, hi - lo), qd.ndarray(qd.f64, hi - lo)
    bl.from_numpy(b[lo:hi])
    solver = ctx.cg(5000, 1e-10)
    solver.solve(xs, bl)
    torch.cuda.synchronize()
    its = torch.tensor([solver.iterations], device=dev)
    allits = [torch.zeros_like(its) for _ in range(world)]
    dist.all_gather(allits, its)
    assert len({int(t.item()) for t in allits}) == 1, f"ranks disagree on iterations: {allits}"
    co, ci, cv = o.build_matrix(buf, n, n, False)
    xo, oko, ito, _ = o.cg(co, ci, cv, b, 5000, 1e-10)
    assert solver.success and oko and abs(solver.iterations - ito) <= max(3, 0.02 * ito), (solver.iterations, ito)
    rel = np.linalg.norm(xs.to_numpy() - xo[lo:hi]) / np.linalg.norm(xo[lo:hi])
    assert rel < 1e-7, rel
    if rank == 0:
        print(f"fem dist ok: world={world} rows={n} nnz={int(ip[-1])} offsets={ro.tolist()} ghosts(rank0)={ctx.n_ghost} "
              f"iterations={solver.iterations} (oracle {ito}) rel={rel:.2e}")
    del solver, ctx
    dist.barrier()


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
    os.environ["QDSP_ALLREDUCE"] = "p2p"
    changing_x_case(rank, world, dev, n)
    one_way_ghost_case(rank, world, dev)
    fem_case(rank, world, dev)
    if rank == 0:
        print(f"DIST_GPU_OK world={world} n={n}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
