#!/usr/bin/env python
"""bench.py -- the headline benchmark of BASELINE.json: CG iterations/s (+ SpMV GB/s against the HBM roofline) on the
3-D 7-point Poisson matrix 256^3 (16.7 M rows, 117 M non-zeros), fp32, row-block partitioned over 1/2/4/8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [...]                          # the reference's Eigen CPU path on the host cores
    torchrun --nproc-per-node N ... bench.py --gpus N ...           # one rank per GPU

One "step" = one fixed-budget SparseCG solve of `--iters-per-step` iterations (Eigen/Jacobi recurrence, the convergence
exit disabled so that every step does the same work).  value = CG iterations per second over all timed steps,
inputs resident in HBM; e2e = the same through the host-buffer C-ABI call (b H2D, x D2H inside the timed region).
Prints ONE JSON line on rank 0.  The matrix (1.0 GB) is ~8x the 126 MB L2, so no L2 flush is needed between steps.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# ---------------------------------------------------------------------------------------------------------------------
# workload: C2 of SURVEY 8(d)
# ---------------------------------------------------------------------------------------------------------------------
def poisson3d_triplets_device(n, lo, hi, device):
    """7-point Laplacian rows of z-planes [lo, hi) of an n^3 grid, global numbering r = (i*n + j)*n + k; diag 6,
    off-diag -1 (Dirichlet).  Emitted diagonal by diagonal (not row-sorted) so the device build's sort does real work.
    Same matrix as tests/workloads.poisson3d_triplets, generated with torch on the device (plumbing)."""
    import torch

    i = torch.arange(lo, hi, device=device, dtype=torch.int64).view(-1, 1, 1)
    j = torch.arange(n, device=device, dtype=torch.int64).view(1, -1, 1)
    k = torch.arange(n, device=device, dtype=torch.int64).view(1, 1, -1)
    r = ((i * n + j) * n + k).reshape(-1)
    rows, cols, vals = [], [], []
    for d, (di, dj, dk) in enumerate(((0, 0, 0), (-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1))):
        ii, jj, kk = i + di, j + dj, k + dk
        ok = ((ii >= 0) & (ii < n) & (jj >= 0) & (jj < n) & (kk >= 0) & (kk < n)).reshape(-1)
        c = ((ii * n + jj) * n + kk).reshape(-1)
        rows.append(r[ok])
        cols.append(c[ok])
        vals.append(torch.full((int(ok.sum()),), 6.0 if d == 0 else -1.0, device=device, dtype=torch.float32))
    return torch.cat(rows), torch.cat(cols), torch.cat(vals)


def stencil_probe(n, lo, hi, device):
    """Parity probe for the assembled operator: x(g) = small integers keyed on the global row id g, and y = A x for the
    z-planes [lo, hi) in closed form from the stencil (6 x - the in-bounds neighbours).  Every partial sum is a small
    integer, so the fp32 product is exact whatever the summation order: the GPU result must be bit-equal."""
    import torch

    i = torch.arange(lo, hi, device=device, dtype=torch.int64).view(-1, 1, 1)
    j = torch.arange(n, device=device, dtype=torch.int64).view(1, -1, 1)
    k = torch.arange(n, device=device, dtype=torch.int64).view(1, 1, -1)

    def xval(ii, jj, kk):
        g = (ii * n + jj) * n + kk
        return ((g * 40503) % 17 - 8).to(torch.float32)

    x = xval(i, j, k)
    y = 6.0 * x
    for di, dj, dk in ((-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1)):
        ii, jj, kk = i + di, j + dj, k + dk
        ok = (ii >= 0) & (ii < n) & (jj >= 0) & (jj < n) & (kk >= 0) & (kk < n)
        y = y - torch.where(ok, xval(ii.clamp(0, n - 1), jj.clamp(0, n - 1), kk.clamp(0, n - 1)), torch.zeros((), device=device))
    return x.reshape(-1).contiguous(), y.reshape(-1).contiguous()


def live_spmv_traffic(n, timeout_s=240, compress=False):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE spmv_kernel launch, captured now, on this box, by running
    `bench.py --spmv-once` under ncu in a child process (after the timed regions; nothing timed runs under the profiler).
    -> (bytes, source string) or (None, why)."""
    import shutil

    ncu = shutil.which("ncu") or "/usr/local/cuda/bin/ncu"
    if not os.path.exists(ncu):
        return None, "ncu not on this box"
    cmd = [ncu, "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum", "--clock-control", "none", "--print-units",
           "base", "-k", "regex:spmv_kernel", "--launch-skip", "2", "--launch-count", "1", "--csv", sys.executable,
           os.path.join(ROOT, "bench.py"), "--spmv-once", "--grid", str(n)]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    env["QDSP_SPMV_COMPRESS"] = "1" if compress else "0"  # canonical 4-byte indices unless asked for the coded stream
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout_s, env=env, cwd=ROOT)
    except (OSError, subprocess.TimeoutExpired) as e:
        return None, f"ncu child failed: {type(e).__name__}"
    total, seen = 0.0, 0
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    import csv

    for row in csv.reader(out.stdout.splitlines()):
        if len(row) < 3:
            continue
        for pos, cell in enumerate(row):
            if cell in ("dram__bytes_read.sum", "dram__bytes_write.sum") and pos + 2 < len(row):
                try:
                    total += float(row[pos + 2].replace(",", "")) * scale.get(row[pos + 1], 1.0)
                    seen += 1
                except ValueError:
                    pass
    if seen != 2:
        return None, f"ncu child rc={out.returncode}, no dram metrics parsed"
    return total, "live: ncu dram__bytes_read.sum + dram__bytes_write.sum of one spmv_kernel launch, child process of this run"


def spmv_once(args):
    """Child of live_spmv_traffic: assemble the workload matrix and launch the SpMV kernel four times."""
    import torch

    import quadrants_b200 as qd

    qd.init(arch=qd.cuda, device=0)
    n = args.grid
    dev = torch.device("cuda", 0)
    r, c, v = poisson3d_triplets_device(n, 0, n, dev)
    K = qd.linalg.SparseMatrixBuilder(n ** 3, n ** 3, max_num_triplets=int(r.shape[0]), dtype=qd.f32)
    K.ptr.insert_triplets(r, c, v)
    A = K.build()
    x = qd.ndarray(qd.f32, n ** 3)
    x.tensor.uniform_(-1.0, 1.0)
    y = qd.ndarray(qd.f32, n ** 3)
    for _ in range(4):
        A.matrix.spmv(None, x, y)
    torch.cuda.synchronize()
    return 0


def spmv_bytes(n, m, nnz, sv):
    """Algorithmic bytes of one SpMV: every array once."""
    return nnz * (sv + 4) + (n + 1) * 4 + m * sv + n * sv


def cg_iter_bytes(n, nnz, sv):
    """Algorithmic bytes of one Jacobi-PCG iteration: SpMV + 11 vector passes."""
    return nnz * (sv + 4) + (n + 1) * 4 + 13 * n * sv


# ---------------------------------------------------------------------------------------------------------------------
# clocks during the timed region
# ---------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ts, line in self.samples:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's Eigen path (oracle/_ref), else the C port (oracle/qd_oracle.c)
# ---------------------------------------------------------------------------------------------------------------------
def poisson3d_csr_host(n):
    """The C2 matrix as CSR on the host (numpy; symmetric, so the same arrays are Eigen's col-major storage)."""
    idx = np.arange(n, dtype=np.int64)
    i, j, k = np.meshgrid(idx, idx, idx, indexing="ij")
    r = ((i * n + j) * n + k).ravel()
    cols = np.full((n ** 3, 7), -1, dtype=np.int64)
    vals = np.zeros((n ** 3, 7), dtype=np.float32)
    # ascending column order inside a row: -n^2, -n, -1, 0, +1, +n, +n^2
    for s, (di, dj, dk) in enumerate(((-1, 0, 0), (0, -1, 0), (0, 0, -1), (0, 0, 0), (0, 0, 1), (0, 1, 0), (1, 0, 0))):
        ii, jj, kk = i + di, j + dj, k + dk
        ok = ((ii >= 0) & (ii < n) & (jj >= 0) & (jj < n) & (kk >= 0) & (kk < n)).ravel()
        cols[ok, s] = (r + (di * n + dj) * n + dk)[ok]
        vals[ok, s] = 6.0 if s == 3 else -1.0
    keep = cols >= 0
    indptr = np.zeros(n ** 3 + 1, dtype=np.int32)
    np.cumsum(keep.sum(axis=1), out=indptr[1:])
    return indptr, cols[keep].astype(np.int32), vals[keep]


class CpuCg:
    """The reference's Eigen Jacobi-PCG on the host cores (oracle/_ref), else the C port (oracle/qd_oracle.c).
    As shipped the reference is single-threaded (no -fopenmp in its cmake) on a col-major matrix, the only storage
    SparseCG accepts (conjugate_gradient.h:40); the best-effort variant lets Eigen thread the row-major product with
    OpenMP on every core.  Both are timed; the faster one is reported."""

    def __init__(self, csr, n_rows):
        import oracle

        self.n = n_rows
        self.b = np.ones(n_rows, np.float32)
        self.cores = os.cpu_count() or 1
        indptr, indices, vals = csr
        self.variants = {}
        if oracle.EigenRef.available():
            self.kind = "reference"
            ref = oracle.EigenRef()
            m = ref.matrix_from_compressed(indptr, indices, vals, n_rows, n_rows, row_major=False)
            self.variants["shipped"] = (1, lambda it, m=m: m.cg(self.b, it, 1e-30)[2])
            if oracle.EigenRef.available(omp=True) and self.cores > 1:
                refo = oracle.EigenRef(omp=True)
                refo.set_threads(self.cores)
                mo = refo.matrix_from_compressed(indptr, indices, vals, n_rows, n_rows, row_major=True)
                self.variants["omp"] = (self.cores, lambda it, mo=mo: mo.cg(self.b, it, 1e-30, allow_row_major=True)[2])
        else:
            self.kind = "port"
            orc = oracle.Oracle()
            self.variants["shipped"] = (1, lambda it: orc.cg(indptr, indices, vals, self.b, it, 1e-30)[2])

    def rate(self, name, iters):
        t0 = time.perf_counter()
        done = self.variants[name][1](iters)
        return done / (time.perf_counter() - t0)

    def baseline(self, iters):
        rates = {k: self.rate(k, iters) for k in self.variants}
        best = max(rates, key=rates.get)
        what = ("Eigen 3.4.90 ConjugateGradient (oracle/_ref, the reference's flags)" if self.kind == "reference"
                else "C restatement of Eigen's Jacobi-PCG (oracle/qd_oracle.c)")
        sample = f"{iters} Jacobi-PCG iterations of {what} on the full workload matrix; " + "; ".join(
            f"{'1 thread col-major as the reference ships' if k == 'shipped' else f'{self.cores} OpenMP threads row-major, best effort'}: {v:.3f} it/s"
            for k, v in rates.items())
        return {"value": rates[best], "unit": "iters/s", "cores": self.variants[best][0], "kind": self.kind,
                "sample": sample, "variant": best}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the best-effort Eigen variant is meant to use every host core
    # (libgomp reads the variable when oracle/_ref is loaded, so set it first)
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    import oracle

    oracle.build()
    n = args.grid
    csr = poisson3d_csr_host(n)
    n_rows = n ** 3
    cpu = CpuCg(csr, n_rows)
    it = args.cpu_iters
    base = cpu.baseline(it)            # untimed warm-up that also picks the faster variant
    name = base["variant"]
    for _ in range(max(0, args.warmup - 1)):
        cpu.rate(name, it)
    t0 = time.perf_counter()
    steps = 0
    for _ in range(args.steps):
        cpu.variants[name][1](it)
        steps += 1
        if time.perf_counter() - t0 > 150:   # bounded: the whole arm must end within a few minutes
            break
    dt = time.perf_counter() - t0
    value = steps * it / dt
    base["value"] = value
    line = {"impl": "reference", "metric": "cg_iters_per_s", "value": value, "unit": "iters/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"3D 7-point Poisson {n}^3 fp32 Jacobi-PCG (C2), full matrix; one step = {it} "
                                   "iterations on the host cores (bounded sample of the GPU arm's step)",
                       "rows": n_rows, "nnz": int(csr[1].shape[0]), "iters_per_step": it},
            "cpu_baseline": base,
            "e2e": {"value": value, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import torch.distributed as dist

    import quadrants_b200 as qd
    from quadrants_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries the ONE JSON line and nothing else: whatever the libraries print (NCCL's version banner ...) goes
    # to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    qd.init(arch=qd.cuda, device=local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    n = args.grid
    n_rows = n ** 3
    sv = 4
    ips = args.iters_per_step

    # ---- assemble through the public API: builder -> device CSR (K2-K4 replacement) ---------------------------------
    if world > 1:
        from quadrants_b200 import distributed as qdist

        ro = qdist.even_row_offsets(n_rows, world, multiple=n * n)  # slabs of whole z-planes
        lo, hi = int(ro[rank]), int(ro[rank + 1])
        r, c, v = poisson3d_triplets_device(n, lo // (n * n), hi // (n * n), dev)
        K = qd.linalg.SparseMatrixBuilder(hi - lo, n_rows, max_num_triplets=int(r.shape[0]), dtype=qd.f32)
        K.ptr.insert_triplets(r - lo, c, v)  # this rank's rows, global column ids
        del r, c, v
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        A = K.build()
        e1.record()
        torch.cuda.synchronize()
        build_ms = e0.elapsed_time(e1)
        del K
        nnz_t = torch.tensor([A.matrix.nnz()], device=dev, dtype=torch.int64)
        dist.all_reduce(nnz_t)
        nnz_total = int(nnz_t.item())
        ctx = qdist.RowBlockContext(rank, world).localize(A, ro, use_nccl=os.environ.get("QDSP_ALLREDUCE") == "nccl")
        bg = qd.ndarray(qd.f32, hi - lo)
        bg.fill(1.0)
        xg = qd.ndarray(qd.f32, hi - lo)
        dcg = ctx.cg(ips, 1e-30)

        cg = argparse.Namespace(cg_solver=dcg)  # same call shape as the single-GPU solver below
    else:
        r, c, v = poisson3d_triplets_device(n, 0, n, dev)
        ntrip = int(r.shape[0])
        K = qd.linalg.SparseMatrixBuilder(n_rows, n_rows, max_num_triplets=ntrip, dtype=qd.f32)
        K.ptr.insert_triplets(r, c, v)
        del r, c, v
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        A = K.build()
        e1.record()
        torch.cuda.synchronize()
        build_ms = e0.elapsed_time(e1)
        del K
        nnz_total = A.matrix.nnz()
        bg = qd.ndarray(qd.f32, n_rows)
        bg.fill(1.0)
        xg = qd.ndarray(qd.f32, n_rows)
        cg = qd.linalg.SparseCG(A, bg, max_iter=ips, atol=1e-30)
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def step_device():
        cg.cg_solver.run_fixed(xg, bg, ips)

    # ---- value: K steps, inputs resident in HBM -----------------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    l0 = _lib.launch_count()
    t_wall0 = time.time()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    barrier()
    t_wall1 = time.time()
    launches = _lib.launch_count() - l0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop(t_wall0, t_wall1)
    ms_per_step = ms_total / args.steps
    value = ips * args.steps / (ms_total * 1e-3)

    n_loc = (hi - lo) if world > 1 else n_rows
    plane_lo, plane_hi = (lo // (n * n), hi // (n * n)) if world > 1 else (0, n)

    def spmv_call(x, y):
        if world > 1:
            ctx.spmv(x, y)
        else:
            A.matrix.spmv(None, x, y)

    def all_ranks(obj):
        if world == 1:
            return [obj]
        box = [None] * world
        dist.all_gather_object(box, obj)
        return box

    # ---- parity gate 1 (before any rate is reported): the assembled (and, for N > 1, partitioned + halo-exchanged)
    # operator against the closed-form stencil, bit for bit, on every rank ---------------------------------------------
    xp_t, yp_t = stencil_probe(n, plane_lo, plane_hi, dev)
    xs = qd.ndarray(qd.f32, n_loc)
    ys = qd.ndarray(qd.f32, n_loc)
    xs.tensor.copy_(xp_t)
    ys.tensor.fill_(float("nan"))
    spmv_call(xs, ys)
    torch.cuda.synchronize()
    spmv_exact = bool(torch.equal(ys.tensor, yp_t))
    del xp_t, yp_t
    if not all(all_ranks(spmv_exact)):
        raise SystemExit(f"PARITY FAILURE rank {rank}: SpMV differs from the closed-form stencil; no rate reported")

    # ---- roofline of the dominant kernel: SpMV (the same CTA code the CG's A.p kernel runs) ---------------------------
    xs.tensor.uniform_(-1.0, 1.0, generator=torch.Generator(device=dev).manual_seed(2 + rank))
    for _ in range(5):
        spmv_call(xs, ys)
    reps = 50
    barrier()
    e0.record(stream)
    for _ in range(reps):
        spmv_call(xs, ys)
    e1.record(stream)
    barrier()
    spmv_ms = e0.elapsed_time(e1) / reps
    nnz_loc = A.matrix.nnz()
    n_cols_loc = n_loc + (ctx.n_ghost if world > 1 else 0)
    alg = spmv_bytes(n_loc, n_cols_loc, nnz_loc, sv)
    # The shipped SpMV streams ONE-BYTE diagonal codes instead of the 4-byte column indices when the matrix allows it
    # (qdsp_csr_spmv_config; the Poisson stencil does): fewer bytes than SURVEY 8(d)'s canonical-CSR model.  The roofline
    # of the line stays the CANONICAL kernel (a copy of the matrix with the codes switched off, timed the same way, N = 1);
    # the coded kernel is reported beside it under roofline.compressed with the bytes it actually has to move.
    info = A.matrix.spmv_info()
    coded = info["index_bytes_per_nnz"] == 1
    moved = alg - 3 * nnz_loc if coded else alg
    compressed = None
    if coded:
        compressed = {"index_bytes_per_nnz": 1, "ms_per_launch": spmv_ms, "bytes_moved_model": moved,
                      "achieved": moved / (spmv_ms * 1e-3) / 1e9,
                      "canonical_equivalent_gbs": alg / (spmv_ms * 1e-3) / 1e9,
                      "note": "same entries, same order, same arithmetic: bit-identical y (parity gate above ran on this kernel)"}
        if world == 1:
            from quadrants_b200 import _core
            ip_t, ix_t, vv_t = A.matrix.csr_tensors()
            h0 = C.c_void_p()
            _lib.check(lib.qdsp_csr_from_device_csr(n_rows, n_rows, nnz_loc, ip_t.data_ptr(), ix_t.data_ptr(), vv_t.data_ptr(),
                                                    _lib.F32, 0, stream.cuda_stream, C.byref(h0)))
            A0 = _core.CuSparseMatrix._wrap(h0)
            A0.spmv_config(0)
            for _ in range(5):
                A0.spmv(None, xs, ys)
            torch.cuda.synchronize()
            e0.record(stream)
            for _ in range(reps):
                A0.spmv(None, xs, ys)
            e1.record(stream)
            torch.cuda.synchronize()
            spmv_ms = e0.elapsed_time(e1) / reps   # the canonical kernel: what the roofline below is about
            del A0, ip_t, ix_t, vv_t
            moved = alg
    per_rank = all_ranks({"rank": rank, "ms_per_launch": spmv_ms, "algorithmic_bytes_per_launch": alg,
                          "bytes_moved_model": moved, "achieved": moved / (spmv_ms * 1e-3) / 1e9})
    spmv_gbs = sum(p["achieved"] for p in per_rank)          # whole job
    per_gpu = min(p["achieved"] for p in per_rank)           # the slowest rank is the one the roofline is about
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    roofline = {"bound": "hbm",
                "kernel": "spmv_kernel<float> (== cg_spmv_dot_kernel body)" if world == 1 else
                          "qdsp_dist_spmv per rank: halo push by peer stores + spmv_kernel<float> + consumed signal",
                "achieved": per_gpu, "peak": peak,
                "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650",
                "unit": "GB/s", "frac": per_gpu / peak, "frac_of_nominal_8000": per_gpu / 8000.0,
                "algorithmic_bytes_per_launch": alg, "ms_per_launch": max(p["ms_per_launch"] for p in per_rank),
                "traffic": None}
    if world > 1:
        roofline["per_rank"] = per_rank
        roofline["note"] = "achieved/frac = the slowest rank's per-GPU figure; spmv_gbs on the line = sum over ranks"
        if coded:
            roofline["note"] += ("; the ranks stream one-byte diagonal codes: achieved = bytes_moved_model / time, "
                                 "canonical-CSR bytes / time is in compressed.canonical_equivalent_gbs")
    if compressed is not None:
        compressed["frac"] = compressed["achieved"] / peak
        compressed["frac_of_nominal_8000"] = compressed["achieved"] / 8000.0
        compressed["traffic"] = None
        roofline["compressed"] = compressed
    del xs, ys

    # ---- e2e: the host-buffer C-ABI call (qdsp_cg_solve_host): b H2D + x D2H inside the timed region; for N > 1 every
    # rank moves its own slice of b and x (pinned), the region is barrier-bracketed and the max over ranks is taken --------
    bh = torch.ones(n_loc, dtype=torch.float32).pin_memory()
    xh = torch.zeros(n_loc, dtype=torch.float32).pin_memory()
    it, err, ok = C.c_int(), C.c_double(), C.c_int()
    h = cg.cg_solver._h

    def step_host():
        _lib.check(lib.qdsp_cg_solve_host(h, xh.data_ptr(), bh.data_ptr(), stream.cuda_stream, C.byref(it),
                                          C.byref(err), C.byref(ok)))

    for _ in range(max(1, args.warmup)):
        step_host()
    barrier()
    e0.record(stream)
    done_iters = 0
    for _ in range(args.steps):
        step_host()
        done_iters += it.value
    e1.record(stream)
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))
    e2e = {"value": done_iters / (e2e_ms * 1e-3), "unit": "iters/s", "h2d_bytes_per_step": n_rows * sv,
           "d2h_bytes_per_step": n_rows * sv, "call": "qdsp_cg_solve_host" + (" on every rank's slice" if world > 1 else ""),
           "iters_per_step": done_iters // args.steps, "final_rel_residual": err.value}

    # ---- parity gate 2: the solve that was just timed.  Every rank must have run the same number of iterations and hold
    # the same residual estimate, and the TRUE residual |b - A x| / |b| recomputed from the returned x must agree with the
    # recurrence's ------------------------------------------------------------------------------------------------------
    seen = all_ranks((done_iters, err.value))
    xd = qd.ndarray(qd.f32, n_loc)
    xd.tensor.copy_(xh, non_blocking=False)
    rd = qd.ndarray(qd.f32, n_loc)
    spmv_call(xd, rd)
    rr = torch.stack([((1.0 - rd.tensor.double()) ** 2).sum(), torch.tensor(float(n_loc), device=dev, dtype=torch.float64)])
    if world > 1:
        dist.all_reduce(rr)
    true_rel = float(torch.sqrt(rr[0] / rr[1]).item())
    cg_ok = (len(set(seen)) == 1 and np.isfinite(true_rel) and abs(true_rel - err.value) <= 2e-2 * err.value + 1e-7)
    if not cg_ok:
        raise SystemExit(f"PARITY FAILURE rank {rank}: per-rank (iterations, error) {seen}; true residual {true_rel} vs "
                         f"reported {err.value}; no rate reported")
    parity = {"spmv_vs_closed_form_stencil": "bit-exact on every rank", "cg_iterations_equal_across_ranks": True,
              "cg_true_rel_residual": true_rel, "cg_reported_rel_residual": err.value}
    del xd, rd

    # ---- CPU baseline on a bounded sample (rank 0, N = 1 only) ----------------------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        import oracle

        oracle.build()
        ip, ix, vv = A.matrix.to_scipy_arrays()
        cpu = CpuCg((ip, ix, vv), n_rows).baseline(args.cpu_iters)
        del ip, ix, vv

    # ---- the kernel to beat: the reference's own cuSPARSE / cuBLAS calls on this GPU, this matrix (N = 1) ---------------
    if world == 1 and not args.no_kernel_to_beat:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_cusparse

            ktb = bench_cusparse.measure(*A.matrix.csr_tensors(), n_rows, n_rows)
            roofline["kernel_to_beat_ms"] = ktb["spmv_csr_alg1_hoisted_ms"]
            roofline["kernel_to_beat"] = ktb
        except Exception as e:  # the bench line must survive a missing library
            roofline["kernel_to_beat_ms"] = None
            roofline["kernel_to_beat"] = {"unavailable": f"{type(e).__name__}: {e}"}

    # ---- DRAM traffic of the dominant kernel, captured now by a child process under ncu (rank 0) ------------------------
    if rank == 0:
        traffic, src = (None, "disabled (--no-live-traffic)") if args.no_live_traffic else live_spmv_traffic(n)
        if traffic is None:
            try:
                rec = json.load(open(os.path.join(ROOT, "profiles", "spmv_traffic.json")))
                traffic, src = rec.get("dram_bytes_per_launch"), f"committed capture profiles/spmv_traffic.json ({src})"
            except (OSError, ValueError):
                pass
        roofline["traffic"], roofline["traffic_source"] = traffic, src
        if world > 1:
            roofline["traffic_note"] = "single-GPU launch over the whole matrix; a rank streams 1/N of it"
        if compressed is not None and not args.no_live_traffic:
            ctraffic, csrc = live_spmv_traffic(n, compress=True)
            compressed["traffic"], compressed["traffic_source"] = ctraffic, csrc

    if rank == 0:
        iter_bytes = cg_iter_bytes(n_rows, nnz_total, sv)
        line = {"metric": "cg_iters_per_s", "value": value, "unit": "iters/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"3D 7-point Poisson {n}^3 fp32 (C2): SparseMatrixBuilder->CSR on device, then "
                                       f"Jacobi-PCG; one step = one {ips}-iteration SparseCG solve",
                           "rows": n_rows, "nnz": int(nnz_total), "iters_per_step": ips,
                           "partition": (f"row-block x{world} (z-slabs), halo by NVLink peer stores, dots all-reduced "
                                         + ("by NCCL" if os.environ.get("QDSP_ALLREDUCE") == "nccl" else
                                            "in-kernel over peer memory")) if world > 1 else "single GPU",
                           "l2": ("no flush: the matrix stream of one iteration (0.59 GB coded / 0.94 GB canonical over all "
                                  "ranks) exceeds the 126 MB L2 of a GPU up to 4 ranks; at 8 ranks a rank's share (73 / 117 MB "
                                  "+ 42 MB of vectors) is L2-sized -- strong scaling of the fixed 256^3 problem"),
                           "csr_build_ms": build_ms},
                "cg_gbs": value * iter_bytes / 1e9, "cg_algorithmic_bytes_per_iter": iter_bytes,
                "spmv_gbs": spmv_gbs, "gpu_launches": int(launches), "clocks": clocks}
        if compressed is not None:
            # the CG's A.p kernel streams the one-byte codes: the bytes an iteration has to move are 3 * nnz below the
            # canonical-CSR model that cg_gbs is quoted on (cg_gbs may therefore exceed the HBM peak)
            line["cg_bytes_moved_per_iter"] = iter_bytes - 3 * int(nnz_total)
            line["cg_gbs_moved"] = value * (iter_bytes - 3 * int(nnz_total)) / 1e9
        line["roofline"] = roofline
        line["e2e"] = e2e
        line["parity_checked"] = True
        line["parity"] = parity
        if cpu:
            line["cpu_baseline"] = cpu
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        # tear the solver and the peer mappings down while every rank is still alive
        torch.cuda.synchronize()
        dcg.close()
        ctx.close()
        dist.barrier()
        torch.cuda.synchronize()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--grid", type=int, default=256, help="n of the n^3 Poisson grid (BASELINE C2: 256)")
    ap.add_argument("--iters-per-step", type=int, default=100)
    ap.add_argument("--cpu-iters", type=int, default=12, help="CG iterations of the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-to-beat", action="store_true", help="skip the cuSPARSE/cuBLAS reference-path timing")
    ap.add_argument("--no-live-traffic", action="store_true", help="skip the ncu child that measures DRAM traffic")
    ap.add_argument("--spmv-once", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    if os.environ.get("QDSP_DEBUG_HANG"):
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["QDSP_DEBUG_HANG"]), exit=True)
    if args.warmup < 3:
        args.warmup = 3
    if args.spmv_once:
        return spmv_once(args)
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
