#!/usr/bin/env python
"""Secondary throughput harness (NOT the driver's bench line): the other BASELINE.json configs on one B200.

    python tools/bench_paths.py [c1] [scan] [sort] [build] [fem] [algos] [solver] [algebra]       # default: all, one JSON line per config

C1  2-D Poisson 512 x 512 fp32: build, SpMV, SparseCG to 1e-6 (the reference's CPU-runnable case)
C3  inclusive scan and key-value sort of 2^28 random i32                     (byte models of SURVEY 8(d))
C4  triplet -> CSR build of 200 M random f32 triplets, ~30 % duplicates, rows = cols = 2^24
C5  fp64 SpMV + Jacobi-PCG on an irregular FEM-like matrix (power-law row lengths)
All through the C ABI (ctypes), CUDA events on the launching stream, inputs resident in HBM and far larger than L2."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import quadrants_b200 as qd
from quadrants_b200 import _lib

PEAK = 6551.7
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except (OSError, KeyError, ValueError):
    pass


def timed(fn, reps, setup=None):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    total = 0.0
    for _ in range(reps):
        if setup:
            setup()
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        total += e0.elapsed_time(e1)
    return total / reps


def emit(name, ms, nbytes, **extra):
    gbs = nbytes / (ms * 1e-3) / 1e9
    print(json.dumps({"config": name, "ms": round(ms, 4), "algorithmic_bytes": int(nbytes), "gbs": round(gbs, 1),
                      "frac_of_measured_peak": round(gbs / PEAK, 4), "frac_of_nominal_8000": round(gbs / 8000, 4),
                      **extra}), flush=True)


def bench_c1(lib, stream):
    """C1 (BASELINE configs[0], the reference's own CPU-runnable case): 2-D 5-point Poisson 512 x 512, build + SpMV +
    SparseCG to 1e-6, fp32.  Everything fits in L2 (16 MB), so this line is launch/latency-bound, not HBM-bound."""
    from tests import workloads as W
    r, c, v = W.poisson2d_triplets(512)
    n = 512 * 512
    rt, ct, vt = torch.from_numpy(r.astype(np.int32)).cuda(), torch.from_numpy(c.astype(np.int32)).cuda(), torch.from_numpy(v).cuda()
    K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=int(r.shape[0]), dtype=qd.f32)
    out = []
    def setup():
        K.ptr.buffer_tensor()[0] = 0
        K.ptr._appended = 0
        K.ptr.insert_triplets(rt, ct, vt)
        out.clear()
    def run():
        out.append(K.build())
    setup(); run()
    ms_build = timed(run, 5, setup=setup)
    A = out[0]
    nnz = A.matrix.nnz()
    x = qd.ndarray(qd.f32, n); x.tensor.uniform_(-1, 1)
    y = qd.ndarray(qd.f32, n)
    spmv = lambda: A.matrix.spmv(None, x, y)
    for _ in range(5):
        spmv()
    reps = 500
    ms = timed(lambda: [spmv() for _ in range(reps)], 3) / reps
    emit("C1 2-D Poisson 512x512 fp32 SpMV (L2-resident)", ms, nnz * 8 + (n + 1) * 4 + 2 * n * 4, rows=n, nnz=int(nnz),
         build_ms=round(ms_build, 3))
    b = qd.ndarray(qd.f32, n); b.fill(1.0)
    cg = qd.linalg.SparseCG(A, b, max_iter=5000, atol=1e-6)
    cg.solve(); torch.cuda.synchronize()
    ms = timed(lambda: cg.solve(), 5)
    emit("C1 2-D Poisson 512x512 fp32 Jacobi-PCG to 1e-6", ms, cg.iterations * (nnz * 8 + (n + 1) * 4 + 13 * n * 4),
         iterations=cg.iterations, rel_residual=cg.error, iters_per_s=round(cg.iterations / ms * 1e3, 1))


def bench_scan(lib, stream):
    n = 1 << 28
    g = torch.Generator(device="cuda").manual_seed(5)
    src = torch.randint(-2**31, 2**31, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
    a = torch.empty_like(src)
    tmp = torch.empty(lib.qdsp_scan_tmp_bytes(n), dtype=torch.uint8, device="cuda")
    run = lambda: _lib.check(lib.qdsp_inclusive_scan_i32(a.data_ptr(), n, tmp.data_ptr(), tmp.numel(), stream))
    for _ in range(3):
        a.copy_(src); run()
    ms = timed(run, 10, setup=lambda: a.copy_(src))
    emit("C3 inclusive scan 2^28 i32", ms, 2 * n * 4, gkeys_per_s=round(n / ms / 1e6, 2))


def bench_sort(lib, stream):
    n = 1 << 28
    g = torch.Generator(device="cuda").manual_seed(4)
    src = torch.randint(-2**31, 2**31, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
    k = torch.empty_like(src)
    v = torch.empty_like(src)
    ar = torch.arange(n, dtype=torch.int32, device="cuda")
    tmp = torch.empty(lib.qdsp_sort_tmp_bytes(n, _lib.KEY_I32, 4), dtype=torch.uint8, device="cuda")
    def setup():
        k.copy_(src); v.copy_(ar)
    run = lambda: _lib.check(lib.qdsp_sort_pairs(k.data_ptr(), v.data_ptr(), n, _lib.KEY_I32, 4, tmp.data_ptr(), tmp.numel(), stream))
    for _ in range(2):
        setup(); run()
    ms = timed(run, 5, setup=setup)
    emit("C3 parallel_sort 2^28 i32 keys + i32 values", ms, n * 4 + 4 * 2 * n * 8, two_touch_bytes=2 * n * 8,
         gpairs_per_s=round(n / ms / 1e6, 2))
    tmpk = torch.empty(lib.qdsp_sort_tmp_bytes(n, _lib.KEY_I32, 0), dtype=torch.uint8, device="cuda")
    runk = lambda: _lib.check(lib.qdsp_sort_pairs(k.data_ptr(), None, n, _lib.KEY_I32, 0, tmpk.data_ptr(), tmpk.numel(), stream))
    setup(); runk()
    ms = timed(runk, 5, setup=setup)
    emit("C3 parallel_sort 2^28 i32 keys only", ms, n * 4 + 4 * 2 * n * 4, gkeys_per_s=round(n / ms / 1e6, 2))


def bench_build(lib, stream):
    rows = cols = 1 << 24
    nu, nd = 140_000_000, 60_000_000
    g = torch.Generator(device="cuda").manual_seed(6)
    r = torch.randint(0, rows, (nu,), dtype=torch.int32, device="cuda", generator=g)
    c = torch.randint(0, cols, (nu,), dtype=torch.int32, device="cuda", generator=g)
    pick = torch.randint(0, nu, (nd,), dtype=torch.int64, device="cuda", generator=g)
    r = torch.cat([r, r[pick]]); c = torch.cat([c, c[pick]])
    perm = torch.randperm(nu + nd, device="cuda", generator=g)
    r, c = r[perm].contiguous(), c[perm].contiguous()
    del perm, pick
    v = torch.empty(nu + nd, dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g)
    t = nu + nd
    K = qd.linalg.SparseMatrixBuilder(rows, cols, max_num_triplets=t, dtype=qd.f32)
    K.ptr.insert_triplets(r, c, v)
    buf = K.ptr.buffer_tensor().clone()      # the filled buffer, restored before every timed build
    del r, c, v
    out = []
    def setup():
        K.ptr.buffer_tensor().copy_(buf)
        out.clear()
    def run():
        out.append(K.build())
    setup(); run(); nnz = out[0].matrix.nnz()
    ms = timed(run, 3, setup=setup)
    s, kb = 4, 8
    P = 6
    model = 3 * t * s + t * (kb + s) + t * kb + P * 2 * t * (kb + s) + t * (kb + s) + nnz * (4 + s) + (rows + 1) * 4
    emit("C4 triplet->CSR build, 200 M f32 triplets (~30 % duplicates), 2^24 x 2^24", ms, model, triplets=t, nnz=int(nnz),
         mtriplets_per_s=round(t / ms / 1e3, 1))


def bench_fem(lib, stream):
    from tests import workloads as W
    nv, nt = 400_000, 2_000_000
    r, c, v, n = W.fem_tet_triplets(nv, nt, seed=7, dtype=np.float64, compact=True)
    K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=int(r.shape[0]), dtype=qd.f64)
    K.ptr.insert_triplets(torch.from_numpy(r).cuda(), torch.from_numpy(c).cuda(), torch.from_numpy(v).cuda())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); A = K.build(); e1.record(); torch.cuda.synchronize()
    ms_build = e0.elapsed_time(e1)
    nnz = A.matrix.nnz()
    ip = A.matrix.to_scipy_arrays()[0]
    lens = np.diff(ip)
    x = qd.ndarray(qd.f64, n); x.tensor.uniform_(-1, 1)
    y = qd.ndarray(qd.f64, n)
    run = lambda: A.matrix.spmv(None, x, y)
    for _ in range(5):
        run()
    reps = 200
    ms = timed(lambda: [run() for _ in range(reps)], 3) / reps
    emit("C5 FEM tet stiffness fp64 SpMV (L2-resident: matrix %.0f MB)" % (nnz * 12 / 1e6), ms, nnz * 12 + (n + 1) * 4 + 2 * n * 8,
         rows=n, nnz=int(nnz), row_len_min=int(lens.min()), row_len_mean=round(float(lens.mean()), 1), row_len_max=int(lens.max()),
         triplets=int(r.shape[0]), build_ms=round(ms_build, 3))
    b = qd.ndarray(qd.f64, n); b.tensor.uniform_(-1, 1)
    cg = qd.linalg.SparseCG(A, b, max_iter=5000, atol=1e-10)
    xs, ok = cg.solve()
    torch.cuda.synchronize()
    ms = timed(lambda: cg.solve(), 3)
    emit("C5 FEM fp64 Jacobi-PCG to 1e-10", ms, cg.iterations * (nnz * 12 + (n + 1) * 4 + 12 * n * 8), iterations=cg.iterations,
         success=bool(ok), rel_residual=cg.error, iters_per_s=round(cg.iterations / ms * 1e3, 1))


def bench_algos(lib, stream):
    """the modern device-wide ops at N = 2^26 (256 MB per 4-byte array, > L2) against the byte models
    of DESIGN.md 3.7/3.8."""
    A = qd.algorithms
    n = 1 << 26
    D = 4
    g = torch.Generator(device="cuda").manual_seed(9)
    cnt = qd.ndarray(qd.i32, 1); cnt.fill(n)
    for dt, sb in ((qd.i32, 4), (qd.f32, 4), (qd.f64, 8)):
        a, o, o1 = qd.ndarray(dt, n), qd.ndarray(dt, n), qd.ndarray(dt, 1)
        if dt is qd.i32:
            a.tensor.copy_(torch.randint(-1000, 1000, (n,), dtype=torch.int32, device="cuda", generator=g))
        else:
            a.tensor.uniform_(-1, 1, generator=g)
        sc = qd.ndarray(qd.u32 if sb == 4 else qd.u64, A.exclusive_scan_scratch_slots(n, D))
        run = lambda: A.reduce_add(a, o1, sc, cnt, dt, D)
        run(); run()
        emit(f"N1 reduce_add 2^26 {dt}", timed(run, 10), n * sb)
        run = lambda: A.exclusive_scan_add(a, o, sc, cnt, dt, D)
        run(); run()
        emit(f"N1 exclusive_scan_add 2^26 {dt} (3-pass, fixed order)", timed(run, 10), 3 * n * sb, two_touch_bytes=2 * n * sb)
        del a, o, o1, sc
    a, o, fl, num = qd.ndarray(qd.i32, n), qd.ndarray(qd.i32, n), qd.ndarray(qd.i32, n), qd.ndarray(qd.i32, 1)
    a.tensor.copy_(torch.arange(n, dtype=torch.int32, device="cuda"))
    fl.tensor.copy_((torch.rand(n, device="cuda", generator=g) < 0.5).to(torch.int32))
    sc = qd.ndarray(qd.u32, A.select_scratch_slots(n))
    run = lambda: A.select(a, fl, o, num, sc, cnt, D)
    run(); run()
    k = int(num.to_numpy()[0])
    emit("N1 select 2^26 i32, 50 % selected", timed(run, 10), n * 4 + 2 * n * 4 + k * 4, selected=k)
    keys = qd.ndarray(qd.i32, n)
    keys.tensor.copy_((torch.arange(n, device="cuda") // 8).to(torch.int32))
    ko, vo = qd.ndarray(qd.i32, n), qd.ndarray(qd.f32, n)
    vals = qd.ndarray(qd.f32, n); vals.tensor.uniform_(-1, 1, generator=g)
    run = lambda: A.reduce_by_key_add(keys, vals, ko, vo, num, sc, cnt, qd.f32, D)
    run(); run()
    r = int(num.to_numpy()[0])
    emit("N1 reduce_by_key_add 2^26 i32 keys / f32 values, runs of 8", timed(run, 10), 2 * n * 4 + n * 4 + r * 8, runs=r)
    del fl, ko, vo, vals
    src = torch.randint(-2**31, 2**31, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
    tk, v, tv = qd.ndarray(qd.i32, n), qd.ndarray(qd.i32, n), qd.ndarray(qd.i32, n)
    ssc = qd.ndarray(qd.u32, A.sort_scratch_slots(n, D))
    n0 = qd.ndarray(qd.i32, ()); n0.fill(n)
    def setup():
        keys.tensor.copy_(src); v.tensor.copy_(a.tensor)
    run = lambda: A.sort(keys, tk, v, tv, ssc, n0, qd.i32, True, 32, D)
    setup(); run()
    emit("N1 sort 2^26 i32 keys + i32 values (device count)", timed(run, 5, setup=setup), n * 4 + 4 * 2 * n * 8,
         gpairs_per_s=round(n / timed(run, 3, setup=setup) / 1e6, 2))
    del tk, v, tv, ssc, keys, src
    t = 60_000_000
    K = qd.linalg.SparseMatrixBuilder(1 << 24, 1 << 24, max_num_triplets=t, dtype=qd.f32)
    r_ = torch.randint(0, 1 << 24, (t,), dtype=torch.int32, device="cuda", generator=g)
    c_ = torch.randint(0, 1 << 24, (t,), dtype=torch.int32, device="cuda", generator=g)
    v_ = torch.empty(t, dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g)
    def reset():
        K.ptr.buffer_tensor()[0] = 0
        K.ptr._appended = 0
    run = lambda: K.ptr.insert_triplets(r_, c_, v_)
    reset(); run()
    emit("N2 bulk append of 60 M f32 triplets (no atomics)", timed(run, 5, setup=reset), t * 12 + t * 12,
         mtriplets_per_s=round(t / timed(run, 3, setup=reset) / 1e3, 1))


def emit_plain(name, **kv):
    print(json.dumps({"config": name, **kv}), flush=True)


def bench_solver(lib, stream):
    """A16: qd.linalg.SparseSolver (device LLT / LU) on 2-D Poisson matrices, fp64; the reference's Eigen SimplicialLLT
    (AMD) on the host cores beside it when oracle/_ref is on the box.  Latency-bound (dependency chain), not HBM-bound:
    reported as times, factor size and multiply-add rate."""
    import time
    from tests import workloads as W
    import oracle
    ref = oracle.EigenRef() if oracle.EigenRef.available() else None
    for nx in (128, 256, 512):
        n = nx * nx
        r, c, v = W.poisson2d_triplets(nx, dtype=np.float64)
        K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=int(r.shape[0]), dtype=qd.f64)
        K.insert_triplets(r.astype(np.int32), c.astype(np.int32), v)
        A = K.build()
        b = qd.ndarray(qd.f64, n); b.tensor.uniform_(-1, 1)
        for st in ("LLT", "LU"):
            s = qd.linalg.SparseSolver(dtype=qd.f64, solver_type=st)
            t0 = time.perf_counter(); s.analyze_pattern(A); t_an = (time.perf_counter() - t0) * 1e3
            s.factorize(A)
            ms_f = timed(lambda: s.factorize(A), 3)
            x = s.solve(b)
            ms_s = timed(lambda: s.solve(b), 5)
            res = (A @ x).tensor - b.tensor
            stt = s.solver.stats()
            extra = {}
            if ref is not None and st == "LLT":
                m = ref.matrix(oracle.make_triplet_buffer(r, c, v, np.float64), n, n, row_major=False)
                t0 = time.perf_counter(); m.direct_solve(b.to_numpy(), "LLT", "AMD"); extra["eigen_host_total_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
            emit_plain(f"A16 SparseSolver {st} fp64, 2-D Poisson {nx}x{nx}", rows=n, factor_nnz=stt["factor_nnz"],
                       etree_height=stt["etree_height"], analyze_host_ms=round(t_an, 2), factorize_ms=round(ms_f, 3),
                       solve_ms=round(ms_s, 3), gmadd_per_s=round(stt["flops"] * (2 if st == "LU" else 1) / 2 / ms_f / 1e6, 2),
                       rel_residual=float(res.norm() / b.tensor.norm()), **extra)


def bench_algebra(lib, stream):
    """N3: A + B, A^T, A B on the device (expand -> onesweep sort -> collapse), fp32, against the bytes each must move."""
    n, nnz = 4_000_000, 40_000_000
    g = torch.Generator(device="cuda").manual_seed(1)
    def mk(seed):
        g.manual_seed(seed)
        r = torch.randint(0, n, (nnz,), device="cuda", generator=g, dtype=torch.int32)
        c = torch.randint(0, n, (nnz,), device="cuda", generator=g, dtype=torch.int32)
        v = torch.rand(nnz, device="cuda", generator=g) * 2 - 1
        K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=nnz, dtype=qd.f32)
        K.insert_triplets(r, c, v)
        return K.build()
    A, B = mk(1), mk(2)
    na, nb = A.matrix.nnz(), B.matrix.nnz()
    out = []
    for name, fn, model in (("A + B", lambda: out.append(A + B), (na + nb) * 8 * 2 + 2 * (n + 1) * 4),
                            ("A^T", lambda: out.append(A.transpose()), na * 8 * 2 + 2 * (n + 1) * 4)):
        fn(); out.clear()
        ms = timed(fn, 3, setup=out.clear)
        emit(f"N3 {name}, {n} x {n}, {na} nnz fp32 (model: read operands + write result once)", ms, model,
             result_nnz=int(out[0].matrix.nnz()))
        out.clear()
    del B
    torch.cuda.empty_cache()
    m2 = 1_000_000
    g.manual_seed(3)
    def mk2(seed, k):
        g.manual_seed(seed)
        r = torch.randint(0, m2, (k,), device="cuda", generator=g, dtype=torch.int32)
        c = torch.randint(0, m2, (k,), device="cuda", generator=g, dtype=torch.int32)
        v = torch.rand(k, device="cuda", generator=g) * 2 - 1
        K = qd.linalg.SparseMatrixBuilder(m2, m2, max_num_triplets=k, dtype=qd.f32)
        K.insert_triplets(r, c, v)
        return K.build()
    P, Q = mk2(4, 8_000_000), mk2(5, 8_000_000)
    fn = lambda: out.append(P @ Q)
    fn(); prod_nnz = int(out[0].matrix.nnz()); out.clear()
    ms = timed(fn, 3, setup=out.clear)
    emit_plain(f"N3 A B, {m2} x {m2}, 8 M nnz each fp32 (~64 M partial products)", ms=round(ms, 3), result_nnz=prod_nnz,
               mproducts_per_s=round(64.0 / ms * 1e3, 1))


def main():
    qd.init(arch=qd.cuda, device=0)
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    which = sys.argv[1:] or ["c1", "scan", "sort", "build", "fem", "algos", "solver", "algebra"]
    for w in which:
        {"c1": bench_c1, "scan": bench_scan, "sort": bench_sort, "build": bench_build, "fem": bench_fem, "algos": bench_algos, "solver": bench_solver,
         "algebra": bench_algebra}[w](lib, stream)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
