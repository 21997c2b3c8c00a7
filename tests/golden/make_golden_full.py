"""Full-size golden digests of the BASELINE.json configs C2, C4 and C5, produced by the reference's own Eigen
(oracle/_ref/libeigen_ref.so = /root/reference/external/eigen compiled by oracle/Makefile with the reference's flags,
performing the calls of quadrants/program/sparse_matrix.cpp:139-152,254-279 and conjugate_gradient.h:36-49).

Run in a container that has /root/reference (about 20 minutes of CPU, < 30 GB of RAM):

    python tests/golden/make_golden_full.py [c2] [c4] [c4d] [c5]

Writes tests/golden/full_<cfg>.json (sha256 digests of the compressed arrays, iteration counts, norms) and
tests/golden/full_<cfg>.npz (strided samples of y / x).  The -m gpu tests of tests/test_gpu_full_size.py regenerate
the same inputs from the same seeds (tests/workloads.py), run the CUDA path and compare: digests bit-exact, samples at
the tolerance written in the test."""
import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402
from tests import workloads as W  # noqa: E402

STRIDE = 4096
C2_STRIDE = 4099   # prime: a power-of-two stride on the 256^3 grid would sample the k = 0 boundary plane only


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).view(np.uint8).data).hexdigest()


def csr_digests(outer, inner, vals):
    return {"nnz": int(inner.shape[0]), "indptr_sha256": sha(outer), "indices_sha256": sha(inner),
            "values_sha256": sha(vals)}


def save(cfg, meta, arrays):
    with open(os.path.join(HERE, f"full_{cfg}.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    if arrays:
        np.savez_compressed(os.path.join(HERE, f"full_{cfg}.npz"), **arrays)
    print(cfg, json.dumps(meta)[:600])


def c2(ref):
    """3-D 7-point Poisson 256^3 fp32: CSR digests, y = A x (seed 2), CG to 1e-6 (budget 1000) on
    b = A x_true (seed 3) in fp32, plus the same solve in fp64 as the yardstick."""
    t0 = time.time()
    n, N = 256, 256 ** 3
    r, c, v = W.poisson3d_triplets(n, np.float32)
    buf = O.make_triplet_buffer(r, c, v, np.float32)
    del r, c, v
    m = ref.matrix(buf, N, N, True)
    outer, inner, vals = m.export()
    meta = {"config": "C2 poisson3d 256^3 f32", "rows": N, "eigen": ref.version(), **csr_digests(outer, inner, vals)}
    print("c2 build", time.time() - t0)
    x = W.c2_spmv_x()
    y = m.spmv(x)
    meta["y_rowmajor_sha256"] = sha(y)
    del m
    mc = ref.matrix_from_compressed(outer, inner, vals, N, N, False)   # symmetric: CSC arrays == CSR arrays
    yc = mc.spmv(x)
    meta["y_colmajor_sha256"] = sha(yc)
    arrays = {"y_rowmajor": y[::C2_STRIDE].copy(), "y_colmajor": yc[::C2_STRIDE].copy()}
    xt = W.c2_cg_xtrue()
    b = mc.spmv(xt)
    meta["b_sha256"] = sha(b)
    t1 = time.time()
    xs, ok, it, err = mc.cg(b, 1000, 1e-6)
    print("c2 cg f32", ok, it, err, time.time() - t1)
    meta["cg_f32"] = {"success": bool(ok), "iterations": int(it), "error": float(err), "tol": 1e-6, "budget": 1000,
                      "x_norm2": float(np.linalg.norm(xs.astype(np.float64))),
                      "rel_err_vs_xtrue": float(np.linalg.norm(xs.astype(np.float64) - xt) / np.linalg.norm(xt.astype(np.float64)))}
    arrays["cg_f32_x"] = xs[::C2_STRIDE].copy()
    del mc
    md = ref.matrix_from_compressed(outer, inner, vals.astype(np.float64), N, N, False)
    t1 = time.time()
    xd, okd, itd, errd = md.cg(b.astype(np.float64), 1000, 1e-6)
    print("c2 cg f64", okd, itd, errd, time.time() - t1)
    meta["cg_f64"] = {"success": bool(okd), "iterations": int(itd), "error": float(errd),
                      "x_norm2": float(np.linalg.norm(xd)),
                      "rel_err_vs_xtrue": float(np.linalg.norm(xd - xt) / np.linalg.norm(xt.astype(np.float64)))}
    meta["cg_f32"]["rel_err_vs_f64"] = float(np.linalg.norm(xs.astype(np.float64) - xd) / np.linalg.norm(xd))
    arrays["cg_f64_x"] = xd[::C2_STRIDE].copy()
    meta["sample_stride"] = C2_STRIDE
    save("c2", meta, arrays)


def c4(ref, dtype=np.float32, nu=140_000_000, nd=60_000_000, tag="c4"):
    """Triplet -> CSR build from 200 M random triplets with 30 % repeats on 2^24 x 2^24:
    Eigen row-major setFromTriplets, digests of the three arrays."""
    t0 = time.time()
    rows = cols = 1 << 24
    r, c, v = W.random_triplets(rows, cols, nu, nd, seed=6, dtype=dtype)
    buf = O.make_triplet_buffer(r, c, v, dtype)
    meta = {"config": f"{tag} random triplets {nu}+{nd} {np.dtype(dtype).name}", "rows": rows, "cols": cols,
            "triplets": int(buf[0]), "buffer_sha256": sha(buf)}
    del r, c, v
    print(tag, "generated", time.time() - t0)
    m = ref.matrix(buf, rows, cols, True)
    outer, inner, vals = m.export()
    meta.update(csr_digests(outer, inner, vals))
    meta["eigen"] = ref.version()
    print(tag, "built", time.time() - t0)
    save(tag, meta, {"values": vals[::STRIDE].copy(), "indices": inner[::STRIDE].copy()})


def c5(ref):
    """FEM tetrahedral stiffness, 2 M tets on 400 k vertices, power-law degree, fp64: CSR digests,
    y = A x samples, Jacobi-PCG to 1e-10 (budget 5000): iteration count and x samples."""
    t0 = time.time()
    r, c, v, n = W.fem_tet_triplets(400_000, 2_000_000, seed=7, dtype=np.float64, compact=True)
    buf = O.make_triplet_buffer(r, c, v, np.float64)
    meta = {"config": "C5 fem tets 2M/400k f64", "rows": int(n), "triplets": int(buf[0]), "buffer_sha256": sha(buf),
            "eigen": ref.version()}
    del r, c, v
    m = ref.matrix(buf, n, n, True)
    outer, inner, vals = m.export()
    meta.update(csr_digests(outer, inner, vals))
    lens = np.diff(outer)
    meta["row_len_min_mean_max"] = [int(lens.min()), float(lens.mean()), int(lens.max())]
    print("c5 built", time.time() - t0)
    x = np.random.default_rng(8).uniform(-1, 1, n)
    y = m.spmv(x)
    meta["y_rowmajor_sha256"] = sha(y)
    del m
    mc = ref.matrix(buf, n, n, False)
    yc = mc.spmv(x)
    meta["y_colmajor_sha256"] = sha(yc)
    b = np.random.default_rng(8).uniform(-1, 1, n)      # b ~ U(-1, 1) seed 8
    t1 = time.time()
    xs, ok, it, err = mc.cg(b, 5000, 1e-10)
    print("c5 cg", ok, it, err, time.time() - t1)
    meta["cg_f64"] = {"success": bool(ok), "iterations": int(it), "error": float(err), "tol": 1e-10, "budget": 5000,
                      "x_norm2": float(np.linalg.norm(xs))}
    meta["sample_stride"] = 64
    save("c5", meta, {"y_rowmajor": y[::64].copy(), "y_colmajor": yc[::64].copy(), "cg_x": xs[::64].copy(),
                      "row_scale": W.row_abs_scale(outer, inner, vals, x)[::64].copy()})


def main():
    ref = O.EigenRef()
    which = sys.argv[1:] or ["c5", "c4d", "c2", "c4"]
    for w in which:
        if w == "c2":
            c2(ref)
        elif w == "c4":
            c4(ref)
        elif w == "c4d":
            c4(ref, np.float64, 35_000_000, 15_000_000, "c4d")
        elif w == "c5":
            c5(ref)


if __name__ == "__main__":
    main()
