"""Synthetic inputs of SURVEY.md 8(d), generated on the host with numpy and handed byte-identically to the oracle and to
the GPU.  Everything returns plain numpy arrays; `oracle.make_triplet_buffer` turns (rows, cols, vals) into the builder
buffer layout."""
from __future__ import annotations

import numpy as np


def poisson2d_triplets(nx: int, ny: int | None = None, dtype=np.float32, split_diag: bool = False):
    """C1: 5-point Laplacian on an nx*ny grid, row r = i*ny + j; diag 4, off-diag -1, Dirichlet (no wrap).
    Emitted in row order.  split_diag emits the diagonal as two triplets of 2.0 (exercises the duplicate sum)."""
    ny = nx if ny is None else ny
    i, j = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    r = (i * ny + j).ravel()
    rows, cols, vals = [], [], []
    for di, dj in ((-1, 0), (0, -1), (0, 0), (0, 1), (1, 0)):
        ii, jj = i + di, j + dj
        ok = ((ii >= 0) & (ii < nx) & (jj >= 0) & (jj < ny)).ravel()
        c = (ii * ny + jj).ravel()
        if di == 0 and dj == 0 and split_diag:
            for _ in range(2):
                rows.append(r); cols.append(c); vals.append(np.full(r.shape, 2.0))
        else:
            rows.append(r[ok]); cols.append(c[ok])
            vals.append(np.full(int(ok.sum()), 4.0 if (di == 0 and dj == 0) else -1.0))
    rows, cols, vals = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
    order = np.argsort(rows, kind="stable")
    return rows[order].astype(np.int64), cols[order].astype(np.int64), vals[order].astype(dtype)


def poisson3d_triplets(n: int, dtype=np.float32, planes: tuple[int, int] | None = None):
    """C2: 7-point Laplacian on n^3, r = (i*n + j)*n + k; diag 6, off-diag -1.  `planes=(lo, hi)` restricts to the rows of
    z-planes i in [lo, hi) (global numbering kept) -- what one rank of a row-block partition assembles.
    Emitted diagonal-by-diagonal (NOT row-sorted), so the build's sort does real work."""
    lo, hi = (0, n) if planes is None else planes
    i, j, k = np.meshgrid(np.arange(lo, hi), np.arange(n), np.arange(n), indexing="ij")
    r = ((i * n + j) * n + k).ravel().astype(np.int64)
    rows, cols, vals = [], [], []
    for d, (di, dj, dk) in enumerate(((0, 0, 0), (-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1))):
        ii, jj, kk = i + di, j + dj, k + dk
        ok = ((ii >= 0) & (ii < n) & (jj >= 0) & (jj < n) & (kk >= 0) & (kk < n)).ravel()
        c = ((ii * n + jj) * n + kk).ravel().astype(np.int64)
        rows.append(r[ok]); cols.append(c[ok])
        vals.append(np.full(int(ok.sum()), 6.0 if d == 0 else -1.0, dtype=dtype))
    return np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)


def random_triplets(rows: int, cols: int, n_unique: int, n_dup: int, seed: int = 6, dtype=np.float32):
    """C4: n_unique uniformly drawn coordinates plus n_dup repeats of coordinates of the first set, shuffled;
    values ~ U(-1, 1)."""
    rng = np.random.default_rng(seed)
    r = rng.integers(0, rows, n_unique, dtype=np.int64)
    c = rng.integers(0, cols, n_unique, dtype=np.int64)
    pick = rng.integers(0, n_unique, n_dup)
    r = np.concatenate([r, r[pick]])
    c = np.concatenate([c, c[pick]])
    perm = rng.permutation(r.shape[0])
    v = rng.uniform(-1.0, 1.0, r.shape[0]).astype(dtype)
    return r[perm], c[perm], v


def fem_tet_triplets(n_vertices: int, n_tets: int, seed: int = 7, zipf: float = 0.8, dtype=np.float64,
                     compact: bool = False):
    """C5: random points in the unit cube; tets whose 4 distinct vertices are drawn with probability ~ rank^-zipf
    (power-law vertex degree => power-law row lengths); element matrix = P1 Laplacian stiffness from the actual
    coordinates (PSD) + 1e-2 * lumped mass on the diagonal (=> SPD).  16 triplets per tet, lots of duplicates."""
    rng = np.random.default_rng(seed)
    pts = rng.random((n_vertices, 3))
    w = np.arange(1, n_vertices + 1, dtype=np.float64) ** (-zipf)
    cdf = np.cumsum(w / w.sum())
    tets = np.searchsorted(cdf, rng.random((n_tets, 4))).clip(0, n_vertices - 1)
    # repair repeated vertices inside a tet by shifting (keeps the power law, guarantees 4 distinct vertices)
    for _ in range(8):
        s = np.sort(tets, axis=1)
        bad = (s[:, 1:] == s[:, :-1]).any(axis=1)
        if not bad.any():
            break
        tets[bad] = (tets[bad] + rng.integers(1, n_vertices, (int(bad.sum()), 4))) % n_vertices
    s = np.sort(tets, axis=1)
    tets = tets[~(s[:, 1:] == s[:, :-1]).any(axis=1)]
    p = pts[tets]                                  # (T, 4, 3)
    d = p[:, 1:, :] - p[:, :1, :]                  # (T, 3, 3) edge matrix
    vol = np.abs(np.linalg.det(d)) / 6.0
    good = vol > 1e-9
    tets, d, vol = tets[good], d[good], vol[good]
    dinv = np.linalg.inv(d)                        # rows of D^-1 give grad of barycentric 1..3
    g = np.zeros((tets.shape[0], 4, 3))
    g[:, 1:, :] = np.transpose(dinv, (0, 2, 1))
    g[:, 0, :] = -g[:, 1:, :].sum(axis=1)
    ke = np.einsum("tad,tbd->tab", g, g) * vol[:, None, None]
    ke[:, np.arange(4), np.arange(4)] += 1e-2 * vol[:, None] / 4.0
    if compact:
        # drop the vertices no tet touches (their rows would be empty => a singular matrix): renumber the used ones
        used, tets = np.unique(tets, return_inverse=True)
        tets = tets.reshape(-1, 4)
        n_vertices = int(used.shape[0])
    rows = np.repeat(tets, 4, axis=1).ravel()
    cols = np.tile(tets, (1, 4)).ravel()
    return rows.astype(np.int64), cols.astype(np.int64), ke.reshape(-1).astype(dtype), n_vertices


def spd_dense_10(seed: int = 0, dtype=np.float32):
    """G5 (tests/python/test_sparse_cg.py:13-39): 10x10 A = M M^T, b = 1..10, as triplets of the dense pattern."""
    rng = np.random.default_rng(seed)
    m = rng.random((10, 10))
    a = (m @ m.T).astype(dtype)
    i, j = np.meshgrid(np.arange(10), np.arange(10), indexing="ij")
    return i.ravel(), j.ravel(), a.ravel(), a, np.arange(1, 11).astype(dtype)


def c2_spmv_x(n: int = 256 ** 3):
    """x ~ U(-1, 1), seed 2, fp32."""
    return np.random.default_rng(2).uniform(-1.0, 1.0, n).astype(np.float32)


def c2_cg_xtrue(n: int = 256 ** 3):
    """b = A x_true with x_true ~ U(-1, 1), seed 3, fp32."""
    return np.random.default_rng(3).uniform(-1.0, 1.0, n).astype(np.float32)


def row_abs_scale(indptr, indices, values, x):
    """sum_j |a_ij x_j| per row in fp64: the scale of the per-row SpMV tolerance (BASELINE.json north_star)."""
    prod = np.abs(values.astype(np.float64) * x[indices].astype(np.float64))
    csum = np.concatenate([[0.0], np.cumsum(prod)])
    return csum[indptr[1:]] - csum[indptr[:-1]]


def solver_cases():
    """A16 (SparseSolver) inputs, shared by tests/golden/make_solver_golden.py (reference's Eigen) and the GPU tests.
    Yields (name, n, rows, cols, vals_f64, b_f64, symmetric).  The 10x10 ones are the reference's own test matrix
    (tests/python/test_sparse_linear_solver.py:22-29: A A^T + I, b = 1..n)."""
    for seed in range(3):
        rng = np.random.default_rng(100 + seed)
        m = rng.random((10, 10))
        a = m @ m.T + np.eye(10)
        i, j = np.meshgrid(np.arange(10), np.arange(10), indexing="ij")
        yield f"dense10_s{seed}", 10, i.ravel(), j.ravel(), a.ravel(), np.arange(1, 11, dtype=np.float64), True
    r, c, v = poisson2d_triplets(24, dtype=np.float64)
    n = 24 * 24
    yield "poisson2d_24", n, r, c, v.astype(np.float64), 1.0 + np.arange(n) / n, True
    r, c, v, nv = fem_tet_triplets(300, 1500, seed=11, compact=True)
    # duplicates are summed by the builder; the lower and upper triangles receive the same sums in the same order
    yield "fem_300", nv, r, c, v, np.cos(np.arange(nv) * 0.37), True
    rng = np.random.default_rng(42)
    n = 120
    r = rng.integers(0, n, 900)
    c = rng.integers(0, n, 900)
    keep = r != c
    r, c = r[keep], c[keep]
    v = rng.uniform(-1.0, 1.0, r.shape[0])
    rowsum = np.zeros(n)
    np.add.at(rowsum, r, np.abs(v))
    colsum = np.zeros(n)
    np.add.at(colsum, c, np.abs(v))
    d = 1.0 + np.maximum(rowsum, colsum)  # strictly diagonally dominant by rows and columns: the row matching of the LU leaves it alone
    yield ("unsym_dd_120", n, np.concatenate([r, np.arange(n)]), np.concatenate([c, np.arange(n)]),
           np.concatenate([v, d]), np.sin(np.arange(n) * 0.11) + 2.0, False)
