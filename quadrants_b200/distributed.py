"""Row-block partitioned SpMV / CG over several B200 (one process per GPU), SURVEY.md 8(e).

The reference is single-device (quadrants/rhi/cuda/cuda_context.cpp:16-26), so there is no reference API to mirror here;
the names follow the domain: a *partition* of the rows into contiguous blocks, *ghost* columns owned by other ranks,
the *halo* of x that travels before every product.

torch.distributed is plumbing only (set-up time): it carries the ghost-id lists, the CUDA IPC handles of the
peer-visible buffers and, optionally, the NCCL unique id.  The data path -- halo stores and the all-reduce of the CG
scalars -- runs inside the kernels over NVLink peer memory (csrc/dist_dev.cuh).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def even_row_offsets(n_rows: int, world: int, multiple: int = 1) -> np.ndarray:
    """Contiguous row blocks of (almost) equal size; block boundaries are multiples of `multiple` (e.g. one z-plane)."""
    units = n_rows // multiple
    assert units * multiple == n_rows, "n_rows must be a multiple of `multiple`"
    cuts = [(units * g) // world * multiple for g in range(world + 1)]
    return np.asarray(cuts, dtype=np.int64)


def nnz_balanced_row_offsets(indptr: np.ndarray, world: int) -> np.ndarray:
    """Boundaries chosen by nnz balance (prefix sum of row lengths split at nnz*g/world): for power-law matrices (C5)."""
    indptr = np.asarray(indptr, dtype=np.int64)
    nnz = int(indptr[-1])
    cuts = [int(np.searchsorted(indptr, nnz * g // world, side="left")) for g in range(world + 1)]
    cuts[0], cuts[-1] = 0, indptr.shape[0] - 1
    return np.maximum.accumulate(np.asarray(cuts, dtype=np.int64))


def plan_host(rank: int, world: int, row_offsets, indices_global: np.ndarray, ghost_base: int = -1):
    """qdsp_dist_plan_host: -> (indices_local, ghost_ids, owner_offsets).  Pure host code, no GPU needed.
    Ghost j gets local column id ghost_base + j (default: n_local)."""
    lib = _lib.load()
    ro = np.ascontiguousarray(row_offsets, dtype=np.int64)
    idx = np.ascontiguousarray(indices_global, dtype=np.int32).copy()
    # (a dense owner contributes all of its rows, at most twice the ghosts actually referenced)
    ghosts = np.empty(max(1, min(2 * idx.shape[0], int(ro[-1]))), np.int32)
    n_ghost = C.c_int64()
    owner = np.zeros(world + 1, np.int32)
    _lib.check(lib.qdsp_dist_plan_host(rank, world, ro.ctypes.data, idx.ctypes.data, idx.shape[0], int(ghost_base),
                                       ghosts.ctypes.data,
                                       ghosts.shape[0], C.byref(n_ghost), owner.ctypes.data))
    return idx, ghosts[: n_ghost.value].copy(), owner


def exchange_requests(rank: int, world: int, ghost_base: int, ghost_ids: np.ndarray, owner_offsets: np.ndarray,
                      group=None):
    """Every rank tells every owner which of its rows it needs and where to put them.
    -> list of (peer, ids, dst_offset, dst_offset_z): what THIS rank must send; dst_offset_z is the slot of the same rows
    in the requester's second ghost segment (zbase = ghost_base + n_ghost rounded up to 32 elements).
    Works on any backend (gloo on CPU, nccl on GPU)."""
    import torch.distributed as dist

    mine = {}
    zbase = (ghost_base + int(len(ghost_ids)) + 31) & ~31
    for q in range(world):
        a, b = int(owner_offsets[q]), int(owner_offsets[q + 1])
        if q != rank and b > a:
            mine[q] = (np.ascontiguousarray(ghost_ids[a:b]), ghost_base + a, zbase + a)
    if world == 1:
        return []
    everyone = [None] * world
    dist.all_gather_object(everyone, mine, group=group)
    sends = []
    for s in range(world):
        if s != rank and rank in everyone[s]:
            ids, dst, dst_z = everyone[s][rank]
            sends.append((s, np.ascontiguousarray(ids, dtype=np.int32), int(dst), int(dst_z)))
    return sends


class RowBlockContext:
    """One rank's end of the partition.  Usage (every rank, same order):

        ctx = RowBlockContext(rank, world)
        ctx.localize(A, row_offsets)        # A: this rank's rows with GLOBAL column ids -> local numbering + peers mapped
        ctx.spmv(x_local, y_local)          # halo push + product
        cg = ctx.cg(max_iter, tol);  cg.solve(x_local, b_local)
    """

    def __init__(self, rank: int, world: int, group=None):
        self.rank, self.world, self.group = int(rank), int(world), group
        self._lib = _lib.load()
        self._h = C.c_void_p()
        _lib.check(self._lib.qdsp_dist_create(self.rank, self.world, C.byref(self._h)))
        self.A = None
        self.n_local = self.n_ghost = 0

    def close(self):
        """Unmaps the peers' buffers and frees the communicator.  Call it (after closing the solvers built on this
        context) while every rank is still alive; every rank must do so."""
        if self._h.value:
            self._lib.qdsp_dist_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def localize(self, A, row_offsets, use_nccl: bool = False):
        import torch
        import torch.distributed as dist

        from ._core import _stream

        m = getattr(A, "matrix", A)  # linalg.SparseMatrix or _core.CuSparseMatrix
        ro = np.ascontiguousarray(row_offsets, dtype=np.int64)
        assert ro.shape[0] == self.world + 1
        nl, gb, ng = C.c_int(), C.c_int(), C.c_int()
        owner = np.zeros(self.world + 1, np.int32)
        _lib.check(self._lib.qdsp_dist_localize(self._h, m._need(), ro.ctypes.data, _stream(), C.byref(nl), C.byref(gb),
                                                C.byref(ng), owner.ctypes.data))
        self.n_local, self.ghost_base, self.n_ghost, self.owner_offsets = nl.value, gb.value, ng.value, owner
        m._cols = self.ghost_base + self.n_ghost
        if hasattr(A, "m"):
            A.m = m._cols
        p = C.c_void_p()
        _lib.check(self._lib.qdsp_dist_ghost_ids(self._h, C.byref(p)))
        ghost_ids = (np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_int32)), shape=(self.n_ghost,)).copy()
                     if self.n_ghost else np.empty(0, np.int32))
        self.ghost_ids = ghost_ids
        for peer, ids, dst, dst_z in exchange_requests(self.rank, self.world, self.ghost_base, ghost_ids, owner,
                                                       self.group):
            _lib.check(self._lib.qdsp_dist_set_send_list(self._h, peer, ids.ctypes.data, ids.shape[0], dst, dst_z))
        handle = (C.c_ubyte * 64)()
        _lib.check(self._lib.qdsp_dist_comm_handle(self._h, handle))
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle), group=self.group)
        else:
            handles = [bytes(handle)]
        blob = b"".join(handles)
        _lib.check(self._lib.qdsp_dist_attach(self._h, blob))
        if use_nccl and self.world > 1:
            box = [None]
            if self.rank == 0:
                uid = (C.c_ubyte * 128)()
                _lib.check(self._lib.qdsp_dist_nccl_unique_id(uid))
                box[0] = bytes(uid)
            dist.broadcast_object_list(box, src=0, group=self.group)
            _lib.check(self._lib.qdsp_dist_nccl_init(self._h, box[0]))
        self.A = A
        self._m = m
        if self.world > 1:
            dist.barrier(group=self.group)  # nobody pushes a halo before every peer has mapped every buffer
        torch.cuda.synchronize()
        return self

    def spmv(self, x_local, y_local):
        from ._core import _stream

        _lib.check(self._lib.qdsp_dist_spmv(self._h, self._m._need(), x_local.data_ptr(), y_local.data_ptr(), _stream()))

    def cg(self, max_iter: int, tol: float, verbose: bool = False):
        return DistCG(self, max_iter, tol, verbose)


class DistCG:
    """Jacobi-PCG (the reference's Eigen semantics) over the partition; x and b are the rank's owned slices."""

    def __init__(self, ctx: RowBlockContext, max_iter: int, tol: float, verbose: bool = False):
        self.ctx = ctx
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self.iterations, self.error, self.success = 0, 0.0, False
        _lib.check(self._lib.qdsp_cg_create_dist(ctx._m._need(), ctx._h, int(max_iter), float(np.float32(tol)),
                                                 _lib.CG_EIGEN_JACOBI, int(bool(verbose)), C.byref(self._h)))

    def close(self):
        if self._h.value:
            self._lib.qdsp_cg_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def solve(self, x, b):
        from ._core import _stream

        it, err, ok = C.c_int(), C.c_double(), C.c_int()
        _lib.check(self._lib.qdsp_cg_solve(self._h, x.data_ptr(), b.data_ptr(), _stream(), C.byref(it), C.byref(err),
                                           C.byref(ok)))
        self.iterations, self.error, self.success = it.value, err.value, bool(ok.value)
        return x, self.success

    def run_fixed(self, x, b, iters: int):
        from ._core import _stream

        _lib.check(self._lib.qdsp_cg_run_fixed(self._h, x.data_ptr(), b.data_ptr(), int(iters), _stream()))
