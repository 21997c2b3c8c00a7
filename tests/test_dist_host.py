"""Host-side logic of the multi-GPU path on CPU: the partition plan (C ABI, pure host) and the ghost-list exchange
protocol under torch.distributed gloo with world_size 2 and 3.  The halo and the product are emulated with numpy from
exactly the lists the GPU path would use, and compared with the undistributed product."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import scipy.sparse as sp

from quadrants_b200 import distributed as qdist
from tests import workloads as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _csr(n=12):
    r, c, v = W.poisson3d_triplets(n, dtype=np.float64)
    return sp.coo_matrix((v, (r, c)), shape=(n ** 3, n ** 3)).tocsr()


def test_row_offsets():
    ro = qdist.even_row_offsets(256 ** 3, 8, multiple=256 * 256)
    assert ro[0] == 0 and ro[-1] == 256 ** 3 and np.all(np.diff(ro) == 32 * 256 * 256)
    ro = qdist.even_row_offsets(12 ** 3, 5, multiple=144)
    assert ro[0] == 0 and ro[-1] == 12 ** 3 and np.all(np.diff(ro) % 144 == 0) and np.all(np.diff(ro) > 0)
    ip = np.array([0, 1, 2, 3, 103, 104, 105, 106, 206], dtype=np.int32)
    ro = qdist.nnz_balanced_row_offsets(ip, 2)
    assert ro[0] == 0 and ro[-1] == 8 and 3 <= ro[1] <= 5


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_plan_host_single_process(world):
    A = _csr(8)
    n = A.shape[0]
    ro = qdist.even_row_offsets(n, world, multiple=64)
    x = np.random.default_rng(0).uniform(-1, 1, n)
    y = A @ x
    for rank in range(world):
        lo, hi = int(ro[rank]), int(ro[rank + 1])
        Al = A[lo:hi]
        idx, ghosts, owner = qdist.plan_host(rank, world, ro, Al.indices)
        assert owner[0] == 0 and owner[-1] == ghosts.shape[0] and owner[rank] == owner[rank + 1]
        assert np.all(np.diff(ghosts) > 0) and not np.any((ghosts >= lo) & (ghosts < hi))
        for q in range(world):
            g = ghosts[owner[q]:owner[q + 1]]
            assert np.all((g >= ro[q]) & (g < ro[q + 1]))
        xl = np.concatenate([x[lo:hi], x[ghosts]])
        Aloc = sp.csr_matrix((Al.data, idx, Al.indptr), shape=(hi - lo, hi - lo + ghosts.shape[0]))
        assert np.array_equal(Aloc @ xl, y[lo:hi])


def test_plan_host_errors():
    from quadrants_b200 import _lib

    with pytest.raises(_lib.QdspError):
        qdist.plan_host(0, 9, np.arange(10), np.zeros(1, np.int32))
    with pytest.raises(_lib.QdspError):
        qdist.plan_host(0, 2, np.array([0, 4, 8]), np.array([9], np.int32))  # column outside the matrix
    with pytest.raises(_lib.QdspError):
        qdist.plan_host(0, 2, np.array([0, 8, 4]), np.array([1], np.int32))  # offsets must ascend


WORKER = r"""
import os, sys
import numpy as np, scipy.sparse as sp, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
from quadrants_b200 import distributed as qdist
from tests import workloads as W
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 10
r, c, v = W.poisson3d_triplets(n, dtype=np.float64)
A = sp.coo_matrix((v, (r, c)), shape=(n**3, n**3)).tocsr()
ro = qdist.even_row_offsets(n**3, world, multiple=n*n)
lo, hi = int(ro[rank]), int(ro[rank+1])
Al = A[lo:hi]
idx, ghosts, owner = qdist.plan_host(rank, world, ro, Al.indices)
sends = qdist.exchange_requests(rank, world, hi - lo, ghosts, owner)
x = np.random.default_rng(1).uniform(-1, 1, n**3)
# emulate the halo: every rank "pushes" the rows it was asked for into the requester's operand vector
vec = np.concatenate([x[lo:hi], np.full(ghosts.shape[0], np.nan)])
zbase = ((hi - lo) + ghosts.shape[0] + 31) & ~31
vec = np.concatenate([vec, np.full(zbase - vec.shape[0] + ghosts.shape[0], np.nan)])   # [owned | ghosts | pad | z-ghosts]
z = np.random.default_rng(2).uniform(-1, 1, n**3)       # stands for D^-1 r' of the whole system
outbox = {{peer: (dst, x[ids], dst_z, z[ids]) for peer, ids, dst, dst_z in sends}}
for peer, ids, dst, dst_z in sends:
    assert np.all(np.diff(ids) > 0) and np.all((ids >= lo) & (ids < hi))
inbox = [None] * world
dist.all_gather_object(inbox, outbox)
for s in range(world):
    if s != rank and rank in inbox[s]:
        dst, vals, dst_z, zvals = inbox[s][rank]
        vec[dst:dst + vals.shape[0]] = vals
        assert dst_z - zbase == dst - (hi - lo), "z slot must mirror the ghost slot"
        vec[dst_z:dst_z + zvals.shape[0]] = zvals
ng = ghosts.shape[0]
assert not np.isnan(vec[:hi - lo + ng]).any(), "a ghost entry was never delivered"
assert not np.isnan(vec[zbase:zbase + ng]).any(), "a z-ghost entry was never delivered"
# the z channel: ghost copies of p updated locally equal what the owners compute (same operands, same operation)
beta = 0.37
p_new = z + beta * x
ghost_new = vec[zbase:zbase + ng] + beta * vec[hi - lo:hi - lo + ng]
assert np.array_equal(ghost_new, p_new[ghosts]), "locally updated ghosts differ from the owners' values"
vec = vec[:hi - lo + ng]
Aloc = sp.csr_matrix((Al.data, idx, Al.indptr), shape=(hi - lo, vec.shape[0]))
y = Aloc @ vec
assert np.array_equal(y, (A @ x)[lo:hi]), "distributed product differs"
# the all-reduced CG scalars: every rank sums the contributions in rank order -> identical bits
part = float(y @ x[lo:hi])
parts = [None] * world
dist.all_gather_object(parts, part)
tot = 0.0
for p in parts:
    tot += p
assert abs(tot - float((A @ x) @ x)) <= 1e-9 * abs(tot)
dist.barrier()
if rank == 0:
    print("DIST_HOST_OK", world)
"""


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_exchange_protocol_gloo(tmp_path, world):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env, cwd=ROOT)
    assert r.returncode == 0 and f"DIST_HOST_OK {world}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_plan_host_dense_owner():
    """A rank that reads at least half of an owner's rows takes the whole slice (one contiguous send run for the
    owner); a sparsely read owner keeps its exact ghost list."""
    ro = np.array([0, 4, 12, 20])
    # rank 0 reads columns 5, 6, 8, 11 of rank 1 (4 of 8 rows: dense) and column 13 of rank 2 (1 of 8: sparse)
    cols = np.array([0, 5, 6, 2, 8, 11, 13, 3], np.int32)
    idx, ghosts, owner = qdist.plan_host(0, 3, ro, cols)
    assert ghosts.tolist() == list(range(4, 12)) + [13]
    assert owner.tolist() == [0, 0, 8, 9]
    back = np.where(idx < 4, idx, ghosts[np.maximum(idx - 4, 0)])
    assert back.tolist() == cols.tolist()
