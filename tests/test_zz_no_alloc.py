"""qdsp_inclusive_scan_i32_static: the prefix sum for hosts that may not allocate device memory.

Who needs it: an application that gave all of its device memory to one fixed arena at start-up (so there is no allocator
left to draw the scan's tmp buffer from), or a fixed-rate loop in which cudaMalloc / cudaFree are banned because they stall
the whole device.  For those the scan's tile descriptors live in a ~1 MB array compiled into libqdsp.so (one copy per
GPU), and the call takes only the data pointer, the length and the stream.

What is checked here, through the C ABI and bit for bit (it is integer work): the static-scratch entry point returns the
same sums as qdsp_inclusive_scan_i32 with a caller-owned tmp buffer, on two calls in a row that share the array (the
second one must not be disturbed by what the first left behind), and a length beyond qdsp_scan_static_max_n() is refused
with a message instead of overrunning the array.

The file name makes pytest collect it last: under `pytest -x` a failure here then hides no other test.  It passes alone
in a fresh process (it is then the first thing that touches the GPU, so it also covers the per-device address look-up of
the first call) and after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    import quadrants_b200 as qd
    from quadrants_b200 import _lib

    qd.init(arch=qd.cuda)
    return _lib.load()


def _scan_with_tmp(lib, a):
    """The existing entry point: scratch comes from the caller (here a torch allocation)."""
    import torch

    from quadrants_b200 import _lib

    t = torch.from_numpy(a).cuda()
    tmp = torch.empty(lib.qdsp_scan_tmp_bytes(a.size), dtype=torch.uint8, device="cuda")
    _lib.check(lib.qdsp_inclusive_scan_i32(t.data_ptr(), a.size, tmp.data_ptr(), tmp.numel(),
                                           torch.cuda.current_stream().cuda_stream))
    return t.cpu().numpy()


def _scan_static(lib, t):
    """The entry point under test: no tmp argument, nothing allocated inside the call."""
    import torch

    from quadrants_b200 import _lib

    _lib.check(lib.qdsp_inclusive_scan_i32_static(t.data_ptr(), t.numel(), torch.cuda.current_stream().cuda_stream))


def test_static_scratch_scan_matches_the_tmp_buffer_scan_twice_in_a_row(lib):
    import torch

    rng = np.random.default_rng(20261106)
    # two different lengths: many tiles first, then fewer and ragged, so the second call finds the first one's
    # descriptors (and its ticket counter) still in the shared array and has to clear them itself
    sizes = [(1 << 22) + 4099, 300001]
    inputs = [rng.integers(-2**31, 2**31, n, dtype=np.int64).astype(np.int32) for n in sizes]
    expected = [_scan_with_tmp(lib, a) for a in inputs]
    for a, e in zip(inputs, expected):  # host-side sanity of the yardstick itself (wrapping cumulative sum)
        assert np.array_equal(e, np.cumsum(a.astype(np.int64)).astype(np.int32))
    # the device inputs exist before the calls under test; the two calls run back to back on one stream
    on_device = [torch.from_numpy(a).cuda() for a in inputs]
    torch.cuda.synchronize()
    for t in on_device:
        _scan_static(lib, t)
    torch.cuda.synchronize()
    for t, e, n in zip(on_device, expected, sizes):
        got = t.cpu().numpy()
        assert np.array_equal(got, e), f"n={n}: first mismatch at {int(np.argmax(got != e))}"


def test_static_scratch_scan_same_array_again_and_tiny_lengths(lib):
    import torch

    # a third and fourth use of the shared array in this process, one of them a single partial tile; and n == 0 is a no-op
    for n in (1, 8193):
        a = np.arange(1, n + 1, dtype=np.int32)
        t = torch.from_numpy(a).cuda()
        _scan_static(lib, t)
        assert np.array_equal(t.cpu().numpy(), _scan_with_tmp(lib, a))
    assert lib.qdsp_inclusive_scan_i32_static(None, 0, torch.cuda.current_stream().cuda_stream) == 0


def test_static_scratch_scan_refuses_lengths_beyond_the_array(lib):
    import torch

    limit = lib.qdsp_scan_static_max_n()
    assert limit == 1 << 26
    t = torch.zeros(16, dtype=torch.int32, device="cuda")  # never touched: the length check comes first
    st = lib.qdsp_inclusive_scan_i32_static(t.data_ptr(), limit + 1, torch.cuda.current_stream().cuda_stream)
    assert st == 1  # QDSP_ERR_INVALID
    msg = lib.qdsp_last_error().decode()
    assert str(limit) in msg and "qdsp_inclusive_scan_i32" in msg, msg
    torch.cuda.synchronize()
    assert int(t.sum()) == 0


def test_static_scratch_scan_at_exactly_the_limit(lib):
    import torch

    n = lib.qdsp_scan_static_max_n()
    a = np.random.default_rng(7).integers(-1000, 1000, n, dtype=np.int64).astype(np.int32)
    t = torch.from_numpy(a).cuda()
    _scan_static(lib, t)
    assert np.array_equal(t.cpu().numpy(), _scan_with_tmp(lib, a))
