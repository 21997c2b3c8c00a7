"""The reference's modern device-wide algorithm library with its names, argument lists and error
behaviour, over the kernels of libqdsp:

    reduce_{add,min,max}(arr, out, scratch, n, dtype, log256_max_n)              _reduce.py:448-507
    exclusive_scan_{add,min,max}(arr, out, scratch, n, dtype, log256_max_n)      _scan.py:496-541
    select(arr, flags, out, num_out, scratch, n, log256_max_n)                   _select.py:124-147
    reduce_by_key_add(keys_in, values_in, keys_out, values_out, num_runs, scratch, n, value_dtype, log256_max_n)
                                                                                 _reduce_by_key.py:145-177
    sort(keys, tmp_keys, values, tmp_values, scratch, n, key_dtype, has_values, end_bit, log256_max_n)
                                                                                 _radix_sort.py:277-344
    {reduce,exclusive_scan,select,reduce_by_key,sort}_scratch_slots(...)         same files

In the reference these are @qd.func's composed into a caller's @qd.kernel; there is no JIT here, so they are host
calls that enqueue the kernels on the current stream.  What the reference's design is for survives: `n` may be a
device-resident count (a 1-element / 0-d i32 ndarray -- the reference's `count[0]` / `n[()]`), which the kernels read on
the device, so no call ever reads a count back to the host and the launch topology depends only on the array length.
`log256_max_n` is validated like the reference does and otherwise only bounds a host-side `n`: the kernels here are
3-pass (tile reduce, one-CTA pass over <= N/2048 partials, tile emit), not a log256-deep staircase.

Scratch: `*_scratch_slots` return max(the reference's formula, what these kernels need), so a buffer sized by either
works and the u32-for-4-byte / u64-for-8-byte typing rule of the reference carries over.
"""
from __future__ import annotations

from .. import _lib
from ..lang import Ndarray, f32, f64, i32, i64, u32, u64, _require_prog

BLOCK_DIM = 256  # the reference's tile, only used by the scratch-slot formulas below (_reduce.py:42)
_MAX_LOG256_MAX_N = 4  # _reduce.py:58

_ELEM = {"i32": _lib.ELEM_I32, "u32": _lib.ELEM_U32, "f32": _lib.ELEM_F32, "i64": _lib.ELEM_I64, "u64": _lib.ELEM_U64,
         "f64": _lib.ELEM_F64}
_SUPPORTED = (i32, u32, f32, i64, u64, f64)
_OWN_TILE = 2048  # smallest tile of device_algos.cu (8-byte elements); the 4-byte tile is 4096
_OWN_PAD = 64  # 256-byte alignment of the scratch areas, in 4-byte slots


def _validate_log256_max_n(log256_max_n):
    """_reduce.py:69-82"""
    if log256_max_n > _MAX_LOG256_MAX_N:
        raise ValueError(
            f"log256_max_n={log256_max_n} exceeds the maximum of {_MAX_LOG256_MAX_N}. Quadrants tensors are indexed "
            f"by signed 32-bit integers, so they hold at most 2**31 - 1 elements; BLOCK_DIM ** {_MAX_LOG256_MAX_N} "
            f"== 256 ** {_MAX_LOG256_MAX_N} == 2**32 already covers that range, so a larger depth can never address "
            f"a valid tensor.")


def _depth_for_n(n: int) -> int:
    """Smallest D >= 1 with 256**D >= n (_reduce.py:315-323, _radix_sort.py:347-355)."""
    d, cap = 1, BLOCK_DIM
    while cap < n:
        cap *= BLOCK_DIM
        d += 1
    return d


def _cdiv(a, b):
    return (a + b - 1) // b


def _own_partial_slots(n):
    return _cdiv(max(int(n), 0), _OWN_TILE) + _OWN_PAD


# ---- scratch sizing ------------------------------------------------------------------------------------------------------
def reduce_scratch_slots(n, log256_max_n: int = None) -> int:
    """Slots (u32 for 4-byte dtypes, u64 for 8-byte ones) `reduce_*` needs for a length-n input (_reduce.py:343-369)."""
    if log256_max_n is None:
        log256_max_n = _depth_for_n(n)
    _validate_log256_max_n(log256_max_n)
    cursor, cur = 0, n
    for _ in range(log256_max_n - 1):
        cur = _cdiv(cur, BLOCK_DIM)
        cursor += cur
    return max(cursor, _own_partial_slots(n), 1)


def exclusive_scan_scratch_slots(n, log256_max_n: int = None) -> int:
    """_scan.py:317-345"""
    if log256_max_n is None:
        log256_max_n = _depth_for_n(n)
    _validate_log256_max_n(log256_max_n)
    cursor = 0
    if log256_max_n > 1:
        cursor = cur = _cdiv(n, BLOCK_DIM)
        for _ in range(log256_max_n - 2):
            cur = _cdiv(cur, BLOCK_DIM)
            cursor += cur
    return max(cursor, _own_partial_slots(n), 1)


def _scan_total_scratch_slots(n, partials_cursor):
    """Cursor after the partials staircase of an in-place scan of n elements (_scan.py:190-206, _reduce.py:85-113)."""
    cursor, t = partials_cursor, n
    for _ in range(8):
        t = _cdiv(t, BLOCK_DIM)
        cursor += t * (t > 1)
    return cursor


def select_scratch_slots(n: int) -> int:
    """u32 slots (_select.py:150-166: n write indices + the scan partials; these kernels need only the partials)."""
    b0 = _cdiv(n, BLOCK_DIM)
    ref = _scan_total_scratch_slots(b0, n + b0) if n > 0 else 0
    return max(ref, _own_partial_slots(n), 1)


def reduce_by_key_scratch_slots(n: int) -> int:
    """u32 slots (_reduce_by_key.py:179-197)."""
    if n <= 0:
        ref = 0
    elif n <= BLOCK_DIM:
        ref = n
    else:
        b0 = _cdiv(n, BLOCK_DIM)
        ref = _scan_total_scratch_slots(b0, n + b0)
    return max(ref, _own_partial_slots(n), 1)


def sort_scratch_slots(n, log256_max_n: int = None) -> int:
    """u32 slots, whatever the key width (_radix_sort.py:358-398).  The onesweep here needs its 8 digit histograms,
    8 tickets and one look-back word per (tile, digit): 2112 + 256 * ceil(n / 2048) slots, which exceeds the
    reference's ~n only for small n."""
    if log256_max_n is None:
        log256_max_n = _depth_for_n(n)
    _validate_log256_max_n(log256_max_n)
    hist_len = _cdiv(n, BLOCK_DIM) * BLOCK_DIM
    cursor = nn = hist_len
    for _ in range(log256_max_n - 1):
        nn = _cdiv(nn, BLOCK_DIM)
        cursor += nn
    own = 64 + 8 * 256 + 256 * max(_cdiv(max(int(n), 0), 2048), 1) + _OWN_PAD
    return max(cursor, own, 1)


# ---- argument plumbing ---------------------------------------------------------------------------------------------------
def _elem_of(dtype, what):
    if dtype not in _SUPPORTED:
        raise NotImplementedError(f"device {what} dtype {dtype} not supported")
    return _ELEM[dtype.name]


def _nd(x, name):
    if not isinstance(x, Ndarray):
        raise RuntimeError(f"{name}: expected a 1-D qd.field / qd.ndarray")
    return x


def _count(n, length, log256_max_n, who):
    """(capacity, device pointer or None).  A host int bounds the capacity itself; a device count is read by the kernels."""
    _validate_log256_max_n(log256_max_n)
    if isinstance(n, Ndarray):
        if n.dtype is not i32 or n.tensor.shape[0] < 1:
            raise RuntimeError(f"{who}: a device-resident n must be an i32 ndarray with one element")
        cap = min(length, BLOCK_DIM ** log256_max_n)
        return cap, n.data_ptr()
    n = int(n)
    if n < 0 or n > length:
        raise ValueError(f"{who}: n={n} outside [0, {length}]")
    if n > BLOCK_DIM ** log256_max_n:
        raise ValueError(f"{who}: n={n} exceeds the capacity 256**{log256_max_n} the call is sized for")
    return n, None


def _scratch_bytes(scratch):
    s = _nd(scratch, "scratch")
    return s.data_ptr(), s.tensor.numel() * s.tensor.element_size()


def _check_dtype(arr, dtype, who):
    if arr.dtype is not dtype:
        raise RuntimeError(f"{who}: tensor dtype {arr.dtype} does not match the dtype argument {dtype}")


def _reduce(op, arr, out, scratch, n, dtype, log256_max_n, who):
    elem = _elem_of(dtype, "reduce")
    arr, out = _nd(arr, who), _nd(out, who)
    _check_dtype(arr, dtype, who)
    _check_dtype(out, dtype, who)
    if out.tensor.shape[0] < 1:
        raise RuntimeError(f"{who}: out must hold one element")
    cap, nptr = _count(n, arr.tensor.shape[0], log256_max_n, who)
    sptr, sbytes = _scratch_bytes(scratch)
    _lib.check(_lib.load().qdsp_reduce(arr.data_ptr(), out.data_ptr(), cap, nptr, elem, op, sptr, sbytes,
                                       _require_prog().stream()))


def reduce_add(arr, out, scratch, n, dtype, log256_max_n):
    """out[0] = sum(arr[0:n]); integer sums wrap, float sums use a fixed tree order (reproducible)."""
    _reduce(_lib.OP_ADD, arr, out, scratch, n, dtype, log256_max_n, "reduce_add")


def reduce_min(arr, out, scratch, n, dtype, log256_max_n):
    """out[0] = min(arr[0:n]); the dtype's largest value (+inf for floats) when n == 0."""
    _reduce(_lib.OP_MIN, arr, out, scratch, n, dtype, log256_max_n, "reduce_min")


def reduce_max(arr, out, scratch, n, dtype, log256_max_n):
    """out[0] = max(arr[0:n]); the dtype's smallest value (-inf for floats, 0 for unsigned) when n == 0."""
    _reduce(_lib.OP_MAX, arr, out, scratch, n, dtype, log256_max_n, "reduce_max")


def _scan(op, arr, out, scratch, n, dtype, log256_max_n, who):
    elem = _elem_of(dtype, "scan")
    arr, out = _nd(arr, who), _nd(out, who)
    if out is arr or out.data_ptr() == arr.data_ptr():
        raise ValueError(f"{who}: out must be a distinct buffer from arr (no in-place scan)")
    _check_dtype(arr, dtype, who)
    _check_dtype(out, dtype, who)
    if out.tensor.shape[0] < arr.tensor.shape[0]:
        raise RuntimeError(f"{who}: arr and out must have the same shape")
    cap, nptr = _count(n, arr.tensor.shape[0], log256_max_n, who)
    sptr, sbytes = _scratch_bytes(scratch)
    _lib.check(_lib.load().qdsp_exclusive_scan(arr.data_ptr(), out.data_ptr(), cap, nptr, elem, op, sptr, sbytes,
                                               _require_prog().stream()))


def exclusive_scan_add(arr, out, scratch, n, dtype, log256_max_n):
    """out[i] = sum(arr[0:i]), out[0] = 0."""
    _scan(_lib.OP_ADD, arr, out, scratch, n, dtype, log256_max_n, "exclusive_scan_add")


def exclusive_scan_min(arr, out, scratch, n, dtype, log256_max_n):
    """out[i] = min(arr[0:i]), out[0] = the dtype's largest value."""
    _scan(_lib.OP_MIN, arr, out, scratch, n, dtype, log256_max_n, "exclusive_scan_min")


def exclusive_scan_max(arr, out, scratch, n, dtype, log256_max_n):
    """out[i] = max(arr[0:i]), out[0] = the dtype's smallest value."""
    _scan(_lib.OP_MAX, arr, out, scratch, n, dtype, log256_max_n, "exclusive_scan_max")


def select(arr, flags, out, num_out, scratch, n, log256_max_n):
    """Stream compaction: out[0:num_out[0]] = arr[i] for flags[i] == 1, in input order; flags must be exactly 0/1."""
    arr, flags, out, num_out = _nd(arr, "select"), _nd(flags, "select"), _nd(out, "select"), _nd(num_out, "select")
    if flags.dtype is not i32:
        raise RuntimeError("select: flags must be a qd.i32 tensor")
    if num_out.dtype is not i32 or num_out.tensor.shape[0] < 1:
        raise RuntimeError("select: num_out must be a 1-element qd.i32 tensor")
    if out.dtype is not arr.dtype:
        raise RuntimeError("select: out must have the dtype of arr")
    length = arr.tensor.shape[0]
    if flags.tensor.shape[0] < length or out.tensor.shape[0] < length:
        raise RuntimeError("select: flags and out must be at least as long as arr")
    cap, nptr = _count(n, length, log256_max_n, "select")
    sptr, sbytes = _scratch_bytes(scratch)
    _lib.check(_lib.load().qdsp_select(arr.data_ptr(), flags.data_ptr(), out.data_ptr(), num_out.data_ptr(), cap, nptr,
                                       arr.dtype.np_dtype.itemsize, sptr, sbytes, _require_prog().stream()))


def reduce_by_key_add(keys_in, values_in, keys_out, values_out, num_runs, scratch, n, value_dtype, log256_max_n):
    """Collapses each run of equal consecutive keys into (key, sum of values); num_runs[0] = number of runs."""
    who = "reduce_by_key_add"
    keys_in, values_in = _nd(keys_in, who), _nd(values_in, who)
    keys_out, values_out, num_runs = _nd(keys_out, who), _nd(values_out, who), _nd(num_runs, who)
    for t, what in ((keys_in.dtype, "key"), (value_dtype, "value")):
        if t not in (i32, u32, f32):
            raise NotImplementedError(f"reduce_by_key_add: {what} dtype {t} not supported (i32, u32, f32)")
    _check_dtype(values_in, value_dtype, who)
    _check_dtype(values_out, value_dtype, who)
    if keys_out.dtype is not keys_in.dtype:
        raise RuntimeError(f"{who}: keys_out must have the dtype of keys_in")
    if num_runs.dtype is not i32 or num_runs.tensor.shape[0] < 1:
        raise RuntimeError(f"{who}: num_runs must be a 1-element qd.i32 tensor")
    length = keys_in.tensor.shape[0]
    if min(values_in.tensor.shape[0], keys_out.tensor.shape[0], values_out.tensor.shape[0]) < length:
        raise RuntimeError(f"{who}: values_in, keys_out and values_out must be at least as long as keys_in")
    cap, nptr = _count(n, length, log256_max_n, who)
    sptr, sbytes = _scratch_bytes(scratch)
    _lib.check(_lib.load().qdsp_reduce_by_key_add(
        keys_in.data_ptr(), values_in.data_ptr(), keys_out.data_ptr(), values_out.data_ptr(), num_runs.data_ptr(), cap,
        nptr, _ELEM[keys_in.dtype.name], _ELEM[value_dtype.name], sptr, sbytes, _require_prog().stream()))


def sort(keys, tmp_keys, values, tmp_values, scratch, n, key_dtype, has_values, end_bit, log256_max_n):
    """Ascending stable LSD radix sort of the low `end_bit` bits of keys[0:n] in place (values permuted alike when
    has_values); tmp_keys / tmp_values are the ping-pong buffers, `n` a 0-d i32 ndarray (or an int)."""
    if key_dtype not in _SUPPORTED:
        raise NotImplementedError(f"sort key dtype {key_dtype} not supported")
    keys, tmp_keys = _nd(keys, "sort"), _nd(tmp_keys, "sort")
    _check_dtype(keys, key_dtype, "sort")
    _check_dtype(tmp_keys, key_dtype, "sort")
    width = key_dtype.np_dtype.itemsize * 8
    if end_bit <= 0 or end_bit % 8 or end_bit > width or (end_bit // 8) % 2:
        raise ValueError(f"sort: end_bit={end_bit} must be a positive multiple of 8, at most the key width {width}, "
                         "giving an even number of digit passes")
    length = keys.tensor.shape[0]
    if tmp_keys.tensor.shape[0] < length:
        raise RuntimeError("sort: tmp_keys must have the shape of keys")
    vbytes, vptr, tvptr = 0, None, None
    if has_values:
        values, tmp_values = _nd(values, "sort"), _nd(tmp_values, "sort")
        if values.dtype not in _SUPPORTED:
            raise NotImplementedError(f"sort value dtype {values.dtype} not supported")
        if tmp_values.dtype is not values.dtype:
            raise RuntimeError("sort: tmp_values must have the dtype of values")
        if min(values.tensor.shape[0], tmp_values.tensor.shape[0]) < length:
            raise RuntimeError("sort: values and tmp_values must have the shape of keys")
        vbytes, vptr, tvptr = values.dtype.np_dtype.itemsize, values.data_ptr(), tmp_values.data_ptr()
    cap, nptr = _count(n, length, log256_max_n, "sort")
    sptr, sbytes = _scratch_bytes(scratch)
    _lib.check(_lib.load().qdsp_sort(keys.data_ptr(), tmp_keys.data_ptr(), vptr, tvptr, cap, nptr,
                                     _ELEM[key_dtype.name], vbytes, int(end_bit), sptr, sbytes,
                                     _require_prog().stream()))


__all__ = [
    "exclusive_scan_add", "exclusive_scan_max", "exclusive_scan_min", "exclusive_scan_scratch_slots",
    "reduce_add", "reduce_by_key_add", "reduce_by_key_scratch_slots", "reduce_max", "reduce_min",
    "reduce_scratch_slots", "select", "select_scratch_slots", "sort", "sort_scratch_slots",
]
