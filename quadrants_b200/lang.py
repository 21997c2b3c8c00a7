"""The sliver of `quadrants.lang` the sparse path touches: qd.init, dtypes, ndarrays / 1-D fields and the runtime
object whose `.prog` the linalg classes hand to the C++ layer (python/quadrants/lang/impl.py get_runtime()).

Device memory is PyTorch's (plumbing only): an Ndarray is a contiguous CUDA torch tensor; the C ABI receives its raw
pointer, exactly like Program::get_ndarray_data_ptr_as_int (quadrants/program/program.cpp:485-494)."""
from __future__ import annotations

import numpy as np

from . import _lib


class QuadrantsRuntimeError(RuntimeError):
    """python/quadrants/lang/exception.py"""


class DataType:
    def __init__(self, name, np_dtype, torch_name, storage_np=None):
        self.name, self.np_dtype, self._torch_name = name, np.dtype(np_dtype), torch_name
        # dtype of the torch tensor that holds the bytes (u64 rides in an int64 tensor: same bits, full torch support)
        self.storage_np = np.dtype(storage_np if storage_np is not None else np_dtype)

    @property
    def torch_dtype(self):
        import torch

        return getattr(torch, self._torch_name)

    def __repr__(self):
        return self.name

    def __call__(self, v):
        return self.np_dtype.type(v)


f32 = float32 = DataType("f32", np.float32, "float32")
f64 = float64 = DataType("f64", np.float64, "float64")
i32 = int32 = DataType("i32", np.int32, "int32")
u32 = uint32 = DataType("u32", np.uint32, "uint32")
i64 = int64 = DataType("i64", np.int64, "int64")
u64 = uint64 = DataType("u64", np.uint64, "int64", storage_np=np.int64)
_ALL = {t.name: t for t in (f32, f64, i32, u32, i64, u64)}


def to_dtype(x) -> DataType:
    if isinstance(x, DataType):
        return x
    if x is float:
        return f32
    if x is int:
        return i32
    d = np.dtype(x)
    for t in _ALL.values():
        if t.np_dtype == d:
            return t
    raise QuadrantsRuntimeError(f"unsupported dtype {x}")


class Arch:
    """_qd_core.Arch: only cuda exists here (no multi-backend dispatch, no CPU fallback)."""
    cuda = "cuda"
    x64 = "x64"
    arm64 = "arm64"
    vulkan = "vulkan"


cuda = Arch.cuda
cpu = x64 = Arch.x64
gpu = Arch.cuda


class _Config:
    def __init__(self, arch, device):
        self.arch, self.device = arch, device


class Program:
    """Stand-in for _qd_core.Program: resolves ndarrays to raw device pointers and owns the device binding."""

    def __init__(self, arch, device):
        self._config = _Config(arch, device)

    def config(self):
        return self._config

    def get_ndarray_data_ptr_as_int(self, nd) -> int:
        return nd.data_ptr()

    def stream(self) -> int:
        import torch

        return torch.cuda.current_stream().cuda_stream

    def create_sparse_matrix(self, n, m, dtype, storage_format="col_major"):
        """export_lang.cpp:518-526 -> make_cu_sparse_matrix on arch=cuda (the storage format does not apply: CSR)."""
        from . import _core

        return _core.make_cu_sparse_matrix(n, m if m else n, dtype)

    def make_sparse_matrix_from_ndarray(self, sm, ndarray):
        """export_lang.cpp:527-532 / sparse_matrix.cpp:320-342: N x 3 float (row, col, value) triplets -> sm."""
        import torch

        trip = ndarray.tensor.reshape(-1).view(-1, 3)
        sm.build_csr_from_coo(trip[:, 0].to(torch.int32).contiguous(), trip[:, 1].to(torch.int32).contiguous(),
                              trip[:, 2].to(sm.get_data_type().torch_dtype).contiguous())

    def synchronize(self):
        import torch

        torch.cuda.synchronize()


class _Runtime:
    prog: Program | None = None


_runtime = _Runtime()


def get_runtime():
    return _runtime


def init(arch=cuda, device: int | None = None, **_ignored):
    """qd.init(arch=qd.cuda).  Any other arch is refused: this package is the CUDA path of the sparse stack only."""
    import os

    import torch

    if arch != Arch.cuda:
        raise QuadrantsRuntimeError(
            f"quadrants_b200 implements arch=qd.cuda only (got {arch}); the reference's Eigen CPU path is its oracle, "
            "not a fallback")
    if not torch.cuda.is_available():
        raise QuadrantsRuntimeError("qd.init(arch=qd.cuda): no CUDA device visible and there is no CPU fallback")
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count()
    torch.cuda.set_device(device)
    _lib.check(_lib.load().qdsp_init(device))
    _runtime.prog = Program(arch, device)
    return _runtime.prog


def reset():
    _runtime.prog = None


def sync():
    import torch

    torch.cuda.synchronize()


def _require_prog() -> Program:
    if _runtime.prog is None:
        raise QuadrantsRuntimeError("call qd.init(arch=qd.cuda) first")
    return _runtime.prog


class Ndarray:
    """1-D dense device array (qd.ndarray / ScalarNdarray).  `.arr` is what the reference passes to C++."""

    def __init__(self, dtype, arr_shape, offset: int = 0, _tensor=None):
        import torch

        prog = _require_prog()
        self.dtype = to_dtype(dtype)
        if isinstance(arr_shape, (int, np.integer)):
            arr_shape = (int(arr_shape),)
        self.shape = tuple(int(s) for s in arr_shape)
        self.element_shape = ()
        self.offset = int(offset)  # qd.field(..., offset=k): logical index i lives at storage index i - k
        if _tensor is not None:
            self.tensor = _tensor
        else:
            n = int(np.prod(self.shape)) if self.shape else 1
            self.tensor = torch.zeros(n, dtype=self.dtype.torch_dtype, device=f"cuda:{prog.config().device}")
        self.arr = self

    def data_ptr(self) -> int:
        return self.tensor.data_ptr()

    def to_numpy(self):
        return self.tensor.detach().cpu().numpy().view(self.dtype.np_dtype).reshape(self.shape + self.element_shape)

    def from_numpy(self, a):
        import torch

        a = np.ascontiguousarray(a, dtype=self.dtype.np_dtype).reshape(-1)
        if a.shape[0] != self.tensor.shape[0]:
            raise QuadrantsRuntimeError(f"from_numpy: shape mismatch {a.shape} vs {self.shape + self.element_shape}")
        self.tensor.copy_(torch.from_numpy(a.view(self.dtype.storage_np)))

    def fill(self, v):
        if self.dtype.storage_np != self.dtype.np_dtype:
            v = int(np.array([v], dtype=self.dtype.np_dtype).view(self.dtype.storage_np)[0])
        self.tensor.fill_(v)

    def _idx(self, i):
        if isinstance(i, tuple):
            i = i[0]
        j = int(i) - self.offset
        if not 0 <= j < self.tensor.shape[0]:
            raise IndexError(i)
        return j

    def __getitem__(self, i):
        if self.element_shape:
            k = self.element_shape[0]
            return self.tensor[int(i) * k:(int(i) + 1) * k].tolist()
        return self.tensor[self._idx(i)].item()

    def __setitem__(self, i, v):
        if self.element_shape:
            import torch

            k = self.element_shape[0]
            self.tensor[int(i) * k:(int(i) + 1) * k] = torch.as_tensor(v, dtype=self.tensor.dtype)
            return
        self.tensor[self._idx(i)] = v

    def __len__(self):
        return self.shape[0]


ScalarNdarray = Ndarray


def ndarray(dtype, shape):
    return Ndarray(dtype, shape)


class Vector:
    """qd.Vector.ndarray(n, dtype, shape): a dense AOS array of n-vectors (python/quadrants/lang/matrix.py
    Vector.ndarray) -- what `SparseMatrix.build_from_ndarray` takes (tests/python/test_sparse_matrix.py:70-85).  Storage is
    one contiguous tensor of shape*n scalars; `a[i] = (..)` / `a[i]` address whole elements."""

    @staticmethod
    def ndarray(n, dtype, shape, layout=None):
        if isinstance(shape, (int, np.integer)):
            shape = (int(shape),)
        total = int(np.prod(shape)) * int(n)
        a = Ndarray(dtype, total)
        a.shape = tuple(int(s) for s in shape)
        a.element_shape = (int(n),)
        return a


def field(dtype, shape, offset=0):
    """1-D dense field; enough of qd.field for parallel_sort / PrefixSumExecutor callers."""
    if isinstance(offset, (tuple, list)):
        offset = offset[0]
    return Field(dtype, shape, offset=offset)


class Field(Ndarray):
    """qd.field(...) result.  Same dense device storage as Ndarray; a distinct type because the reference's Python layer
    dispatches on it (e.g. SparseSolver.solve: Field -> numpy result, Ndarray -> Ndarray result,
    python/quadrants/linalg/sparse_solver.py:103-110)."""
