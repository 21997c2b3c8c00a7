"""quadrants_b200 -- B200-native implementation of the sparse linear-algebra hot path of Quadrants.

    import quadrants_b200 as qd
    qd.init(arch=qd.cuda)
    K = qd.linalg.SparseMatrixBuilder(n, n, max_num_triplets=..., dtype=qd.f32)
    ...
    A = K.build();  y = A @ x;  x, ok = qd.linalg.SparseCG(A, b, max_iter=2000, atol=1e-6).solve()
    qd.algorithms.parallel_sort(keys, values);  qd.algorithms.PrefixSumExecutor(n).run(arr)

Same names, arguments and error behaviour as python/quadrants/linalg and python/quadrants/algorithms/_algorithms.py;
behind them hand-written sm_100a CUDA reached through the C ABI of include/qdsp.h.  No CPU fallback.
"""
from . import _lib  # noqa: F401
from .lang import (Arch, Field, Ndarray, QuadrantsRuntimeError, ScalarNdarray, Vector, cpu, cuda, f32, f64, field, float32,  # noqa: F401
                   float64, get_runtime, gpu, i32, i64, init, int32, int64, ndarray, reset, sync, u32, u64, uint32,
                   uint64, x64)
from . import _core  # noqa: F401,E402
from . import linalg  # noqa: F401,E402
from . import algorithms  # noqa: F401,E402

__version__ = "0.1.0"
