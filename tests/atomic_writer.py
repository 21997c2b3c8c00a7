"""Test helper: a CUDA kernel that fills a SparseMatrixBuilder buffer EXACTLY the way the reference's JIT kernels do --
`insert_triplet_f32 / _f64` (quadrants/runtime/llvm/runtime_module/internal_functions.h:12-39):

    id = atomic_add(buf[0], 1);  buf[1 + 3 id .. 3 + 3 id] = (i, j, bits(v))

compiled with NVRTC (cuda-python) at test time and launched on torch's current stream through the driver API.  The
arrival order is whatever the hardware's atomics make it (nondeterministic), which is the point: SURVEY 8(a) A2."""
import numpy as np

SRC = r"""
extern "C" __global__ void insert_f32(int *buf, const int *rows, const int *cols, const float *vals, int n) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  int id = atomicAdd(buf, 1);
  buf[1 + 3 * id] = rows[t];
  buf[2 + 3 * id] = cols[t];
  buf[3 + 3 * id] = __float_as_int(vals[t]);
}
extern "C" __global__ void insert_f64(long long *buf, const int *rows, const int *cols, const double *vals, int n) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  long long id = (long long)atomicAdd((unsigned long long *)buf, 1ull);
  buf[1 + 3 * id] = rows[t];
  buf[2 + 3 * id] = cols[t];
  buf[3 + 3 * id] = __double_as_longlong(vals[t]);
}
"""


def compile_ptx() -> bytes:
    from cuda.bindings import nvrtc

    def ck(r):
        if int(r[0]) != 0:
            raise RuntimeError(f"nvrtc error {r[0]}")
        return r[1:] if len(r) > 2 else (r[1] if len(r) == 2 else None)

    prog = ck(nvrtc.nvrtcCreateProgram(SRC.encode(), b"atomic_writer.cu", 0, [], []))
    opts = [b"--gpu-architecture=compute_100"]
    res = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    if int(res[0]) != 0:
        n = ck(nvrtc.nvrtcGetProgramLogSize(prog))
        log = b" " * n
        nvrtc.nvrtcGetProgramLog(prog, log)
        raise RuntimeError(log.decode())
    n = ck(nvrtc.nvrtcGetPTXSize(prog))
    ptx = b" " * n
    ck(nvrtc.nvrtcGetPTX(prog, ptx))
    return ptx


class AtomicWriter:
    def __init__(self):
        from cuda.bindings import driver

        self.driver = driver
        ptx = compile_ptx()
        err, self.module = driver.cuModuleLoadData(np.char.array(ptx).ctypes.data)
        assert int(err) == 0, err
        self.fn = {}
        for name in ("insert_f32", "insert_f64"):
            err, f = driver.cuModuleGetFunction(self.module, name.encode())
            assert int(err) == 0, err
            self.fn[name] = f

    def insert(self, buffer_ptr: int, rows, cols, vals):
        """rows / cols: int32 CUDA tensors; vals: float32 or float64 CUDA tensor; buffer_ptr: the builder buffer."""
        import torch

        n = int(rows.numel())
        name = "insert_f32" if vals.dtype == torch.float32 else "insert_f64"
        args = [np.array([buffer_ptr], dtype=np.uint64), np.array([rows.data_ptr()], dtype=np.uint64),
                np.array([cols.data_ptr()], dtype=np.uint64), np.array([vals.data_ptr()], dtype=np.uint64),
                np.array([n], dtype=np.int32)]
        argp = np.array([a.ctypes.data for a in args], dtype=np.uint64)
        stream = torch.cuda.current_stream().cuda_stream
        (err,) = self.driver.cuLaunchKernel(self.fn[name], (n + 255) // 256, 1, 1, 256, 1, 1, 0, stream,
                                            argp.ctypes.data, 0)
        assert int(err) == 0, err
