// common.cuh -- shared host/device helpers for the qdsp kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <exception>
#include <stdexcept>
#include <string>

#include "../../include/qdsp.h"

namespace qdsp {

// ---- errors: thrown inside the library, turned into qdsp_status + qdsp_last_error() at the ABI ---------------------
struct Error : std::runtime_error {
  int status;
  Error(int s, const std::string &m) : std::runtime_error(m), status(s) {}
};

#define QDSP_CUDA(expr)                                                                              \
  do {                                                                                               \
    cudaError_t e__ = (expr);                                                                        \
    if (e__ != cudaSuccess)                                                                          \
      throw ::qdsp::Error(QDSP_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__) + " (" + \
                                             __FILE__ + ":" + std::to_string(__LINE__) + ")");      \
  } while (0)

#define QDSP_REQUIRE(cond, status, msg)                   \
  do {                                                    \
    if (!(cond)) throw ::qdsp::Error((status), (msg));    \
  } while (0)

extern std::atomic<int64_t> g_launch_count;
inline void count_launch(int n = 1) { g_launch_count.fetch_add(n, std::memory_order_relaxed); }

#define QDSP_LAUNCH_CHECK()            \
  do {                                 \
    ::qdsp::count_launch();            \
    QDSP_CUDA(cudaGetLastError());     \
  } while (0)

struct DeviceProps {
  int device = -1, sm_count = 148, cc_major = 10, cc_minor = 0;
  size_t l2_bytes = 0, hbm_bytes = 0;
};
const DeviceProps &device_props();  // of the current device; cached per device

// ---- RAII device memory ----------------------------------------------------------------------------------------------
// Scratch pool (scratch_pool_* in abi.cu): the multi-GB temporaries of a CSR build / sort (packed keys, the radix ping-pong
// pair, the collapse spares) are taken from and returned to a per-process cache instead of cudaMalloc / cudaFree -- a
// fresh multi-GB cudaMalloc maps new physical pages and was measured at 12-99 ms per build, more than the kernels.
// Blocks are only handed back after the owning call has synchronised its stream, so reuse needs no event.  The cache is
// capped (QDSP_SCRATCH_POOL_MB, default 16384; 0 disables), dropped by qdsp_release_scratch() and on any out-of-memory.
void *scratch_pool_acquire(size_t bytes, size_t *got);
void scratch_pool_release(void *p, size_t bytes);
void scratch_pool_trim();
void scratch_pool_info(size_t *cached, size_t *cap_now);

class DeviceBuffer {
 public:
  DeviceBuffer() = default;
  explicit DeviceBuffer(size_t bytes) { alloc(bytes); }
  DeviceBuffer(const DeviceBuffer &) = delete;
  DeviceBuffer &operator=(const DeviceBuffer &) = delete;
  DeviceBuffer(DeviceBuffer &&o) noexcept : p_(o.p_), n_(o.n_), cap_(o.cap_), pooled_(o.pooled_) { o.p_ = nullptr; o.n_ = 0; }
  DeviceBuffer &operator=(DeviceBuffer &&o) noexcept {
    if (this != &o) { release(); p_ = o.p_; n_ = o.n_; cap_ = o.cap_; pooled_ = o.pooled_; o.p_ = nullptr; o.n_ = 0; }
    return *this;
  }
  ~DeviceBuffer() { release(); }
  void alloc(size_t bytes) {
    release();
    if (bytes == 0) bytes = 16;
    cudaError_t e = cudaMalloc(&p_, bytes);
    if (e == cudaErrorMemoryAllocation) {  // give the cached scratch back and try once more
      (void)cudaGetLastError();
      scratch_pool_trim();
      e = cudaMalloc(&p_, bytes);
    }
    QDSP_CUDA(e);
    n_ = bytes;
    pooled_ = false;
  }
  // scratch that dies inside the call that made it (after that call has synchronised its stream)
  void alloc_scratch(size_t bytes) {
    release();
    if (bytes == 0) bytes = 16;
    p_ = scratch_pool_acquire(bytes, &cap_);
    n_ = bytes;
    pooled_ = true;
  }
  void release() {
    if (p_) {
      // unwinding: kernels of the failed call may still be running on the block; cudaFree synchronises, the pool does not
      if (pooled_ && std::uncaught_exceptions() == 0) scratch_pool_release(p_, cap_);
      else cudaFree(p_);
    }
    p_ = nullptr;
    n_ = 0;
  }
  template <typename T = void>
  T *get() const { return static_cast<T *>(p_); }
  size_t bytes() const { return n_; }

 private:
  void *p_ = nullptr;
  size_t n_ = 0, cap_ = 0;
  bool pooled_ = false;
};
inline DeviceBuffer scratch_buffer(size_t bytes) {
  DeviceBuffer b;
  b.alloc_scratch(bytes);
  return b;
}

inline int ceil_div(int64_t a, int64_t b) { return static_cast<int>((a + b - 1) / b); }
inline int bits_for(int64_t n) {  // number of bits needed to represent values in [0, n)
  int b = 0;
  while ((int64_t{1} << b) < n) ++b;
  return b;
}

#ifdef __CUDACC__
// ---- device helpers ------------------------------------------------------------------------------------------------
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

// inclusive warp scan (wrapping for integers)
template <typename T>
__device__ __forceinline__ T warp_inclusive_scan(T v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T n = __shfl_up_sync(FULL, v, o);
    if (lane_id() >= (unsigned)o) v += n;
  }
  return v;
}

// streaming (read-once) 128-bit load that does not pollute L1
__device__ __forceinline__ int4 ld_stream_v4(const void *p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ int2 ld_stream_v2(const void *p) {
  int2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ int ld_stream_s32(const void *p) {
  int r;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

// decoupled look-back descriptors: one 64-bit word = {status:2 | value:62}, read/written with relaxed gpu-scope
// accesses; status and value travel together so no fence is needed between them.
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned *p) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(unsigned *p, unsigned v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
#endif  // __CUDACC__

}  // namespace qdsp
