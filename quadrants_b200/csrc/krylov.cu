// krylov.cu -- dense vector kernels behind qd.linalg.MatrixFreeCG / MatrixFreeBICGSTAB.
//
// The reference writes these solvers in its DSL (python/quadrants/linalg/matrixfree_cg.py:23-310): every dot product is a
// JIT reduction kernel returning a host scalar, every vector update a separate kernel (update_x, update_r, update_p, ...).
// Here: one deterministic dot kernel that forms up to two products in a single pass (fixed tree: per-thread strided sums
// -> warp shuffle -> block -> the last block adds the block sums in index order; accumulated in double for both dtypes),
// one linear-combination kernel out = sum_i c_i x_i (i <= 4, out may alias an input), and the fused CG step
// x += a p, r -= a Ap, r.r in one pass.  All HBM-bound streaming kernels: grid-stride loops of coalesced element-wide
// loads (a warp reads 128 / 256 contiguous bytes per instruction), grid = a multiple of the SM count.
#include "host.hpp"

namespace qdsp {

namespace {
constexpr int KV_THREADS = 256;

struct DotScratch {
  double *partials;  // 2 * grid
  unsigned *ticket;
  double *out;       // 2 doubles
};

template <typename T>
__device__ __forceinline__ void block_dot_finish(double s0, double s1, const DotScratch &sc) {
  __shared__ double sh0[KV_THREADS / 32], sh1[KV_THREADS / 32];
  __shared__ bool last;
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { sh0[warp] = s0; sh1[warp] = s1; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int w = 0; w < KV_THREADS / 32; ++w) { a += sh0[w]; b += sh1[w]; }
    sc.partials[2 * blockIdx.x] = a;
    sc.partials[2 * blockIdx.x + 1] = b;
    __threadfence();
    last = atomicAdd(sc.ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last) {
    // the last block adds the block sums: thread t takes blocks t, t + 256, ... in index order, then the same fixed
    // warp / block tree as above -- the order depends on the grid only, so the result is reproducible run to run
    __threadfence();
    double a = 0, b = 0;
    for (unsigned g = threadIdx.x; g < gridDim.x; g += KV_THREADS) {
      a += __ldcg(sc.partials + 2 * g);
      b += __ldcg(sc.partials + 2 * g + 1);
    }
    a = warp_sum(a);
    b = warp_sum(b);
    __syncthreads();  // sh0 / sh1 were read by thread 0 above
    if (lane == 0) { sh0[warp] = a; sh1[warp] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
      a = 0;
      b = 0;
      for (int w = 0; w < KV_THREADS / 32; ++w) { a += sh0[w]; b += sh1[w]; }
      sc.out[0] = a;
      sc.out[1] = b;
      *sc.ticket = 0u;
    }
  }
}

// out[0] = a.b, out[1] = c.d (c == nullptr: 0)
template <typename T>
__global__ void __launch_bounds__(KV_THREADS)
vec_dot2_kernel(const T *__restrict__ a, const T *__restrict__ b, const T *__restrict__ c, const T *__restrict__ d,
                int64_t n, DotScratch sc) {
  double s0 = 0, s1 = 0;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(KV_THREADS) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * KV_THREADS) {
    s0 += static_cast<double>(a[i]) * static_cast<double>(b[i]);
    if (c) s1 += static_cast<double>(c[i]) * static_cast<double>(d[i]);
  }
  block_dot_finish<T>(s0, s1, sc);
}

struct LinComb {
  const void *x[4];
  double c[4];
  int k;
};

template <typename T>
__global__ void __launch_bounds__(KV_THREADS) vec_lincomb_kernel(T *out, int64_t n, LinComb lc) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(KV_THREADS) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * KV_THREADS) {
    T acc = T(0);
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < lc.k) acc += static_cast<T>(lc.c[j]) * static_cast<const T *>(lc.x[j])[i];
    out[i] = acc;
  }
}

// x += alpha p;  r -= alpha ap;  out[0] = r.r
template <typename T>
__global__ void __launch_bounds__(KV_THREADS)
vec_cg_step_kernel(T *x, T *r, const T *__restrict__ p, const T *__restrict__ ap, T alpha, int64_t n, DotScratch sc) {
  double s0 = 0;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(KV_THREADS) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * KV_THREADS) {
    x[i] += alpha * p[i];
    const T ri = r[i] - alpha * ap[i];
    r[i] = ri;
    s0 += static_cast<double>(ri) * static_cast<double>(ri);
  }
  block_dot_finish<T>(s0, 0.0, sc);
}

int kv_grid(int64_t n) {
  const int cap = device_props().sm_count * 8;
  return static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(cap, (n + KV_THREADS * 4 - 1) / (KV_THREADS * 4))));
}

// per-device scratch of the dot kernels (partials for the largest grid, ticket, result); calls on one stream serialise
struct KvScratch {
  DeviceBuffer buf;
  DotScratch sc{};
};
KvScratch &kv_scratch() {
  static thread_local KvScratch s[16];
  int dev = 0;
  QDSP_CUDA(cudaGetDevice(&dev));
  KvScratch &k = s[dev & 15];
  if (k.buf.bytes() == 0) {
    const size_t grid = static_cast<size_t>(device_props().sm_count) * 8;
    k.buf.alloc(sizeof(double) * 2 * grid + 64);
    QDSP_CUDA(cudaMemset(k.buf.get(), 0, k.buf.bytes()));
    k.sc.partials = k.buf.get<double>();
    k.sc.out = k.sc.partials + 2 * grid;
    k.sc.ticket = reinterpret_cast<unsigned *>(k.sc.out + 2);
  }
  return k;
}

void fetch(const DotScratch &sc, double *out, int count, cudaStream_t stream) {
  double h[2];
  QDSP_CUDA(cudaMemcpyAsync(h, sc.out, sizeof(h), cudaMemcpyDeviceToHost, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  for (int i = 0; i < count; ++i) out[i] = h[i];
}
}  // namespace

void vec_dot2(const void *a, const void *b, const void *c, const void *d, int64_t n, int dtype, cudaStream_t stream,
              double *out) {
  QDSP_REQUIRE(a && b && out && n >= 0, QDSP_ERR_INVALID, "vec_dot2: null argument");
  QDSP_REQUIRE((c == nullptr) == (d == nullptr), QDSP_ERR_INVALID, "vec_dot2: c and d go together");
  KvScratch &k = kv_scratch();
  if (dtype == QDSP_F64)
    vec_dot2_kernel<double><<<kv_grid(n), KV_THREADS, 0, stream>>>(static_cast<const double *>(a), static_cast<const double *>(b),
                                                                  static_cast<const double *>(c), static_cast<const double *>(d), n, k.sc);
  else
    vec_dot2_kernel<float><<<kv_grid(n), KV_THREADS, 0, stream>>>(static_cast<const float *>(a), static_cast<const float *>(b),
                                                                 static_cast<const float *>(c), static_cast<const float *>(d), n, k.sc);
  QDSP_LAUNCH_CHECK();
  fetch(k.sc, out, c ? 2 : 1, stream);
}

void vec_lincomb(void *out, int64_t n, int dtype, int k, const void *const *xs, const double *coefs, cudaStream_t stream) {
  QDSP_REQUIRE(out && xs && coefs && k >= 1 && k <= 4 && n >= 0, QDSP_ERR_INVALID, "vec_lincomb: 1 to 4 terms");
  LinComb lc{};
  lc.k = k;
  for (int i = 0; i < k; ++i) {
    QDSP_REQUIRE(xs[i] != nullptr, QDSP_ERR_INVALID, "vec_lincomb: null term");
    lc.x[i] = xs[i];
    lc.c[i] = coefs[i];
  }
  if (n == 0) return;
  if (dtype == QDSP_F64) vec_lincomb_kernel<double><<<kv_grid(n), KV_THREADS, 0, stream>>>(static_cast<double *>(out), n, lc);
  else vec_lincomb_kernel<float><<<kv_grid(n), KV_THREADS, 0, stream>>>(static_cast<float *>(out), n, lc);
  QDSP_LAUNCH_CHECK();
}

void vec_cg_step(void *x, void *r, const void *p, const void *ap, double alpha, int64_t n, int dtype, cudaStream_t stream,
                 double *rr) {
  QDSP_REQUIRE(x && r && p && ap && rr && n >= 0, QDSP_ERR_INVALID, "vec_cg_step: null argument");
  KvScratch &k = kv_scratch();
  if (dtype == QDSP_F64)
    vec_cg_step_kernel<double><<<kv_grid(n), KV_THREADS, 0, stream>>>(static_cast<double *>(x), static_cast<double *>(r),
                                                                     static_cast<const double *>(p), static_cast<const double *>(ap), alpha, n, k.sc);
  else
    vec_cg_step_kernel<float><<<kv_grid(n), KV_THREADS, 0, stream>>>(static_cast<float *>(x), static_cast<float *>(r),
                                                                    static_cast<const float *>(p), static_cast<const float *>(ap),
                                                                    static_cast<float>(alpha), n, k.sc);
  QDSP_LAUNCH_CHECK();
  fetch(k.sc, rr, 1, stream);
}

}  // namespace qdsp
