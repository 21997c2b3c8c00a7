// dist.cu -- host side of the multi-GPU path: partition plan, peer mapping (CUDA IPC), halo push kernel, NCCL via dlopen.
// The reference is single-device (quadrants/rhi/cuda/cuda_context.cpp:16-26); this layer is new.
#include <dlfcn.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "dist.hpp"
#include "spmv_kernels.cuh"

namespace qdsp {

// ---- plan (pure host) ------------------------------------------------------------------------------------------------
DistPlan dist_plan_host(int rank, int world, const int64_t *row_offsets, int32_t *indices, int64_t nnz,
                        int ghost_base) {
  QDSP_REQUIRE(world >= 1 && world <= DIST_MAX_RANKS && rank >= 0 && rank < world, QDSP_ERR_INVALID,
               "dist plan: rank/world out of range (1..8 ranks)");
  QDSP_REQUIRE(row_offsets != nullptr && (nnz == 0 || indices != nullptr), QDSP_ERR_INVALID, "dist plan: null array");
  for (int q = 0; q < world; ++q)
    QDSP_REQUIRE(row_offsets[q] <= row_offsets[q + 1], QDSP_ERR_INVALID, "dist plan: row offsets must ascend");
  QDSP_REQUIRE(row_offsets[0] == 0 && row_offsets[world] <= 2147483647LL, QDSP_ERR_INVALID,
               "dist plan: global size must fit int32");
  const int64_t lo = row_offsets[rank], hi = row_offsets[rank + 1], total = row_offsets[world];
  DistPlan plan;
  plan.n_local = static_cast<int>(hi - lo);
  plan.ghost_base = ghost_base < 0 ? plan.n_local : ghost_base;
  QDSP_REQUIRE(plan.ghost_base >= plan.n_local, QDSP_ERR_INVALID, "dist plan: ghost_base below n_local");
  std::vector<int32_t> ghosts;
  for (int64_t k = 0; k < nnz; ++k) {
    const int64_t c = indices[k];
    QDSP_REQUIRE(c >= 0 && c < total, QDSP_ERR_INVALID, "dist plan: column index outside the global matrix");
    if (c < lo || c >= hi) ghosts.push_back(static_cast<int32_t>(c));
  }
  std::sort(ghosts.begin(), ghosts.end());
  ghosts.erase(std::unique(ghosts.begin(), ghosts.end()), ghosts.end());
  // Dense owners: when this rank reads at least half of an owner's rows (hub-dominated graphs: FEM meshes with
  // power-law vertex degrees read almost all of x on every rank) it takes ALL of them.  The owner's send list is then
  // one contiguous run, which the halo push moves with 128-bit peer stores instead of one scattered 4/8-byte store per
  // entry (set_send_list: send_contig), and the ghost slot of that owner is simply its slice of x.
  {
    std::vector<int32_t> dense;
    dense.reserve(ghosts.size());
    size_t at = 0;
    for (int q = 0; q < world; ++q) {
      const int64_t q0 = row_offsets[q], q1 = row_offsets[q + 1];
      size_t end = at;
      while (end < ghosts.size() && ghosts[end] < q1) ++end;
      const int64_t have = static_cast<int64_t>(end - at);
      if (q != rank && have > 0 && 2 * have >= q1 - q0) {
        for (int64_t g = q0; g < q1; ++g) dense.push_back(static_cast<int32_t>(g));
      } else {
        dense.insert(dense.end(), ghosts.begin() + static_cast<std::ptrdiff_t>(at), ghosts.begin() + static_cast<std::ptrdiff_t>(end));
      }
      at = end;
    }
    ghosts.swap(dense);
  }
  plan.owner_offsets.assign(world + 1, 0);
  for (int q = 0; q <= world; ++q)
    plan.owner_offsets[q] = static_cast<int32_t>(
        std::lower_bound(ghosts.begin(), ghosts.end(), q == world ? total : row_offsets[q],
                         [](int32_t g, int64_t v) { return static_cast<int64_t>(g) < v; }) - ghosts.begin());
  for (int64_t k = 0; k < nnz; ++k) {
    const int64_t c = indices[k];
    if (c >= lo && c < hi) indices[k] = static_cast<int32_t>(c - lo);
    else indices[k] = plan.ghost_base + static_cast<int32_t>(std::lower_bound(ghosts.begin(), ghosts.end(), static_cast<int32_t>(c)) - ghosts.begin());
  }
  plan.ghost_ids = std::move(ghosts);
  return plan;
}

// ---- context ---------------------------------------------------------------------------------------------------------
DistContext::DistContext(int rank, int world) : rank_(rank), world_(world) {
  QDSP_REQUIRE(world >= 1 && world <= DIST_MAX_RANKS && rank >= 0 && rank < world, QDSP_ERR_INVALID,
               "dist: rank/world out of range (1..8 ranks)");
  dev_.rank = rank;
  dev_.world = world;
  send_lists_.resize(DIST_MAX_RANKS);
}

DistContext::~DistContext() {
  for (void *p : opened_)
    if (p) cudaIpcCloseMemHandle(p);
  if (nccl_comm_) {
    // ncclCommAbort frees the communicator locally without waiting for the other ranks (ncclCommDestroy may block when
    // this runs at interpreter shutdown, after the peers have gone)
    using abort_t = int (*)(void *);
    if (void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD))
      if (auto fn = reinterpret_cast<abort_t>(dlsym(h, "ncclCommAbort"))) fn(nccl_comm_);
  }
}

void DistContext::localize(CsrMatrix &A, const int64_t *row_offsets, cudaStream_t stream) {
  QDSP_REQUIRE(!attached_, QDSP_ERR_STATE, "dist: already attached");
  QDSP_REQUIRE(row_offsets != nullptr, QDSP_ERR_INVALID, "dist: null row offsets");
  QDSP_REQUIRE(A.rows == row_offsets[rank_ + 1] - row_offsets[rank_], QDSP_ERR_INVALID,
               "dist: the matrix must hold exactly this rank's rows");
  QDSP_REQUIRE(A.cols == row_offsets[world_], QDSP_ERR_INVALID, "dist: the matrix must use global column ids");
  std::vector<int32_t> idx(static_cast<size_t>(A.nnz));
  if (A.nnz) {
    QDSP_CUDA(cudaMemcpyAsync(idx.data(), A.indices.get(), sizeof(int32_t) * idx.size(), cudaMemcpyDeviceToHost, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
  }
  // ghosts start on a 128-byte boundary of the operand vector (fp32: 32 elements; also fine for fp64)
  const int n_loc = static_cast<int>(row_offsets[rank_ + 1] - row_offsets[rank_]);
  plan_ = dist_plan_host(rank_, world_, row_offsets, idx.data(), A.nnz, (n_loc + 31) & ~31);
  if (A.nnz) {
    QDSP_CUDA(cudaMemcpyAsync(A.indices.get(), idx.data(), sizeof(int32_t) * idx.size(), cudaMemcpyHostToDevice, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
  }
  A.cols = plan_.ghost_base + n_ghost();
  A.spmv_mode = 0;  // re-plan: the renumbering does not change row lengths, but a plan made earlier is dropped anyway
  spmv_prepare(A, stream);
  dtype_ = A.dtype;
  row0_ = row_offsets[rank_];
  // which row blocks of the SpMV kernel read a ghost column (uniform path only; the stream path waits up front)
  dev_.ghost_blocks = nullptr;
  dev_.ghost_block_rows = 0;
  dev_.block_order = nullptr;
  dev_.block_order_all = nullptr;
  dev_.n_interior = 0;
  if (A.spmv_mode == 1) {
    const int R = SPMV_CONSUMERS / A.spmv_lpr;
    const int nblk = std::max(1, (A.rows + R - 1) / R);
    std::vector<int32_t> ip(static_cast<size_t>(A.rows) + 1);
    QDSP_CUDA(cudaMemcpyAsync(ip.data(), A.indptr.get(), sizeof(int32_t) * ip.size(), cudaMemcpyDeviceToHost, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
    std::vector<unsigned char> flags(nblk, 0);
    for (int b = 0; b < nblk; ++b) {
      const int r1 = std::min(A.rows, (b + 1) * R);
      for (int64_t k = ip[static_cast<size_t>(b) * R]; k < ip[r1]; ++k)
        if (idx[k] >= plan_.ghost_base) { flags[b] = 1; break; }
    }
    ghost_blocks_.alloc(flags.size());
    QDSP_CUDA(cudaMemcpyAsync(ghost_blocks_.get(), flags.data(), flags.size(), cudaMemcpyHostToDevice, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
    dev_.ghost_blocks = ghost_blocks_.get<unsigned char>();
    dev_.ghost_block_rows = R;
    // work order of the SpMV: interior blocks (ascending), then the blocks that need the halo
    std::vector<int32_t> order;
    order.reserve(nblk);
    for (int b = 0; b < nblk; ++b)
      if (!flags[b]) order.push_back(b);
    dev_.n_interior = static_cast<int>(order.size());
    for (int b = 0; b < nblk; ++b)
      if (flags[b]) order.push_back(b);
    block_order_.alloc(sizeof(int32_t) * order.size());
    QDSP_CUDA(cudaMemcpyAsync(block_order_.get(), order.data(), sizeof(int32_t) * order.size(), cudaMemcpyHostToDevice, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
    dev_.block_order = block_order_.get<int>();
    dev_.block_order_all = dev_.block_order;
    // used only when QDSP_HALO_ORDER=1: with the halo signalled early by its owner CTAs the natural order measured
    // equal or slightly faster at 2, 4 and 8 ranks
    const char *e = std::getenv("QDSP_HALO_ORDER");
    if (!(e && e[0] == '1')) dev_.block_order = nullptr;
  }
  const size_t sv = A.value_bytes();
  comm_.alloc(DIST_VECTOR_OFF + sv * (static_cast<size_t>(zbase()) + n_ghost()) + 256);  // [owned | ghosts | z-ghosts]
  QDSP_CUDA(cudaMemsetAsync(comm_.get(), 0, comm_.bytes(), stream));
  ctl_.alloc(64);
  QDSP_CUDA(cudaMemsetAsync(ctl_.get(), 0, 64, stream));
  QDSP_CUDA(cudaStreamSynchronize(stream));
  dev_.self = comm_.get<unsigned char>();
  dev_.counters = ctl_.get<unsigned>();
  dev_.error = reinterpret_cast<int *>(ctl_.get<unsigned>() + 8);
  dev_.ghost_base = plan_.ghost_base;
  dev_.n_ghost = n_ghost();
  dev_.zbase = zbase();
  dev_.zmode = 1;  // until a neighbour registers without a z slot
  dev_.n_recv = 0;
  for (int q = 0; q < world_; ++q)
    if (plan_.owner_offsets[q + 1] > plan_.owner_offsets[q]) dev_.recv_peer[dev_.n_recv++] = q;
}

void DistContext::set_send_list(int peer, const int32_t *global_ids, int count, int dst_offset, int dst_offset_z) {
  QDSP_REQUIRE(peer >= 0 && peer < world_ && peer != rank_, QDSP_ERR_INVALID, "dist: bad peer");
  QDSP_REQUIRE(count >= 0 && (count == 0 || global_ids != nullptr), QDSP_ERR_INVALID, "dist: bad send list");
  QDSP_REQUIRE(comm_.get() != nullptr, QDSP_ERR_STATE, "dist: localize first");
  if (count == 0) return;
  std::vector<int32_t> local(count);
  for (int i = 0; i < count; ++i) {
    const int64_t l = static_cast<int64_t>(global_ids[i]) - row0_;
    QDSP_REQUIRE(l >= 0 && l < plan_.n_local, QDSP_ERR_INVALID, "dist: a requested row is not owned by this rank");
    QDSP_REQUIRE(i == 0 || global_ids[i] > global_ids[i - 1], QDSP_ERR_INVALID, "dist: send list must ascend");
    local[i] = static_cast<int32_t>(l);
  }
  QDSP_REQUIRE(dev_.n_send < DIST_MAX_RANKS, QDSP_ERR_INVALID, "dist: too many send lists");
  DeviceBuffer &buf = send_lists_[dev_.n_send];
  buf.alloc(sizeof(int32_t) * count);
  QDSP_CUDA(cudaMemcpy(buf.get(), local.data(), sizeof(int32_t) * count, cudaMemcpyHostToDevice));
  const int s = dev_.n_send++;
  dev_.send_peer[s] = peer;
  dev_.send_list[s] = buf.get<int>();
  dev_.send_count[s] = count;
  dev_.send_dst[s] = dst_offset;
  dev_.send_dst_z[s] = dst_offset_z;
  if (dst_offset_z < 0) dev_.zmode = 0;
  dev_.send_first[s] = local.front();
  dev_.send_last[s] = local.back();
  dev_.send_contig[s] = (local.back() - local.front() == count - 1) ? 1 : 0;
}

void DistContext::comm_handle(void *out64) const {
  QDSP_REQUIRE(comm_.get() != nullptr, QDSP_ERR_STATE, "dist: localize first");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  QDSP_CUDA(cudaIpcGetMemHandle(&h, comm_.get()));
  std::memcpy(out64, &h, 64);
}

void DistContext::attach(const void *handles) {
  QDSP_REQUIRE(comm_.get() != nullptr, QDSP_ERR_STATE, "dist: localize first");
  QDSP_REQUIRE(!attached_, QDSP_ERR_STATE, "dist: already attached");
  opened_.assign(world_, nullptr);
  for (int q = 0; q < world_; ++q) {
    if (q == rank_) {
      dev_.peer[q] = dev_.self;
      continue;
    }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const char *>(handles) + 64 * q, 64);
    void *p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess)
      throw Error(QDSP_ERR_COMM, std::string("cudaIpcOpenMemHandle(rank ") + std::to_string(q) + "): " + cudaGetErrorString(e));
    opened_[q] = p;
    dev_.peer[q] = static_cast<unsigned char *>(p);
  }
  attached_ = true;
}

int DistContext::error_flag(cudaStream_t stream) const {
  int e = 0;
  if (dev_.error) {
    QDSP_CUDA(cudaMemcpyAsync(&e, dev_.error, sizeof(int), cudaMemcpyDeviceToHost, stream));
    QDSP_CUDA(cudaStreamSynchronize(stream));
  }
  return e;
}

// ---- stand-alone distributed SpMV ---------------------------------------------------------------------------------------
// Successive calls need a cross-rank rendezvous of their own (CG gets one from its all-reduces): the push of call k+1
// stores into a neighbour's ghost segment, and that neighbour may still be inside the SpMV of call k reading it.  Every
// consumer therefore reports "call k read" to the owners of its ghosts when its SpMV kernel has finished
// (consumed flags, one per reader rank, in the owner's comm buffer), and an owner waits for consumed >= k from every
// rank it sends to before it pushes call k+1.  counters[5] = stand-alone calls started on this rank -- a sequence number
// of its own, because counters[1] (the push count that dist_halo_wait compares the halo flags with) also advances in CG.
__device__ __forceinline__ unsigned long long *dist_flag_consumed(unsigned char *buf, int from_rank) {
  return reinterpret_cast<unsigned long long *>(buf + DIST_FLAGS_OFF + 128) + from_rank;
}

__global__ void dist_wait_consumed_kernel(DistDev d) {
  const unsigned calls = *reinterpret_cast<volatile unsigned *>(&d.counters[5]);
  const int s = threadIdx.x;
  if (s < d.n_send) {
    const unsigned long long *f = dist_flag_consumed(d.self, d.send_peer[s]);
    const long long t0 = clock64();
    while (ld_acquire_sys_u64(f) < calls) {
      if (clock64() - t0 > DIST_SPIN_LIMIT) {
        atomicExch(d.error, 2);
        break;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    d.counters[5] = calls + 1u;
    // a rank that receives but sends nothing has no owner CTA to advance the push count its halo wait compares with
    if (d.n_send == 0) d.counters[1] += 1u;
  }
}

__global__ void dist_signal_consumed_kernel(DistDev d) {
  const int r = threadIdx.x;
  if (r >= d.n_recv) return;
  const unsigned calls = *reinterpret_cast<volatile unsigned *>(&d.counters[5]);
  // the SpMV kernel that read the ghosts ended before this kernel began (stream order)
  st_release_sys_u64(dist_flag_consumed(d.peer[d.recv_peer[r]], d.rank), calls);
}

template <typename T>
__global__ void __launch_bounds__(256) dist_load_push_kernel(int n, const T *__restrict__ x, T *__restrict__ vec, DistDev d) {
  const int per = ((n + gridDim.x - 1) / gridDim.x + 1023) & ~1023;
  const int lo = min(n, static_cast<int>(blockIdx.x) * per), hi = min(n, lo + per);
  for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) vec[i] = x[i];
  dist_halo_push_signal<T>(d, vec, lo, hi, per);
}

void DistContext::spmv(const CsrMatrix &A, const void *x_loc, void *y_loc, cudaStream_t stream) {
  QDSP_REQUIRE(attached_, QDSP_ERR_STATE, "dist: attach the peers first");
  QDSP_REQUIRE(A.rows == plan_.n_local && A.cols == plan_.ghost_base + n_ghost() && A.dtype == dtype_, QDSP_ERR_INVALID,
               "dist spmv: the matrix is not the one localized on this context");
  const int n = plan_.n_local;
  const int grid = std::max(1, std::min(ceil_div(n, 1024), device_props().sm_count * 4));
  dist_wait_consumed_kernel<<<1, 32, 0, stream>>>(dev_);
  QDSP_LAUNCH_CHECK();
  if (dtype_ == QDSP_F32)
    dist_load_push_kernel<float><<<grid, 256, 0, stream>>>(n, static_cast<const float *>(x_loc), static_cast<float *>(vector_ptr()), dev_);
  else
    dist_load_push_kernel<double><<<grid, 256, 0, stream>>>(n, static_cast<const double *>(x_loc), static_cast<double *>(vector_ptr()), dev_);
  QDSP_LAUNCH_CHECK();
  launch_spmv(A, vector_ptr(), y_loc, dev_, stream);
  dist_signal_consumed_kernel<<<1, 32, 0, stream>>>(dev_);
  QDSP_LAUNCH_CHECK();
  QDSP_REQUIRE(error_flag(stream) == 0, QDSP_ERR_COMM, "dist spmv: a peer rank did not answer (halo spin timed out)");
}

// ---- NCCL through dlopen ----------------------------------------------------------------------------------------------------
namespace {
struct Id128 { char bytes[128]; };  // ncclUniqueId (NCCL_UNIQUE_ID_BYTES), passed by value
struct NcclApi {
  void *lib = nullptr;
  int (*GetUniqueId)(Id128 *) = nullptr;
  int (*CommInitRank)(void **, int, Id128, int) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
};
NcclApi &nccl() {
  static NcclApi api;
  if (!api.lib) {
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) throw Error(QDSP_ERR_COMM, std::string("dlopen(libnccl.so.2): ") + dlerror());
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(h, "ncclAllReduce"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.GetErrorString)
      throw Error(QDSP_ERR_COMM, "libnccl.so.2 lacks a required symbol");
    api.lib = h;
  }
  return api;
}
void nccl_check(int rc, const char *what) {
  if (rc != 0) throw Error(QDSP_ERR_COMM, std::string(what) + ": " + nccl().GetErrorString(rc));
}
}  // namespace

void DistContext::nccl_unique_id(void *out128) {
  Id128 id{};
  nccl_check(nccl().GetUniqueId(&id), "ncclGetUniqueId");
  std::memcpy(out128, &id, 128);
}

void DistContext::nccl_init(const void *id128) {
  QDSP_REQUIRE(nccl_comm_ == nullptr, QDSP_ERR_STATE, "dist: NCCL already initialised");
  Id128 id;
  std::memcpy(&id, id128, 128);
  nccl_check(nccl().CommInitRank(&nccl_comm_, world_, id, rank_), "ncclCommInitRank");
}

void DistContext::nccl_allreduce_double(double *inout, int count, cudaStream_t stream) {
  QDSP_REQUIRE(nccl_comm_ != nullptr, QDSP_ERR_STATE, "dist: NCCL not initialised");
  // ncclFloat64 = 8, ncclSum = 0 (nccl.h)
  nccl_check(nccl().AllReduce(inout, inout, static_cast<size_t>(count), 8, 0, nccl_comm_, stream), "ncclAllReduce");
}

}  // namespace qdsp
