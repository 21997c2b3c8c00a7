// dist.hpp -- host side of the multi-GPU path (one process per GPU): row-block partition plan, ghost / send lists, the
// peer-mapped comm buffer and the optional NCCL communicator.  Device protocol: dist_dev.cuh.  SURVEY 8(e).
#pragma once
#include <cstdint>
#include <vector>

#include "dist_dev.cuh"
#include "host.hpp"

namespace qdsp {

// Result of planning one rank's row block: which columns are ghosts, who owns them, and the local renumbering.
struct DistPlan {
  int n_local = 0;
  int ghost_base = 0;                  // local column id of the first ghost (>= n_local; 128-byte aligned on the GPU)
  std::vector<int32_t> ghost_ids;      // global column ids of the ghosts, ascending (hence grouped by owner)
  std::vector<int32_t> owner_offsets;  // world + 1: ghosts owned by rank q are ghost_ids[owner_offsets[q] .. [q+1])
};

// Pure host function (no CUDA): rank `rank` owns global rows/cols [row_offsets[rank], row_offsets[rank+1]).
// `indices` (global column ids of the rank's CSR rows) is rewritten in place to the local numbering:
// owned column c -> c - row_offsets[rank]; ghost -> ghost_base + position in ghost_ids (ghost_base < 0: n_local).
DistPlan dist_plan_host(int rank, int world, const int64_t *row_offsets, int32_t *indices, int64_t nnz,
                        int ghost_base = -1);

class DistContext {
 public:
  DistContext(int rank, int world);
  ~DistContext();
  DistContext(const DistContext &) = delete;
  DistContext &operator=(const DistContext &) = delete;

  int rank() const { return rank_; }
  int world() const { return world_; }
  int n_local() const { return plan_.n_local; }
  int n_ghost() const { return static_cast<int>(plan_.ghost_ids.size()); }
  int ghost_base() const { return plan_.ghost_base; }
  // first element of the z-ghost segment of my vector: ghosts rounded up to a 128-byte boundary
  int zbase() const { return (plan_.ghost_base + n_ghost() + 31) & ~31; }
  int dtype() const { return dtype_; }
  const DistPlan &plan() const { return plan_; }

  // A holds this rank's rows with GLOBAL column ids (cols == global size).  Renumbers its columns in place
  // (A.cols becomes n_local + n_ghost), allocates the comm buffer.
  void localize(CsrMatrix &A, const int64_t *row_offsets, cudaStream_t stream);
  // rank `peer` reads `count` of my rows (global ids, ascending) and wants them at element `dst_offset` of its vector
  // dst_offset_z: where the same rows go in its z-ghost segment (< 0: the peer has none; CG then keeps pushing p)
  void set_send_list(int peer, const int32_t *global_ids, int count, int dst_offset, int dst_offset_z = -1);
  void comm_handle(void *out64) const;
  void attach(const void *handles /* world x 64 bytes, rank order */);
  bool ready() const { return attached_; }
  void *vector_ptr() const { return comm_.get<unsigned char>() + DIST_VECTOR_OFF; }
  DistDev dev() const { return dev_; }
  int error_flag(cudaStream_t stream) const;

  // y_loc = A_loc * [x_loc ; ghosts]: copies x into the operand vector, pushes the halo, waits, multiplies
  void spmv(const CsrMatrix &A, const void *x_loc, void *y_loc, cudaStream_t stream);

  // NCCL (loaded with dlopen, like the reference loads its CUDA libraries: quadrants/rhi/cuda/cuda_driver.cpp:172-258)
  static void nccl_unique_id(void *out128);
  void nccl_init(const void *id128);
  bool has_nccl() const { return nccl_comm_ != nullptr; }
  void nccl_allreduce_double(double *inout, int count, cudaStream_t stream);

 private:
  int rank_, world_;
  int dtype_ = QDSP_F32;
  int64_t row0_ = 0;
  DistPlan plan_;
  DeviceBuffer comm_, ctl_, ghost_blocks_, block_order_;
  std::vector<DeviceBuffer> send_lists_;
  std::vector<void *> opened_;
  DistDev dev_;
  bool attached_ = false;
  void *nccl_comm_ = nullptr;
};

}  // namespace qdsp
