// comm::CommunicatorGrid: 2D grid of ranks with row / column / full communicators (reference:
// include/dlaf/communication/communicator_grid.h:37-153, src/communication/communicator_grid.cpp:19-68).
// Here the communicators are NCCL communicators over NVLink (ncclCommSplit of the world communicator); collectives
// are ordered by the CUDA streams they are enqueued on, which is what the reference's round-robin
// CommunicatorPipelines (`npipelines`) provide for MPI - the argument is accepted for source compatibility.
#pragma once
#include <array>
#include <utility>

#include "dlaf/communication/communicator.h"
#include "dlaf/communication/communicator_pipeline.h"
#include "dlafb_engine.h"

// internal hooks of libdlaf_b200 (src/c_api.cu): grid over the MPI-shim communicator
extern "C" int dlafb_grid_create_from_comm(dlafb_grid_t* grid, MPI_Comm comm, int nprow, int npcol, char order);

namespace dlaf::comm {

class CommunicatorGrid {
public:
  CommunicatorGrid(Communicator comm, IndexT_MPI rows, IndexT_MPI cols, common::Ordering ordering,
                   std::size_t npipelines = 3)
      : full_(comm), ordering_(ordering), npipelines_(npipelines) {
    DLAF_ASSERT(rows * cols == comm.size(), rows, cols);
    const int e = dlafb_grid_create_from_comm(&grid_, comm, rows, cols, ordering == common::Ordering::RowMajor ? 'R' : 'C');
    DLAF_ASSERT(e == 0, e);
    int P, Q, pr, pc, wr;
    char o;
    dlafb_grid_info(grid_, &P, &Q, &pr, &pc, &wr, &o);
    size_ = Size2D(P, Q);
    rank_ = Index2D(pr, pc);
  }
  CommunicatorGrid(Communicator comm, const std::array<IndexT_MPI, 2>& size, common::Ordering ordering,
                   std::size_t npipelines = 3)
      : CommunicatorGrid(comm, size[0], size[1], ordering, npipelines) {}

  CommunicatorGrid(CommunicatorGrid&& o) noexcept { *this = std::move(o); }
  CommunicatorGrid& operator=(CommunicatorGrid&& o) noexcept {
    std::swap(grid_, o.grid_);
    full_ = o.full_;
    size_ = o.size_;
    rank_ = o.rank_;
    ordering_ = o.ordering_;
    npipelines_ = o.npipelines_;
    return *this;
  }
  CommunicatorGrid(const CommunicatorGrid&) = delete;
  CommunicatorGrid& operator=(const CommunicatorGrid&) = delete;
  ~CommunicatorGrid() {
    if (grid_)
      dlafb_grid_destroy(grid_);
  }

  /// 2D coordinates of this rank / grid dimensions.
  Index2D rank() const noexcept { return rank_; }
  Size2D size() const noexcept { return size_; }
  /// Rank in the full communicator: row-major linearisation as rankFullCommunicator does
  /// (include/dlaf/communication/communicator_pipeline.h:26).
  IndexT_MPI rankFullCommunicator(const Index2D& index) const noexcept { return index.row() * size_.cols() + index.col(); }
  IndexT_MPI rank_full_communicator(const Index2D& index) const noexcept { return rankFullCommunicator(index); }
  Communicator& fullCommunicator() noexcept { return full_; }
  Communicator& full_communicator() noexcept { return full_; }
  common::Ordering ordering() const noexcept { return ordering_; }
  std::size_t num_pipelines() const noexcept { return npipelines_; }

  /// Pipelines along the row / column / full communicator (communicator_grid.h:110-131 of the reference): see
  /// communicator_pipeline.h for what they carry in this build.
  CommunicatorPipeline<CommunicatorType::Row> row_communicator_pipeline() { return {grid_, rank_, size_}; }
  CommunicatorPipeline<CommunicatorType::Col> col_communicator_pipeline() { return {grid_, rank_, size_}; }
  CommunicatorPipeline<CommunicatorType::Full> full_communicator_pipeline() { return {grid_, rank_, size_}; }

  /// Grid-wide barrier on the device communicator.
  void barrier() const { dlafb_grid_barrier(grid_); }
  /// Handle for the engine ABI.
  dlafb_grid_t handle() const noexcept { return grid_; }

private:
  dlafb_grid_t grid_ = nullptr;
  Communicator full_;
  Size2D size_{1, 1};
  Index2D rank_{0, 0};
  common::Ordering ordering_ = common::Ordering::RowMajor;
  std::size_t npipelines_ = 3;
};

}  // namespace dlaf::comm
