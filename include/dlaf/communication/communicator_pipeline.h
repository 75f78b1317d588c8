// comm::CommunicatorPipeline<CT>: API surface of the reference's communicator pipelines
// (include/dlaf/communication/communicator_pipeline.h:41-148). There a pipeline serialises the tasks that use one
// duplicated MPI communicator; here collectives are enqueued on CUDA streams in program order by every rank (grid.cu,
// dist_device_ops.h), which is the same ordering guarantee without a queue - so the object only carries what callers
// query: the rank / size along its axis, the 2D coordinates, and the row-major rank in the full communicator.
// shared() / exclusive() return the underlying grid handle: the access mode is a no-op in this design.
#pragma once
#include <ostream>

#include "dlaf/common/index2d.h"
#include "dlafb_engine.h"

namespace dlaf::comm {

enum class CommunicatorType { Row, Col, Full };

template <CommunicatorType CT>
class CommunicatorPipeline {
public:
  static constexpr CommunicatorType communicator_type = CT;

  CommunicatorPipeline(dlafb_grid_t grid, Index2D rank_2d, Size2D size_2d) : grid_(grid), rank_2d_(rank_2d), size_2d_(size_2d) {
    if (CT == CommunicatorType::Row) {  // the ranks of my grid row, ordered by column
      rank_ = rank_2d.col();
      size_ = size_2d.cols();
    }
    else if (CT == CommunicatorType::Col) {
      rank_ = rank_2d.row();
      size_ = size_2d.rows();
    }
    else {
      rank_ = rank_2d.row() * size_2d.cols() + rank_2d.col();
      size_ = size_2d.rows() * size_2d.cols();
    }
  }
  CommunicatorPipeline(CommunicatorPipeline&&) = default;
  CommunicatorPipeline& operator=(CommunicatorPipeline&&) = default;
  CommunicatorPipeline(const CommunicatorPipeline&) = delete;
  CommunicatorPipeline& operator=(const CommunicatorPipeline&) = delete;

  IndexT_MPI rank() const noexcept { return rank_; }
  IndexT_MPI size() const noexcept { return size_; }
  Index2D rank_2d() const noexcept { return rank_2d_; }
  Size2D size_2d() const noexcept { return size_2d_; }
  /// Row-major rank in the full communicator (communicator_pipeline.h:26, :83-86).
  IndexT_MPI rank_full_communicator(const Index2D& index) const noexcept {
    return index.row() * size_2d_.cols() + index.col();
  }
  dlafb_grid_t exclusive() noexcept { return grid_; }
  dlafb_grid_t shared() noexcept { return grid_; }
  CommunicatorPipeline sub_pipeline() { return CommunicatorPipeline(grid_, rank_2d_, size_2d_); }
  bool valid() const noexcept { return grid_ != nullptr; }
  void reset() noexcept { grid_ = nullptr; }

  friend std::ostream& operator<<(std::ostream& out, const CommunicatorPipeline& p) {
    return out << "2d rank=" << p.rank_2d_ << ", 2d size=" << p.size_2d_ << ", rank=" << p.rank_ << ", size=" << p.size_;
  }

private:
  dlafb_grid_t grid_ = nullptr;
  IndexT_MPI rank_ = -1, size_ = -1;
  Index2D rank_2d_{-1, -1};
  Size2D size_2d_{-1, -1};
};

}  // namespace dlaf::comm
