// matrix::copy(src, dst, grid): copy of a distributed matrix performing data reshuffling - `dst` may have another block
// size and another source rank than `src` (reference: include/dlaf/matrix/copy.h:118-231; preconditions :124-131: same
// size, same grid, in each dimension the smaller block size divides the larger one). Device matrices; collective over
// the grid; asynchronous on dst's stream (after the work pending on src). The same-distribution overload
// matrix::copy(src, dst) lives in dlaf/matrix/matrix.h.
#pragma once
#include <memory>

#include "dlaf/communication/communicator_grid.h"
#include "dlaf/matrix/matrix.h"
#include "dlafb_matrix.h"

namespace dlaf::matrix {

template <class T>
void copy(Matrix<T, Device::GPU>& src, Matrix<T, Device::GPU>& dst, comm::CommunicatorGrid& grid) {
  DLAF_ASSERT(src.size() == dst.size(), src.size());
  DLAF_ASSERT(src.grid_size() == grid.size() && src.rank_index() == grid.rank(), src.grid_size());
  DLAF_ASSERT(dst.grid_size() == grid.size() && dst.rank_index() == grid.rank(), dst.grid_size());
  const auto mn = [](SizeType a, SizeType b) { return a < b ? a : b; };
  const SizeType mb = mn(src.block_size().rows(), dst.block_size().rows());
  const SizeType nb = mn(src.block_size().cols(), dst.block_size().cols());
  DLAF_ASSERT(src.block_size().rows() % mb == 0 && dst.block_size().rows() % mb == 0, mb);
  DLAF_ASSERT(src.block_size().cols() % nb == 0 && dst.block_size().cols() % nb == 0, nb);
  const comm::Index2D ss = src.distribution().source_rank_index(), ds = dst.distribution().source_rank_index();
  dlafb_redistribute_plan_t plan = nullptr;
  int e = dlafb_redistribute_plan_create(&plan, grid.handle(), dlaf::internal::TypeCode<T>::value, src.size().rows(),
                                         src.size().cols(), src.block_size().rows(), src.block_size().cols(), ss.row(),
                                         ss.col(), dst.block_size().rows(), dst.block_size().cols(), ds.row(), ds.col());
  DLAF_ASSERT(e == 0, e);
  std::shared_ptr<void> keep(plan, [](void* p) {
    dlafb_redistribute_plan_destroy(static_cast<dlafb_redistribute_plan_t>(p));
  });
  cudaEvent_t ev;
  cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventRecord(ev, src.stream());
  cudaStreamWaitEvent(dst.stream(), ev, 0);
  e = dlafb_redistribute_plan_run(plan, src.ptr(), src.ld(), dst.ptr(), dst.ld(), dst.stream());
  DLAF_ASSERT(e == 0, e);
  cudaEventRecord(ev, dst.stream());
  cudaStreamWaitEvent(src.stream(), ev, 0);
  cudaEventDestroy(ev);
  dst.keep_alive(keep);
}

}  // namespace dlaf::matrix
