// dlaf::eigensolver::internal::generalized_to_standard<Backend, Device, T>: public C++ entry of the reduction of a
// Hermitian definite generalized eigenproblem to standard form on local memory, signature and preconditions as in the
// reference (include/dlaf/eigensolver/gen_to_std.h:23-77): mat_a <- inv(B) A inv(B)^H (uplo == Lower) or
// inv(B)^H A inv(B) (uplo == Upper), B being the Cholesky factor stored in mat_b. Asynchronous like the reference: the
// call enqueues the reduction on mat_a's stream (after the work pending on mat_b) and returns;
// mat_a.waitLocalTiles() blocks. mat_b is never modified (the reference may modify its diagonal tiles temporarily).
// The distributed overload (gen_to_std.h:79-140) is collective over the grid: every rank calls it with its local blocks.
#pragma once
#include <memory>

#include "dlaf/communication/communicator_grid.h"
#include "dlaf/matrix/matrix.h"
#include "dlafb_gen_to_std.h"

namespace dlaf::eigensolver::internal {

/// @param mat_a Hermitian matrix, only the tiles of the uplo triangle are read and written.
/// @param mat_b Cholesky factor (Lower: L, Upper: U), only the tiles of the uplo triangle are read.
/// @pre both not distributed, square, same size (N x N), same square block size (NB x NB).
template <Backend backend, Device device, class T>
void generalized_to_standard(blas::Uplo uplo, Matrix<T, device>& mat_a, Matrix<T, device>& mat_b) {
  static_assert(backend == Backend::GPU && device == Device::GPU,
                "this build has no CPU backend: use Backend::GPU with Device::GPU matrices");
  DLAF_ASSERT(mat_a.size().rows() == mat_a.size().cols(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size().rows() == mat_a.block_size().cols(), mat_a.block_size());
  DLAF_ASSERT(mat_b.size().rows() == mat_b.size().cols(), mat_b.size());
  DLAF_ASSERT(mat_b.block_size().rows() == mat_b.block_size().cols(), mat_b.block_size());
  DLAF_ASSERT(mat_a.size() == mat_b.size(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size() == mat_b.block_size(), mat_a.block_size());
  DLAF_ASSERT(mat_a.grid_size() == comm::Size2D(1, 1), mat_a.grid_size());
  DLAF_ASSERT(mat_b.grid_size() == comm::Size2D(1, 1), mat_b.grid_size());
  DLAF_ASSERT(uplo == blas::Uplo::Lower || uplo == blas::Uplo::Upper, static_cast<char>(uplo));
  dlafb_gen_to_std_plan_t plan = nullptr;
  int e = dlafb_gen_to_std_plan_create(&plan, dlaf::internal::TypeCode<T>::value, static_cast<char>(uplo),
                                       mat_a.size().rows(), mat_a.block_size().rows());
  DLAF_ASSERT(e == 0, e);
  std::shared_ptr<void> keep(plan, [](void* p) {
    dlafb_gen_to_std_plan_destroy(static_cast<dlafb_gen_to_std_plan_t>(p));
  });
  // order the reduction after the work pending on the factor, and later work on the factor after the reduction
  cudaEvent_t ev;
  cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventRecord(ev, mat_b.stream());
  cudaStreamWaitEvent(mat_a.stream(), ev, 0);
  e = dlafb_gen_to_std_plan_run(plan, mat_a.ptr(), mat_a.ld(), mat_b.ptr(), mat_b.ld(), mat_a.stream());
  DLAF_ASSERT(e == 0, e);
  cudaEventRecord(ev, mat_a.stream());
  cudaStreamWaitEvent(mat_b.stream(), ev, 0);
  cudaEventDestroy(ev);
  mat_a.keep_alive(keep);
}

/// Distributed on @p grid. @pre mat_a and mat_b are distributed on grid with the same source rank, square, same size
/// and the same square block size.
template <Backend backend, Device device, class T>
void generalized_to_standard(comm::CommunicatorGrid& grid, blas::Uplo uplo, Matrix<T, device>& mat_a,
                             Matrix<T, device>& mat_b) {
  static_assert(backend == Backend::GPU && device == Device::GPU,
                "this build has no CPU backend: use Backend::GPU with Device::GPU matrices");
  DLAF_ASSERT(mat_a.size().rows() == mat_a.size().cols(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size().rows() == mat_a.block_size().cols(), mat_a.block_size());
  DLAF_ASSERT(mat_a.size() == mat_b.size(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size() == mat_b.block_size(), mat_a.block_size());
  DLAF_ASSERT(mat_a.grid_size() == grid.size() && mat_a.rank_index() == grid.rank(), mat_a.grid_size());
  DLAF_ASSERT(mat_b.grid_size() == grid.size() && mat_b.rank_index() == grid.rank(), mat_b.grid_size());
  DLAF_ASSERT(mat_a.distribution().source_rank_index() == mat_b.distribution().source_rank_index(),
              mat_a.distribution().source_rank_index());
  DLAF_ASSERT(uplo == blas::Uplo::Lower || uplo == blas::Uplo::Upper, static_cast<char>(uplo));
  const comm::Index2D src = mat_a.distribution().source_rank_index();
  dlafb_gen_to_std_dist_plan_t plan = nullptr;
  int e = dlafb_gen_to_std_dist_plan_create(&plan, grid.handle(), dlaf::internal::TypeCode<T>::value,
                                            static_cast<char>(uplo), mat_a.size().rows(), mat_a.block_size().rows(),
                                            src.row(), src.col());
  DLAF_ASSERT(e == 0, e);
  std::shared_ptr<void> keep(plan, [](void* p) {
    dlafb_gen_to_std_dist_plan_destroy(static_cast<dlafb_gen_to_std_dist_plan_t>(p));
  });
  cudaEvent_t ev;
  cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventRecord(ev, mat_b.stream());
  cudaStreamWaitEvent(mat_a.stream(), ev, 0);
  e = dlafb_gen_to_std_dist_plan_run(plan, mat_a.ptr(), mat_a.ld(), mat_b.ptr(), mat_b.ld(), mat_a.stream());
  DLAF_ASSERT(e == 0, e);
  cudaEventRecord(ev, mat_a.stream());
  cudaStreamWaitEvent(mat_b.stream(), ev, 0);
  cudaEventDestroy(ev);
  mat_a.keep_alive(keep);
}

}  // namespace dlaf::eigensolver::internal
