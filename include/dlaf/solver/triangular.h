// dlaf::triangular_solver<Backend, Device, T>: public C++ entry of the triangular solver on local memory, signature
// and preconditions as in the reference (include/dlaf/solver/triangular.h:25-101): solves op(A) X = alpha B
// (side == Left) or X op(A) = alpha B (side == Right), X overwriting B. Asynchronous like the reference: the call
// enqueues the solve on mat_b's stream (after the work pending on mat_a) and returns; mat_b.waitLocalTiles() blocks.
// The distributed overload (triangular.h:103-178) is collective over the grid: every rank calls it with its local
// blocks; the NCCL broadcasts / reductions along the grid's rows and columns are enqueued on mat_b's stream.
#pragma once
#include <memory>

#include "dlaf/communication/communicator_grid.h"
#include "dlaf/matrix/matrix.h"
#include "dlafb_solver.h"

namespace dlaf {

/// @param mat_a triangular matrix: only the tiles of the uplo triangle are read (not the diagonal elements when
/// diag == Unit). @pre not distributed, square, size (M x M) for side == Left or (N x N) for side == Right, square
/// blocks of the size of mat_b's row blocks (Left) or column blocks (Right).
/// @param mat_b (M x N): B on entry, X on exit. @pre not distributed.
template <Backend backend, Device device, class T>
void triangular_solver(blas::Side side, blas::Uplo uplo, blas::Op op, blas::Diag diag, T alpha,
                       const Matrix<T, device>& mat_a, Matrix<T, device>& mat_b) {
  static_assert(backend == Backend::GPU && device == Device::GPU,
                "this build has no CPU backend: use Backend::GPU with Device::GPU matrices");
  DLAF_ASSERT(mat_a.size().rows() == mat_a.size().cols(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size().rows() == mat_a.block_size().cols(), mat_a.block_size());
  DLAF_ASSERT(mat_a.grid_size() == comm::Size2D(1, 1), mat_a.grid_size());
  DLAF_ASSERT(mat_b.grid_size() == comm::Size2D(1, 1), mat_b.grid_size());
  const SizeType m = mat_b.size().rows(), n = mat_b.size().cols();
  if (side == blas::Side::Left) {
    DLAF_ASSERT(mat_a.size().rows() == m, mat_a.size());
    DLAF_ASSERT(mat_a.block_size().rows() == mat_b.block_size().rows(), mat_a.block_size());
  }
  else {
    DLAF_ASSERT(mat_a.size().rows() == n, mat_a.size());
    DLAF_ASSERT(mat_a.block_size().rows() == mat_b.block_size().cols(), mat_a.block_size());
  }
  dlafb_trsolve_plan_t plan = nullptr;
  int e = dlafb_triangular_solver_plan_create(&plan, dlaf::internal::TypeCode<T>::value, static_cast<char>(side),
                                              static_cast<char>(uplo), static_cast<char>(op), static_cast<char>(diag),
                                              m, n, mat_a.block_size().rows());
  DLAF_ASSERT(e == 0, e);
  std::shared_ptr<void> keep(plan, [](void* p) {
    dlafb_triangular_solver_plan_destroy(static_cast<dlafb_trsolve_plan_t>(p));
  });
  // order the solve after the work pending on A, and later work on A after the solve
  cudaEvent_t ev;
  cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventRecord(ev, mat_a.stream());
  cudaStreamWaitEvent(mat_b.stream(), ev, 0);
  e = dlafb_triangular_solver_plan_run(plan, &alpha, mat_a.ptr(), mat_a.ld(), mat_b.ptr(), mat_b.ld(), mat_b.stream());
  DLAF_ASSERT(e == 0, e);
  cudaEventRecord(ev, mat_b.stream());
  cudaStreamWaitEvent(mat_a.stream(), ev, 0);
  cudaEventDestroy(ev);
  mat_b.keep_alive(keep);
}

/// Distributed on @p grid. @pre mat_a and mat_b are distributed on grid with the same source rank; mat_a square with
/// square blocks of the size of mat_b's row blocks (Left) or column blocks (Right).
template <Backend backend, Device device, class T>
void triangular_solver(comm::CommunicatorGrid& grid, blas::Side side, blas::Uplo uplo, blas::Op op, blas::Diag diag,
                       T alpha, const Matrix<T, device>& mat_a, Matrix<T, device>& mat_b) {
  static_assert(backend == Backend::GPU && device == Device::GPU,
                "this build has no CPU backend: use Backend::GPU with Device::GPU matrices");
  DLAF_ASSERT(mat_a.size().rows() == mat_a.size().cols(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size().rows() == mat_a.block_size().cols(), mat_a.block_size());
  DLAF_ASSERT(mat_a.grid_size() == grid.size() && mat_a.rank_index() == grid.rank(), mat_a.grid_size());
  DLAF_ASSERT(mat_b.grid_size() == grid.size() && mat_b.rank_index() == grid.rank(), mat_b.grid_size());
  DLAF_ASSERT(mat_a.distribution().source_rank_index() == mat_b.distribution().source_rank_index(),
              mat_a.distribution().source_rank_index());
  const SizeType m = mat_b.size().rows(), n = mat_b.size().cols();
  if (side == blas::Side::Left) {
    DLAF_ASSERT(mat_a.size().rows() == m, mat_a.size());
    DLAF_ASSERT(mat_a.block_size().rows() == mat_b.block_size().rows(), mat_a.block_size());
  }
  else {
    DLAF_ASSERT(mat_a.size().rows() == n, mat_a.size());
    DLAF_ASSERT(mat_a.block_size().rows() == mat_b.block_size().cols(), mat_a.block_size());
  }
  const comm::Index2D src = mat_a.distribution().source_rank_index();
  dlafb_trsolve_dist_plan_t plan = nullptr;
  int e = dlafb_triangular_solver_dist_plan_create(&plan, grid.handle(), dlaf::internal::TypeCode<T>::value,
                                                   static_cast<char>(side), static_cast<char>(uplo),
                                                   static_cast<char>(op), static_cast<char>(diag), m, n,
                                                   mat_b.block_size().rows(), mat_b.block_size().cols(), src.row(),
                                                   src.col());
  DLAF_ASSERT(e == 0, e);
  std::shared_ptr<void> keep(plan, [](void* p) {
    dlafb_triangular_solver_dist_plan_destroy(static_cast<dlafb_trsolve_dist_plan_t>(p));
  });
  cudaEvent_t ev;
  cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventRecord(ev, mat_a.stream());
  cudaStreamWaitEvent(mat_b.stream(), ev, 0);
  e = dlafb_triangular_solver_dist_plan_run(plan, &alpha, mat_a.ptr(), mat_a.ld(), mat_b.ptr(), mat_b.ld(),
                                            mat_b.stream());
  DLAF_ASSERT(e == 0, e);
  cudaEventRecord(ev, mat_b.stream());
  cudaStreamWaitEvent(mat_a.stream(), ev, 0);
  cudaEventDestroy(ev);
  mat_b.keep_alive(keep);
}

}  // namespace dlaf
