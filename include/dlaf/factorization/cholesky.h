// dlaf::cholesky_factorization<Backend, Device, T>: public C++ entry of the Cholesky path, signatures and
// preconditions as in the reference (include/dlaf/factorization/cholesky.h:39-79). Asynchronous like the
// reference: the call enqueues the factorization on the matrix' stream and returns; mat.waitLocalTiles() blocks.
// A tile that is not positive definite terminates the program at the next waitLocalTiles() (the reference traps
// on the device, src/cusolver/assert_info.cu:19-47), printing the ScaLAPACK-style info.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <memory>

#include "dlaf/matrix/matrix.h"
#include "dlafb_engine.h"

namespace dlaf {
namespace factorization::internal {

template <Backend backend, Device device, class T>
struct Cholesky {
  static_assert(backend == Backend::GPU && device == Device::GPU,
                "this build has no CPU backend: use Backend::GPU with a Device::GPU matrix (MatrixMirror copies a "
                "host matrix to the device)");

  static void call(dlafb_grid_t grid, blas::Uplo uplo, Matrix<T, device>& mat_a) {
    const matrix::Distribution& d = mat_a.distribution();
    // the library keeps the plan for (grid, type, uplo, size, block size, source rank) between calls: creating one is a
    // collective with allocations and peer mappings, destroying one synchronises the device
    dlafb_chol_plan_t plan = nullptr;
    int e = dlafb_cholesky_plan_acquire(&plan, grid, dlaf::internal::TypeCode<T>::value, static_cast<char>(uplo),
                                        d.size().rows(), d.block_size().rows(), d.source_rank_index().row(),
                                        d.source_rank_index().col());
    DLAF_ASSERT(e == 0, e);
    std::shared_ptr<void> keep(plan, [](void* p) { dlafb_cholesky_plan_release(static_cast<dlafb_chol_plan_t>(p)); });
    e = dlafb_cholesky_plan_run(plan, mat_a.ptr(), mat_a.ld(), mat_a.stream());
    DLAF_ASSERT(e == 0, e);
    mat_a.keep_alive(keep, [plan] {
      int info = 0;
      dlafb_cholesky_plan_info(plan, &info);
      if (info != 0) {
        std::fprintf(stderr, "[ERROR] cholesky_factorization: the leading minor of order %d is not positive definite\n",
                     info);
        std::abort();
      }
    });
  }
  static void call_L(Matrix<T, device>& mat_a) { call(local_grid(), blas::Uplo::Lower, mat_a); }
  static void call_U(Matrix<T, device>& mat_a) { call(local_grid(), blas::Uplo::Upper, mat_a); }
  static void call_L(comm::CommunicatorGrid& grid, Matrix<T, device>& mat_a) {
    call(grid.handle(), blas::Uplo::Lower, mat_a);
  }
  static void call_U(comm::CommunicatorGrid& grid, Matrix<T, device>& mat_a) {
    call(grid.handle(), blas::Uplo::Upper, mat_a);
  }

private:
  static dlafb_grid_t local_grid() {
    static dlafb_grid_t g = [] {
      dlafb_grid_t h = nullptr;
      const int e = dlafb_grid_create_local(&h);
      DLAF_ASSERT(e == 0, e);
      return h;
    }();
    return g;
  }
};

}  // namespace factorization::internal

/// Local matrix: A = L L^H (uplo = Lower) or A = U^H U (uplo = Upper); only the uplo triangle is referenced and
/// overwritten. @pre square size, square block size, not distributed.
template <Backend backend, Device device, class T>
void cholesky_factorization(blas::Uplo uplo, Matrix<T, device>& mat_a) {
  DLAF_ASSERT(mat_a.size().rows() == mat_a.size().cols(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size().rows() == mat_a.block_size().cols(), mat_a.block_size());
  DLAF_ASSERT(mat_a.grid_size() == comm::Size2D(1, 1), mat_a.grid_size());
  if (uplo == blas::Uplo::Lower)
    factorization::internal::Cholesky<backend, device, T>::call_L(mat_a);
  else
    factorization::internal::Cholesky<backend, device, T>::call_U(mat_a);
}

/// Distributed matrix on @p grid. Collective: every rank of the grid calls it. @pre mat_a is distributed on grid.
template <Backend backend, Device device, class T>
void cholesky_factorization(comm::CommunicatorGrid& grid, blas::Uplo uplo, Matrix<T, device>& mat_a) {
  DLAF_ASSERT(mat_a.size().rows() == mat_a.size().cols(), mat_a.size());
  DLAF_ASSERT(mat_a.block_size().rows() == mat_a.block_size().cols(), mat_a.block_size());
  DLAF_ASSERT(mat_a.grid_size() == grid.size() && mat_a.rank_index() == grid.rank(), mat_a.grid_size());
  if (uplo == blas::Uplo::Lower)
    factorization::internal::Cholesky<backend, device, T>::call_L(grid, mat_a);
  else
    factorization::internal::Cholesky<backend, device, T>::call_U(grid, mat_a);
}

// pre-0.2.0 spelling (CHANGELOG 0.2.0, #942 renamed the algorithms to snake case)
namespace factorization {
template <Backend backend, Device device, class T>
void cholesky(blas::Uplo uplo, Matrix<T, device>& mat_a) {
  cholesky_factorization<backend, device, T>(uplo, mat_a);
}
template <Backend backend, Device device, class T>
void cholesky(comm::CommunicatorGrid& grid, blas::Uplo uplo, Matrix<T, device>& mat_a) {
  cholesky_factorization<backend, device, T>(grid, uplo, mat_a);
}
}  // namespace factorization

}  // namespace dlaf
