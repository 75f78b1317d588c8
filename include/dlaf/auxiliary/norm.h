// dlaf::auxiliary::max_norm<Backend, Device, T>(grid, rank, uplo, A): max norm of a distributed matrix, signature and
// semantics of the reference (include/dlaf/auxiliary/norm.h:35-58): uplo == Lower (square matrix, square tiles; valid for
// sy / he / lower-triangular data) or General; Upper is not implemented there either; 0 for an empty matrix. The
// reference implements it for Backend::MC only (auxiliary/norm/mc.h); here it runs on the device. Collective over the
// grid, synchronous. Every rank receives the value (the reference returns it on @p rank).
#pragma once
#include "dlaf/communication/communicator_grid.h"
#include "dlaf/matrix/matrix.h"
#include "dlafb_matrix.h"

namespace dlaf::auxiliary {

template <Backend backend, Device device, class T>
dlaf::BaseType<T> max_norm(comm::CommunicatorGrid& grid, comm::Index2D /*rank*/, blas::Uplo uplo,
                           Matrix<T, device>& mat_a) {
  static_assert(backend == Backend::GPU && device == Device::GPU,
                "this build has no CPU backend: use Backend::GPU with Device::GPU matrices");
  DLAF_ASSERT(mat_a.grid_size() == grid.size() && mat_a.rank_index() == grid.rank(), mat_a.grid_size());
  DLAF_ASSERT(uplo == blas::Uplo::Lower || uplo == blas::Uplo::General, static_cast<char>(uplo));
  if (mat_a.size().isEmpty())
    return {0};
  if (uplo == blas::Uplo::Lower) {
    DLAF_ASSERT(mat_a.size().rows() == mat_a.size().cols(), mat_a.size());
    DLAF_ASSERT(mat_a.block_size().rows() == mat_a.block_size().cols(), mat_a.block_size());
  }
  const comm::Index2D src = mat_a.distribution().source_rank_index();
  double result = 0;
  const int e = dlafb_max_norm(grid.handle(), dlaf::internal::TypeCode<T>::value, static_cast<char>(uplo),
                               mat_a.size().rows(), mat_a.size().cols(), mat_a.block_size().rows(),
                               mat_a.block_size().cols(), src.row(), src.col(), mat_a.ptr(), mat_a.ld(), mat_a.stream(),
                               &result);
  DLAF_ASSERT(e == 0, e);
  return static_cast<dlaf::BaseType<T>>(result);
}

}  // namespace dlaf::auxiliary
