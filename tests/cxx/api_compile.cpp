// TEST INFRASTRUCTURE: instantiates every public C++ front end of include/dlaf/ for the four element types, so that
// the CPU test-suite (tests/test_cxx_headers.py) catches a header that no longer compiles. Never run.
#include "dlaf/auxiliary/norm.h"
#include "dlaf/common/range2d.h"
#include "dlaf/factorization.h"
#include "dlaf/matrix/index.h"
#include "dlaf/matrix/matrix_mirror.h"
#include "dlaf/solver.h"
#include "dlaf/util_matrix.h"
#include "dlaf/common/round_robin.h"
#include "dlaf/common/timer.h"
#include "dlaf/communication/init.h"
#include "dlaf/tune.h"
#include "dlaf/eigensolver/gen_to_std.h"
#include "dlaf/factorization/cholesky.h"
#include "dlaf/matrix/copy.h"
#include "dlaf/solver/triangular.h"

#include <complex>

template <class T>
void instantiate(dlaf::comm::CommunicatorGrid& grid) {
  using namespace dlaf;
  Matrix<T, Device::GPU> a(LocalElementSize(8, 8), TileElementSize(4, 4));
  Matrix<T, Device::GPU> b(LocalElementSize(8, 8), TileElementSize(4, 4));
  cholesky_factorization<Backend::GPU, Device::GPU, T>(blas::Uplo::Lower, a);
  cholesky_factorization<Backend::GPU, Device::GPU, T>(grid, blas::Uplo::Lower, a);
  triangular_solver<Backend::GPU, Device::GPU, T>(blas::Side::Left, blas::Uplo::Lower, blas::Op::NoTrans,
                                                  blas::Diag::NonUnit, T(1), a, b);
  triangular_solver<Backend::GPU, Device::GPU, T>(grid, blas::Side::Right, blas::Uplo::Upper, blas::Op::ConjTrans,
                                                  blas::Diag::Unit, T(1), a, b);
  eigensolver::internal::generalized_to_standard<Backend::GPU, Device::GPU, T>(blas::Uplo::Lower, b, a);
  eigensolver::internal::generalized_to_standard<Backend::GPU, Device::GPU, T>(grid, blas::Uplo::Upper, b, a);
  Matrix<T, Device::GPU> c(GlobalElementSize(8, 8), TileElementSize(2, 2), grid);
  Matrix<T, Device::GPU> d(GlobalElementSize(8, 8), TileElementSize(4, 4), grid);
  matrix::copy(c, d, grid);
  d.waitLocalTiles();
  (void)auxiliary::max_norm<Backend::GPU, Device::GPU, T>(grid, comm::Index2D(0, 0), blas::Uplo::Lower, d);
  matrix::util::set_random_hermitian_positive_definite(a);
  matrix::util::set_random_hermitian(b);
  matrix::util::set0(c);
  Matrix<T, Device::CPU> h(LocalElementSize(8, 8), TileElementSize(4, 4));
  matrix::util::set_random_hermitian_positive_definite(h);
  {
    matrix::MatrixMirror<T, Device::GPU, Device::CPU> mirror(h);
    cholesky_factorization<Backend::GPU, Device::GPU, T>(blas::Uplo::Lower, mirror.get());
  }
  (void)(matrix::square_size(a) && matrix::square_blocksize(a) && matrix::local_matrix(a) && matrix::equal_process_grid(c, grid));
  SizeType count = 0;
  for (auto ij : common::iterate_range2d(a.distribution().local_nr_tiles()))
    count += a.read(ij).size().rows();
  for (auto ij : common::iterate_range2d(LocalTileIndex(1, 0), LocalTileIndex(2, 2)))
    count += ij.col();
  (void)count;
  a.waitLocalTiles();
  b.waitLocalTiles();
}

void instantiate_misc(dlaf::comm::CommunicatorGrid& grid) {
  auto row = grid.row_communicator_pipeline();
  auto col = grid.col_communicator_pipeline();
  auto full = grid.full_communicator_pipeline();
  (void)(row.rank() + col.size() + full.rank_full_communicator(grid.rank()) + (row.valid() ? 1 : 0));
  dlaf::common::RoundRobin<int> rr(2, 7);
  (void)(rr.nextResource() + rr.currentResource());
  dlaf::common::Timer<> t;
  (void)t.elapsed();
  dlaf::getTuneParameters().sgemm_wide = 0;
  dlaf::getTuneParameters().apply();
}

void instantiate_all(dlaf::comm::CommunicatorGrid& grid) {
  instantiate_misc(grid);
  instantiate<float>(grid);
  instantiate<double>(grid);
  instantiate<std::complex<float>>(grid);
  instantiate<std::complex<double>>(grid);
}
