// TEST INFRASTRUCTURE, RUN ON THE DEVICE by tests/test_cxx_frontend_gpu.py: the public C++ front end of the path
// executed the way a caller of the reference would use it. For the reference's local test sizes
// (test/unit/factorization/test_cholesky.cpp:54-62), both uplo and the four element types:
//   Matrix<T, Device::CPU> filled tile by tile with the analytical A (getCholeskySetters,
//   test/include/dlaf_test/matrix/util_generic_lapack.h:27-68)
//   -> MatrixMirror<T, Device::GPU, Device::CPU> -> cholesky_factorization<Backend::GPU, Device::GPU, T>
//   (local overload, and the CommunicatorGrid overload on a 1 x 1 grid) -> mirror copy-back -> compared element-wise
//   with the analytical factor at 4 (m+1) error, unreferenced triangle untouched (-9.9).
// Also: a Device::GPU matrix factorized in place + waitLocalTiles + copy to the host; the pre-0.2.0 spelling
// factorization::cholesky; repeated calls (the plan cache behind the front end). Prints "checks=N failures=M".
#include <cmath>
#include <complex>
#include <cstdio>
#include <limits>
#include <type_traits>

#include "dlaf/communication/communicator_grid.h"
#include "dlaf/factorization/cholesky.h"
#include "dlaf/init.h"
#include "dlaf/matrix/matrix.h"
#include "dlaf/matrix/matrix_mirror.h"

using namespace dlaf;

static long checks = 0, failures = 0;

template <class T>
struct Info {
  using Real = T;
  static constexpr bool cplx = false;
  static T make(double mag, double) { return static_cast<T>(mag); }
};
template <class R>
struct Info<std::complex<R>> {
  using Real = R;
  static constexpr bool cplx = true;
  static std::complex<R> make(double mag, double phase) {
    return std::complex<R>(static_cast<R>(mag * std::cos(phase)), static_cast<R>(mag * std::sin(phase)));
  }
};

// t_ij = 2^-|i-j| e^{I(j-i)}; a_ij = (4^(min+1) - 1) / (3 2^(i+j)) e^{I(j-i)}; -9.9 outside the referenced triangle
template <class T>
T el(bool factor, blas::Uplo uplo, SizeType i, SizeType j) {
  if ((uplo == blas::Uplo::Lower && i < j) || (uplo == blas::Uplo::Upper && i > j))
    return Info<T>::make(-9.9, 0.0);
  const double mn = (double)std::min(i, j);
  const double mag = factor ? std::exp2(-std::fabs((double)(i - j)))
                            : (std::exp2(2 * mn + 2 - (double)(i + j)) - std::exp2(-(double)(i + j))) / 3.0;
  return Info<T>::make(mag, (double)(j - i));
}

template <class T>
void fill(Matrix<T, Device::CPU>& m, blas::Uplo uplo) {
  const auto& d = m.distribution();
  const SizeType mb = d.block_size().rows(), nb = d.block_size().cols();
  for (SizeType jl = 0; jl < d.local_nr_tiles().cols(); ++jl)
    for (SizeType il = 0; il < d.local_nr_tiles().rows(); ++il) {
      const LocalTileIndex lt(il, jl);
      const GlobalTileIndex gt = d.global_tile_index(lt);
      auto tile = m.readwrite(lt);
      for (SizeType c = 0; c < tile.size().cols(); ++c)
        for (SizeType r = 0; r < tile.size().rows(); ++r)
          tile(TileElementIndex(r, c)) = el<T>(false, uplo, gt.row() * mb + r, gt.col() * nb + c);
    }
}

template <class T>
long compare(Matrix<T, Device::CPU>& m, blas::Uplo uplo, SizeType n) {
  using R = typename Info<T>::Real;
  const double err = (Info<T>::cplx ? 8.0 : 2.0) * (double)std::numeric_limits<R>::epsilon();  // util_types.h:40,:62
  const double tol = 4.0 * (double)(n + 1) * err;
  const auto& d = m.distribution();
  const SizeType mb = d.block_size().rows(), nb = d.block_size().cols();
  long bad = 0;
  for (SizeType jl = 0; jl < d.local_nr_tiles().cols(); ++jl)
    for (SizeType il = 0; il < d.local_nr_tiles().rows(); ++il) {
      const LocalTileIndex lt(il, jl);
      const GlobalTileIndex gt = d.global_tile_index(lt);
      auto tile = m.read(lt);
      for (SizeType c = 0; c < tile.size().cols(); ++c)
        for (SizeType r = 0; r < tile.size().rows(); ++r) {
          const T want = el<T>(true, uplo, gt.row() * mb + r, gt.col() * nb + c);
          const T got = *tile.ptr(TileElementIndex(r, c));
          const double diff = std::abs(got - want), rel = diff / std::abs(want);
          if (!(diff <= tol || rel <= tol))  // CHECK_MATRIX_NEAR: absolute OR relative
            ++bad;
        }
    }
  return bad;
}

enum class Path { MirrorLocal, MirrorGrid, DeviceMatrix, OldSpelling };

template <class T>
void one_case(comm::CommunicatorGrid& grid, blas::Uplo uplo, SizeType n, SizeType nb, Path path) {
  Matrix<T, Device::CPU> host(LocalElementSize(n, n), TileElementSize(nb, nb));
  fill(host, uplo);
  switch (path) {
    case Path::MirrorLocal: {
      matrix::MatrixMirror<T, Device::GPU, Device::CPU> mirror(host);
      cholesky_factorization<Backend::GPU, Device::GPU, T>(uplo, mirror.get());
    } break;
    case Path::MirrorGrid: {
      matrix::MatrixMirror<T, Device::GPU, Device::CPU> mirror(host);
      cholesky_factorization<Backend::GPU, Device::GPU, T>(grid, uplo, mirror.get());
    } break;
    case Path::DeviceMatrix: {
      Matrix<T, Device::GPU> dev(host.distribution());
      matrix::copy(host, dev);
      cholesky_factorization<Backend::GPU, Device::GPU, T>(uplo, dev);
      dev.waitLocalTiles();
      matrix::copy(dev, host);
      host.waitLocalTiles();
    } break;
    case Path::OldSpelling: {
      matrix::MatrixMirror<T, Device::GPU, Device::CPU> mirror(host);
      factorization::cholesky<Backend::GPU, Device::GPU, T>(uplo, mirror.get());
    } break;
  }
  const long bad = compare(host, uplo, n);
  ++checks;
  if (bad != 0) {
    ++failures;
    std::fprintf(stderr, "FAILED type=%c uplo=%c n=%ld nb=%ld path=%d: %ld elements off\n",
                 dlaf::internal::TypeCode<T>::value, uplo == blas::Uplo::Lower ? 'L' : 'U', (long)n, (long)nb, (int)path, bad);
  }
}

template <class T>
void all_cases(comm::CommunicatorGrid& grid) {
  const SizeType sizes[][2] = {{0, 2}, {5, 8}, {34, 34}, {4, 3}, {16, 10}, {34, 13}, {32, 5}};  // test_cholesky.cpp:54-62
  for (blas::Uplo uplo : {blas::Uplo::Lower, blas::Uplo::Upper}) {
    for (const auto& s : sizes) {
      one_case<T>(grid, uplo, s[0], s[1], Path::MirrorLocal);
      one_case<T>(grid, uplo, s[0], s[1], Path::MirrorGrid);
    }
    // the TMA + DMMA path (double: nb % 128 == 0) and a ragged last tile, through every entry; twice: cached plan
    for (int rep = 0; rep < 2; ++rep) {
      one_case<T>(grid, uplo, 640, 128, Path::DeviceMatrix);
      one_case<T>(grid, uplo, 700, 128, Path::MirrorGrid);
      one_case<T>(grid, uplo, 96, 32, Path::OldSpelling);
    }
  }
}

int main() {
  ScopedInitializer init;
  comm::Communicator world(MPI_COMM_WORLD);
  comm::CommunicatorGrid grid(world, 1, 1, common::Ordering::RowMajor);
  all_cases<float>(grid);
  all_cases<double>(grid);
  all_cases<std::complex<float>>(grid);
  all_cases<std::complex<double>>(grid);
  std::printf("checks=%ld failures=%ld\n", checks, failures);
  return failures != 0;
}
