// matrix::util: matrix generators and predicates the reference's miniapps and callers use
// (reference: include/dlaf/util_matrix.h). On this path the generators run on the device:
//   set_random_hermitian_positive_definite(mat)   util_matrix.h:499-501 there: Hermitian, |off-diagonal| <= 1, real
//       diagonal = u + 2n. Same distribution of values, NOT the same random stream (the reference seeds one
//       std::mt19937_64 per tile on the host; here a counter-based generator hashes the global indices, so the matrix
//       is likewise independent of the distribution and of the number of ranks).
//   set_random_hermitian(mat)                     the same generator (its diagonal is real and dominant as well)
//   set0(mat)                                     util_matrix.h: all local tiles to zero
// and the predicates the front ends assert with (square_size, square_blocksize, equal_process_grid, local_matrix).
#pragma once
#include <complex>

#include "dlaf/communication/communicator_grid.h"
#include "dlaf/matrix/matrix.h"
#include "dlafb_kernels.h"

namespace dlaf::matrix {

template <class T, Device D>
bool square_size(const Matrix<T, D>& m) noexcept { return m.size().rows() == m.size().cols(); }
template <class T, Device D>
bool square_blocksize(const Matrix<T, D>& m) noexcept { return m.block_size().rows() == m.block_size().cols(); }
template <class T, Device D>
bool local_matrix(const Matrix<T, D>& m) noexcept { return m.grid_size() == comm::Size2D(1, 1); }
template <class T, Device D>
bool equal_process_grid(const Matrix<T, D>& m, const comm::CommunicatorGrid& g) noexcept {
  return m.grid_size() == g.size() && m.rank_index() == g.rank();
}

namespace util {
namespace internal {
inline int gen_hpd(float*, void* s, int64_t n, int nb, int r, int c, int rf, int rs, int cf, int cs, void* a, int64_t ld, uint64_t seed) {
  return dlafb_gen_hpd_s(s, n, nb, r, c, rf, rs, cf, cs, static_cast<float*>(a), ld, seed);
}
inline int gen_hpd(double*, void* s, int64_t n, int nb, int r, int c, int rf, int rs, int cf, int cs, void* a, int64_t ld, uint64_t seed) {
  return dlafb_gen_hpd_d(s, n, nb, r, c, rf, rs, cf, cs, static_cast<double*>(a), ld, seed);
}
inline int gen_hpd(std::complex<float>*, void* s, int64_t n, int nb, int r, int c, int rf, int rs, int cf, int cs, void* a, int64_t ld, uint64_t seed) {
  return dlafb_gen_hpd_c(s, n, nb, r, c, rf, rs, cf, cs, a, ld, seed);
}
inline int gen_hpd(std::complex<double>*, void* s, int64_t n, int nb, int r, int c, int rf, int rs, int cf, int cs, void* a, int64_t ld, uint64_t seed) {
  return dlafb_gen_hpd_z(s, n, nb, r, c, rf, rs, cf, cs, a, ld, seed);
}
}  // namespace internal

/// @pre square size, square blocks. Asynchronous on the matrix's stream.
template <class T>
void set_random_hermitian_positive_definite(Matrix<T, Device::GPU>& mat, uint64_t seed = 42) {
  DLAF_ASSERT(square_size(mat), mat.size());
  DLAF_ASSERT(square_blocksize(mat), mat.block_size());
  const Distribution& d = mat.distribution();
  const LocalElementSize ls = d.local_size();
  if (ls.isEmpty())
    return;
  const int e = internal::gen_hpd(static_cast<T*>(nullptr), mat.stream(), d.size().rows(), (int)d.block_size().rows(),
                                  (int)ls.rows(), (int)ls.cols(),
                                  (int)d.template global_tile_from_local_tile<Coord::Row>(0), d.grid_size().rows(),
                                  (int)d.template global_tile_from_local_tile<Coord::Col>(0), d.grid_size().cols(),
                                  mat.ptr(), mat.ld(), seed);
  DLAF_ASSERT(e == 0, e);
}
/// Host matrix: generated on the device in the same distribution, then copied.
template <class T>
void set_random_hermitian_positive_definite(Matrix<T, Device::CPU>& mat, uint64_t seed = 42) {
  Matrix<T, Device::GPU> dev(mat.distribution());
  set_random_hermitian_positive_definite(dev, seed);
  copy(dev, mat);
  mat.waitLocalTiles();
}
template <class T, Device D>
void set_random_hermitian(Matrix<T, D>& mat, uint64_t seed = 43) {
  set_random_hermitian_positive_definite(mat, seed);
}

/// All local elements to zero. Asynchronous on the matrix's stream.
template <class T>
void set0(Matrix<T, Device::GPU>& mat) {
  const LocalElementSize ls = mat.distribution().local_size();
  if (!ls.isEmpty())
    cudaMemset2DAsync(mat.ptr(), mat.ld() * sizeof(T), 0, ls.rows() * sizeof(T), ls.cols(), mat.stream());
}

}  // namespace util
}  // namespace dlaf::matrix
