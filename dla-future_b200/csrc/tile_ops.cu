// Host-side composition of the tile kernels (see tile_ops.cuh). Nothing here synchronises or allocates.
#include "tile_ops.cuh"

#include <cstdlib>
#include <type_traits>

#include "gemm_dmma.cuh"
#include "gemm_simt.cuh"

namespace dlafb {
namespace {

inline bool aligned16(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) & 15) == 0;
}

template <class T>
bool dmma_ok(int, int, const T*, long long, const T*, long long, int) {
  return false;
}
template <>
bool dmma_ok<double>(int M, int N, const double* A, long long lda, const double* B, long long ldb, int K) {
  return M >= 1 && N >= 1 && K > 0 && (K % kBK) == 0 && aligned16(A) && aligned16(B) && (lda % 2) == 0 &&
         (ldb % 2) == 0;
}
template <>
bool dmma_ok<double2>(int M, int N, const double2* A, long long, const double2* B, long long, int K) {
  // the real view (2M x K doubles, leading dimension 2*lda) is always 16-byte aligned with an even leading dimension
  return M >= 1 && N >= 1 && K > 0 && (K % kBK) == 0 && aligned16(A) && aligned16(B) && M < (1 << 29) && N < (1 << 29);
}

// cplx: A, B, C are the real views of complex<double> matrices; M, N, lda, ldb, ldc already doubled by the caller
// (C: 2M rows x N/2 columns of complex elements = N product columns).
cudaError_t gemm_nt_dmma_rect(cudaStream_t stream, int M, int N, int K, double alpha, const double* A, long long lda,
                              const double* B, long long ldb, double beta, double* C, long long ldc, int lower,
                              bool inplace, bool cplx = false) {
  // few blocks = a launch whose duration is one CTA's serial chain of K/16 slices: use 4x more, smaller CTAs
  // (the in-place product needs all N <= 128 columns in one CTA, so it splits the rows only)
  const bool small = ceil_div(M, kBM) * ceil_div(N, kBN) < 120;
  static const GemmShape big_shape = [] {
    const char* e = getenv("DLAFB_BULK_SHAPE");
    return (e && e[0] == '0') ? GemmShape::k128x128 : (e && e[0] == '3') ? GemmShape::k64x128 : kDefaultBulkShape;
  }();
  const GemmShape shape = !small ? big_shape : inplace ? GemmShape::k32x128 : GemmShape::k64x64;
  const int bm = shape_bm(shape), bn = shape_bn(shape);
  CUtensorMap ma, mb;
  int ka = (M % 16 == 0) ? 3 : 2, kb = (N % 16 == 0) ? 3 : 2;
  if (ka == 3 && make_operand_map(&ma, A, M, K, lda, 3, bm) != 0)
    ka = 2;
  if (ka == 2 && make_operand_map(&ma, A, M, K, lda, 2) != 0)
    return cudaErrorInvalidValue;
  if (kb == 3 && make_operand_map(&mb, B, N, K, ldb, 3, bn) != 0)
    kb = 2;
  if (kb == 2 && make_operand_map(&mb, B, N, K, ldb, 2) != 0)
    return cudaErrorInvalidValue;
  GemmParams p{};
  p.C = C;
  p.ldc = ldc;
  p.m_total = M;
  p.n_total = N;
  p.K = K;
  p.alpha = alpha;
  p.beta = beta;
  p.lower = lower;
  p.nbox_a = ka == 3 ? 1 : 8;
  p.nbox_b = kb == 3 ? 1 : 8;
  p.list_scale = 1;
  return launch_gemm_nt_dmma(stream, ma, mb, p, ceil_div(M, bm), ceil_div(N, bn), shape, cplx);
}

template <class T>
cudaError_t gemm_nt_dmma_try(cudaStream_t, int, int, int, typename Tr<T>::Real, const T*, long long, const T*,
                             long long, typename Tr<T>::Real, T*, long long, int, bool) {
  return cudaErrorNotSupported;
}
template <>
cudaError_t gemm_nt_dmma_try<double>(cudaStream_t s, int M, int N, int K, double alpha, const double* A, long long lda,
                                     const double* B, long long ldb, double beta, double* C, long long ldc, int lower,
                                     bool inplace) {
  return gemm_nt_dmma_rect(s, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower, inplace);
}

template <>
cudaError_t gemm_nt_dmma_try<double2>(cudaStream_t s, int M, int N, int K, double alpha, const double2* A, long long lda,
                                      const double2* B, long long ldb, double beta, double2* C, long long ldc,
                                      int lower, bool inplace) {
  return gemm_nt_dmma_rect(s, 2 * M, 2 * N, K, alpha, reinterpret_cast<const double*>(A), 2 * lda,
                           reinterpret_cast<const double*>(B), 2 * ldb, beta, reinterpret_cast<double*>(C), 2 * ldc,
                           lower, inplace, true);
}

}  // namespace

template <class T>
cudaError_t gemm_nt(cudaStream_t stream, int M, int N, int K, typename Tr<T>::Real alpha, const T* A, long long lda,
                    const T* B, long long ldb, typename Tr<T>::Real beta, T* C, long long ldc, int lower) {
  if (M <= 0 || N <= 0)
    return cudaSuccess;
  const bool inplace = static_cast<const void*>(A) == static_cast<const void*>(C);
  if (dmma_ok<T>(M, N, A, lda, B, ldb, K)) {
    if (inplace && N * (Tr<T>::is_complex ? 2 : 1) > kBN)
      return cudaErrorInvalidValue;
    return gemm_nt_dmma_try<T>(stream, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower, inplace);
  }
  if (inplace && N > kSimtTile)
    return cudaErrorInvalidValue;
  SimtParams<T> p{};
  p.C = C;
  p.ldc = ldc;
  p.A = A;
  p.lda = lda;
  p.B = B;
  p.ldb = ldb;
  p.m_total = M;
  p.n_total = N;
  p.K = K;
  p.alpha = alpha;
  p.beta = beta;
  p.lower = lower;
  return launch_gemm_simt<T>(stream, p, ceil_div(M, kSimtTile), ceil_div(N, kSimtTile));
}

template <class T>
int inner_block(const T* a, long long lda, int nb) {
  if (std::is_same<T, double>::value && aligned16(a) && (lda % 2) == 0 && (nb % 128) == 0)
    return 128;
  return 64;
}

template <class T>
size_t potrf_work_elems(int n, int ib) {
  return (size_t)ceil_div(n > 0 ? n : 1, ib) * ib * ib + (size_t)n * n;
}

template <class T>
cudaError_t potrf_tile(cudaStream_t stream, char uplo, int n, T* a, long long lda, T* work, int ib, int* info,
                       int info_base) {
  using R = typename Tr<T>::Real;
  if (n <= 0)
    return cudaSuccess;
  const bool upper = (uplo == 'U' || uplo == 'u');
  cudaError_t e;
  if (n <= ib)
    return launch_potf2<T>(stream, uplo, n, a, lda, work, ib, info, info_base, 3);
  T* l = a;
  long long ldl = lda;
  T* wt = work + (size_t)ceil_div(n, ib) * ib * ib;
  if (upper) {
    // factor the conjugate transpose as a lower tile in the workspace, write the upper triangle back
    if ((e = transpose<T>(stream, n, n, a, lda, wt, n, 1, 'L')) != cudaSuccess)
      return e;
    l = wt;
    ldl = n;
  }
  for (int j0 = 0, j = 0; j0 < n; j0 += ib, ++j) {
    const int w = n - j0 < ib ? n - j0 : ib;
    T* wj = work + (size_t)j * ib * ib;
    T* ljj = l + j0 + (long long)j0 * ldl;
    if ((e = launch_potf2<T>(stream, 'L', w, ljj, ldl, wj, ib, info, info_base + j0, 3)) != cudaSuccess)
      return e;
    const int rem = n - j0 - w;
    if (rem == 0)
      break;
    T* b = ljj + w;
    if ((e = gemm_nt<T>(stream, rem, w, w, R(1), b, ldl, wj, ib, R(0), b, ldl, 0)) != cudaSuccess)
      return e;
    T* c = l + (j0 + w) + (long long)(j0 + w) * ldl;
    if ((e = gemm_nt<T>(stream, rem, rem, w, R(-1), b, ldl, b, ldl, R(1), c, ldl, 1)) != cudaSuccess)
      return e;
  }
  if (upper)
    return transpose<T>(stream, n, n, wt, n, a, lda, 1, 'U');
  return cudaSuccess;
}

template <class T>
cudaError_t trsm_panel(cudaStream_t stream, int nk, const T* L, long long ldl, const T* W, int ib, T* B,
                       long long ldb, int m) {
  using R = typename Tr<T>::Real;
  if (m <= 0 || nk <= 0)
    return cudaSuccess;
  cudaError_t e;
  if constexpr (std::is_same<T, double>::value) {
    if (trsm_panel_fused_d(stream, nk, L, ldl, W, ib, B, ldb, m, nullptr, 0, &e))
      return e;  // one kernel
  }
  for (int j0 = 0, j = 0; j0 < nk; j0 += ib, ++j) {
    const int w = nk - j0 < ib ? nk - j0 : ib;
    T* bj = B + (long long)j0 * ldb;
    if ((e = gemm_nt<T>(stream, m, w, w, R(1), bj, ldb, W + (size_t)j * ib * ib, ib, R(0), bj, ldb, 0)) != cudaSuccess)
      return e;
    const int rem = nk - j0 - w;
    if (rem == 0)
      break;
    if ((e = gemm_nt<T>(stream, m, rem, w, R(-1), bj, ldb, L + (j0 + w) + (long long)j0 * ldl, ldl, R(1),
                        B + (long long)(j0 + w) * ldb, ldb, 0)) != cudaSuccess)
      return e;
  }
  return cudaSuccess;
}

template <class T>
cudaError_t trsm_panel_packed(cudaStream_t stream, int nk, const T* L, long long ldl, const T* W, int ib, T* B,
                              long long ldb, int m, T* packed, int tile) {
  if (m <= 0 || nk <= 0)
    return cudaSuccess;
  cudaError_t e;
  if constexpr (std::is_same<T, double>::value) {
    if (trsm_panel_fused_d(stream, nk, L, ldl, W, ib, B, ldb, m, packed, tile, &e))
      return e;
  }
  if ((e = trsm_panel<T>(stream, nk, L, ldl, W, ib, B, ldb, m)) != cudaSuccess)
    return e;
  return pack_tiles<T>(stream, m, nk, B, ldb, packed, tile);
}

#define DLAFB_INST(T)                                                                                              \
  template cudaError_t gemm_nt<T>(cudaStream_t, int, int, int, Tr<T>::Real, const T*, long long, const T*, long long, \
                                  Tr<T>::Real, T*, long long, int);                                                \
  template int inner_block<T>(const T*, long long, int);                                                           \
  template size_t potrf_work_elems<T>(int, int);                                                                   \
  template cudaError_t potrf_tile<T>(cudaStream_t, char, int, T*, long long, T*, int, int*, int);                  \
  template cudaError_t trsm_panel<T>(cudaStream_t, int, const T*, long long, const T*, int, T*, long long, int);  \
  template cudaError_t trsm_panel_packed<T>(cudaStream_t, int, const T*, long long, const T*, int, T*, long long,  \
                                            int, T*, int);
DLAFB_INST(float)
DLAFB_INST(double)
DLAFB_INST(float2)
DLAFB_INST(double2)

}  // namespace dlafb
