/* Thin C ABI over the sm_100a tile kernels of the Cholesky path ("lower boundary").
 *
 * Every function: plain pointers and sizes only; all matrix pointers are DEVICE pointers owned by the caller,
 * column-major with the given leading dimension; work is ordered on `stream` (a cudaStream_t passed as void*);
 * nothing here allocates, synchronises or throws; the return value is a cudaError_t as int (0 = success).
 * Numerical failure (non positive definite tile) is reported through the device int `info_dev`
 * (first failing column, 1-based, plus info_base; first failure wins) exactly like LAPACK potrf's info.
 *
 * X = s (float), d (double), c (float2 = complex<float>), z (double2 = complex<double>).
 *
 * Reference interfaces these replace (paths relative to the DLA-Future tree):
 *   dlafb_potrf_tile_X  -> tile::potrf / potrfInfo on GPU = cusolverDnXpotrf   include/dlaf/lapack/tile.h:578-606
 *   dlafb_trsm_panel_X  -> tile::trsm on GPU = cublasXtrsm, one call per tile  include/dlaf/blas/tile.h:458-470
 *                          (arguments of impl.h:55-67: Right, Lower, ConjTrans, NonUnit, alpha = 1)
 *   dlafb_gemm_nt_X     -> tile::gemm (NoTrans x ConjTrans, impl.h:82-94) and tile::herk (Lower, NoTrans,
 *                          impl.h:69-80; pass A == B and lower = 1)             include/dlaf/blas/tile.h:370-425
 *   dlafb_laset_X       -> tile::laset GPU                                      src/lapack/gpu/laset.cu:66-120
 *   dlafb_lacpy_X       -> gpulapack::lacpy                                     src/lapack/gpu/lacpy.cu:69-209
 *   dlafb_max_abs_X     -> tile::lange / lantr (Norm::Max)                      include/dlaf/auxiliary/norm/mc.h:35-95
 *   dlafb_gen_hpd_X     -> matrix::util::set_random_hermitian_positive_definite include/dlaf/util_matrix.h:499-501
 *                          (same distribution of values, counter-based stream: timing runs only)
 */
#ifndef DLAFB_KERNELS_H
#define DLAFB_KERNELS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Inner block size (64 or 128) the blocked tile algorithms use for an operand at `a` with leading dimension lda
 * and tile size nb; type is 's', 'd', 'c' or 'z'. */
int dlafb_inner_block(char type, const void* a, int64_t lda, int nb);
/* Workspace, in ELEMENTS, that dlafb_potrf_tile_X needs for an n x n tile. */
size_t dlafb_potrf_work_elems(char type, int n, int ib);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate
 * device memory (code under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * dlafb_potrf_tile_d with its workspace in a file-scope device array of the library (2.5 MiB held by every GPU that
 * loads the library until the process exits) instead of a caller-supplied buffer: in-place Cholesky of the uplo
 * triangle of the n x n double tile a, info as above. n <= DLAFB_POTRF_NOALLOC_MAX_N, larger n is rejected
 * (cudaErrorInvalidValue = 1 and a message on stderr). Calls no cudaMalloc / cudaFree / allocating routine. Every
 * call shares the array (the part used is cleared first, on `stream`): calls whose work overlaps on the GPU race -
 * keep them ordered (one stream, or events between streams). */
#define DLAFB_POTRF_NOALLOC_MAX_N 512
int dlafb_potrf_tile_noalloc_d(void* stream, char uplo, int n, double* a, int64_t lda, int* info_dev, int info_base);

#define DLAFB_DECLARE_TILE_API(X, T)                                                                               \
  /* In-place Cholesky of the uplo ('L'/'U') triangle of the n x n tile a. On return work[j*ib*ib ..] (ld = ib)   \
   * holds inv(L_jj) for the j-th ib x ib diagonal block (L = U^H for uplo 'U'). */                                \
  int dlafb_potrf_tile_##X(void* stream, char uplo, int n, T* a, int64_t lda, T* work, int ib, int* info_dev,     \
                           int info_base);                                                                         \
  /* B (m x nk) <- B * L^-H, L the nk x nk lower factor, w = the inverted diagonal blocks left by potrf_tile. */   \
  int dlafb_trsm_panel_##X(void* stream, int nk, const T* l, int64_t ldl, const T* w, int ib, T* b, int64_t ldb,  \
                           int m);                                                                                 \
  /* The same solve, and the solved panel also copied into tile layout: row i of b -> slot i / tile of `packed`        \
   * (contiguous tile x tile column-major tiles; what the row panel of the distributed factorization holds, reference \
   * matrix::Panel with tileLayout, include/dlaf/matrix/panel.h:389-391). One kernel for double when nk % 128 == 0. */ \
  int dlafb_trsm_panel_packed_##X(void* stream, int nk, const T* l, int64_t ldl, const T* w, int ib, T* b,             \
                                  int64_t ldb, int m, T* packed, int tile);                                           \
  /* C (M x N) <- beta*C + alpha * A (M x K) * B (N x K)^H, real alpha/beta; lower != 0: only i >= j is touched    \
   * and (complex) the diagonal is kept real: with A == B this is herk(Lower, NoTrans). */                         \
  int dlafb_gemm_nt_##X(void* stream, int m, int n, int k, double alpha, const T* a, int64_t lda, const T* b,     \
                        int64_t ldb, double beta, T* c, int64_t ldc, int lower);                                   \
  /* out (n x m) <- in (m x n)^T, conjugated when conj != 0; tri 'G' | 'L' | 'U' restricts the part of out written. */ \
  int dlafb_transpose_##X(void* stream, int m, int n, const T* in, int64_t ldin, T* out, int64_t ldout, int conj, \
                          char tri);                                                                               \
  /* off-diagonal <- (alpha_re, alpha_im), diagonal <- (beta_re, beta_im) on the uplo ('G','L','U') part. */      \
  int dlafb_laset_##X(void* stream, char uplo, int m, int n, double alpha_re, double alpha_im, double beta_re,    \
                      double beta_im, T* a, int64_t lda);                                                          \
  int dlafb_lacpy_##X(void* stream, char uplo, int m, int n, const T* a, int64_t lda, T* b, int64_t ldb);         \
  /* max |a_ij| over the uplo part, combined into *out_dev (device double, zero it first; NaN reads as +inf). */   \
  int dlafb_max_abs_##X(void* stream, char uplo, int m, int n, const T* a, int64_t lda, double* out_dev);         \
  /* Synthetic Hermitian positive definite matrix, local block of a block-cyclic distribution: local tile l of   \
   * the rows is global tile row_first + l*row_stride (same for columns). */                                      \
  int dlafb_gen_hpd_##X(void* stream, int64_t n, int nb, int rows, int cols, int row_first, int row_stride,       \
                        int col_first, int col_stride, T* a, int64_t lda, uint64_t seed);

DLAFB_DECLARE_TILE_API(s, float)
DLAFB_DECLARE_TILE_API(d, double)
DLAFB_DECLARE_TILE_API(c, void)
DLAFB_DECLARE_TILE_API(z, void)

#ifdef __cplusplus
}
#endif
#endif
