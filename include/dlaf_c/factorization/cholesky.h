/* Cholesky entry points of the C API; signatures are the reference's
 * (include/dlaf_c/factorization/cholesky.h:31-89, src/c_api/factorization/cholesky.cpp:16-56).
 *
 * `a` is the caller's LOCAL block in column-major order with leading dimension desc.ld / desca[8]. It may be a
 * host pointer (reference behaviour: mirrored to the GPU and back, src/c_api/factorization/cholesky.h:48-55) or,
 * as an extension, a device pointer (factorized in place, no copies). Calls are collective over the grid and
 * synchronous. Only the uplo triangle is referenced and overwritten.
 *
 * dlaf_cholesky_factorization_X returns 0 on success (reference: always 0) or the ScaLAPACK-style info > 0 when
 * the matrix is not positive definite (the reference aborts in that case). dlaf_pXpotrf reports the same value
 * through *info. */
#pragma once
#include <dlaf_c/desc.h>
#include <dlaf_c/utils.h>

DLAF_EXTERN_C int dlaf_cholesky_factorization_s(const int dlaf_context, const char uplo, float* a,
                                                const struct DLAF_descriptor dlaf_desca) DLAF_NOEXCEPT;
DLAF_EXTERN_C int dlaf_cholesky_factorization_d(const int dlaf_context, const char uplo, double* a,
                                                const struct DLAF_descriptor dlaf_desca) DLAF_NOEXCEPT;
DLAF_EXTERN_C int dlaf_cholesky_factorization_c(const int dlaf_context, const char uplo, dlaf_complex_c* a,
                                                const struct DLAF_descriptor dlaf_desca) DLAF_NOEXCEPT;
DLAF_EXTERN_C int dlaf_cholesky_factorization_z(const int dlaf_context, const char uplo, dlaf_complex_z* a,
                                                const struct DLAF_descriptor dlaf_desca) DLAF_NOEXCEPT;

/* ScaLAPACK-like: desca[0] == 1, desca[1] = context registered with dlaf_create_grid_from_blacs_info (or returned by
 * dlaf_create_grid). Factorizes sub(A) = A(ia : ia+n-1, ja : ja+n-1); `a` points to the local block of the WHOLE
 * distributed matrix. The reference requires ia == ja == 1 (src/c_api/factorization/cholesky.h:67-69); this build
 * accepts what ScaLAPACK's p?potrf accepts: (ia - 1) a multiple of desca[4] and (ja - 1) a multiple of desca[5]. */
DLAF_EXTERN_C void dlaf_pspotrf(const char uplo, const int n, float* a, const int ia, const int ja,
                                const int desca[9], int* info) DLAF_NOEXCEPT;
DLAF_EXTERN_C void dlaf_pdpotrf(const char uplo, const int n, double* a, const int ia, const int ja,
                                const int desca[9], int* info) DLAF_NOEXCEPT;
DLAF_EXTERN_C void dlaf_pcpotrf(const char uplo, const int n, dlaf_complex_c* a, const int ia, const int ja,
                                const int desca[9], int* info) DLAF_NOEXCEPT;
DLAF_EXTERN_C void dlaf_pzpotrf(const char uplo, const int n, dlaf_complex_z* a, const int ia, const int ja,
                                const int desca[9], int* info) DLAF_NOEXCEPT;
