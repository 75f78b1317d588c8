/* Common definitions of the C API (reference: include/dlaf_c/utils.h). */
#pragma once

#ifdef __cplusplus
#define DLAF_EXTERN_C extern "C"
#define DLAF_NOEXCEPT noexcept
#else
#define DLAF_EXTERN_C
#define DLAF_NOEXCEPT
#endif

#ifdef __cplusplus
#include <complex>
using dlaf_complex_c = std::complex<float>;
using dlaf_complex_z = std::complex<double>;
#else
#include <complex.h>
typedef float complex dlaf_complex_c;
typedef double complex dlaf_complex_z;
#endif

#include <dlaf_c/desc.h>

/* ScaLAPACK descriptor (desc[0] == 1 dense) -> DLAF descriptor. i, j are ScaLAPACK's 1-based (ia, ja) and are
 * stored 0-based in the result, exactly as the reference does (src/c_api/utils.cpp:25-33, which stops at
 * i == j == 1; here any i, j >= 1 whose 0-based value is a multiple of the block size is accepted downstream). */
DLAF_EXTERN_C struct DLAF_descriptor make_dlaf_descriptor(const int m, const int n, const int i, const int j,
                                                          const int desc[9]) DLAF_NOEXCEPT;
