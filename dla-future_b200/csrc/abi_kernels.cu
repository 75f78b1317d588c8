// extern "C" entry points of include/dlafb_kernels.h.
#include "../../include/dlafb_kernels.h"

#include "tile_ops.cuh"

using namespace dlafb;

namespace {
template <class T>
T make_scalar(double re, double im);
template <>
float make_scalar<float>(double re, double) { return (float)re; }
template <>
double make_scalar<double>(double re, double) { return re; }
template <>
float2 make_scalar<float2>(double re, double im) { return make_float2((float)re, (float)im); }
template <>
double2 make_scalar<double2>(double re, double im) { return make_double2(re, im); }

inline cudaStream_t st(void* s) { return static_cast<cudaStream_t>(s); }
}  // namespace

extern "C" {

int dlafb_inner_block(char type, const void* a, int64_t lda, int nb) {
  switch (type) {
    case 'd': return inner_block<double>(static_cast<const double*>(a), lda, nb);
    case 's': return inner_block<float>(static_cast<const float*>(a), lda, nb);
    case 'c': return inner_block<float2>(static_cast<const float2*>(a), lda, nb);
    case 'z': return inner_block<double2>(static_cast<const double2*>(a), lda, nb);
  }
  return -1;
}

size_t dlafb_potrf_work_elems(char, int n, int ib) { return potrf_work_elems<double>(n, ib); }

#define DLAFB_DEFINE_TILE_API(X, CT, T)                                                                            \
  int dlafb_potrf_tile_##X(void* s, char uplo, int n, CT* a, int64_t lda, CT* work, int ib, int* info, int base) { \
    return (int)potrf_tile<T>(st(s), uplo, n, (T*)a, lda, (T*)work, ib, info, base);                              \
  }                                                                                                                \
  int dlafb_trsm_panel_##X(void* s, int nk, const CT* l, int64_t ldl, const CT* w, int ib, CT* b, int64_t ldb,    \
                           int m) {                                                                                \
    return (int)trsm_panel<T>(st(s), nk, (const T*)l, ldl, (const T*)w, ib, (T*)b, ldb, m);                       \
  }                                                                                                                \
  int dlafb_trsm_panel_packed_##X(void* s, int nk, const CT* l, int64_t ldl, const CT* w, int ib, CT* b,          \
                                  int64_t ldb, int m, CT* packed, int tile) {                                      \
    return (int)trsm_panel_packed<T>(st(s), nk, (const T*)l, ldl, (const T*)w, ib, (T*)b, ldb, m, (T*)packed,     \
                                     tile);                                                                        \
  }                                                                                                                \
  int dlafb_gemm_nt_##X(void* s, int m, int n, int k, double alpha, const CT* a, int64_t lda, const CT* b,        \
                        int64_t ldb, double beta, CT* c, int64_t ldc, int lower) {                                 \
    return (int)gemm_nt<T>(st(s), m, n, k, (Tr<T>::Real)alpha, (const T*)a, lda, (const T*)b, ldb,                \
                           (Tr<T>::Real)beta, (T*)c, ldc, lower);                                                  \
  }                                                                                                                \
  int dlafb_transpose_##X(void* s, int m, int n, const CT* in, int64_t ldin, CT* out, int64_t ldout, int conj,    \
                          char tri) {                                                                              \
    return (int)transpose<T>(st(s), m, n, (const T*)in, ldin, (T*)out, ldout, conj, tri);                         \
  }                                                                                                                \
  int dlafb_laset_##X(void* s, char uplo, int m, int n, double ar, double ai, double br, double bi, CT* a,        \
                      int64_t lda) {                                                                               \
    return (int)laset<T>(st(s), uplo, m, n, make_scalar<T>(ar, ai), make_scalar<T>(br, bi), (T*)a, lda);          \
  }                                                                                                                \
  int dlafb_lacpy_##X(void* s, char uplo, int m, int n, const CT* a, int64_t lda, CT* b, int64_t ldb) {           \
    return (int)lacpy<T>(st(s), uplo, m, n, (const T*)a, lda, (T*)b, ldb);                                        \
  }                                                                                                                \
  int dlafb_max_abs_##X(void* s, char uplo, int m, int n, const CT* a, int64_t lda, double* out) {                \
    return (int)max_abs<T>(st(s), uplo, m, n, (const T*)a, lda, out);                                             \
  }                                                                                                                \
  int dlafb_gen_hpd_##X(void* s, int64_t n, int nb, int rows, int cols, int rf, int rs, int cf, int cs, CT* a,    \
                        int64_t lda, uint64_t seed) {                                                              \
    return (int)gen_hpd<T>(st(s), n, nb, rows, cols, rf, rs, cf, cs, (T*)a, lda, seed);                           \
  }

DLAFB_DEFINE_TILE_API(s, float, float)
DLAFB_DEFINE_TILE_API(d, double, double)
DLAFB_DEFINE_TILE_API(c, void, float2)
DLAFB_DEFINE_TILE_API(z, void, double2)

}  // extern "C"
