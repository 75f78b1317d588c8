// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product; only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load this.
//
// CPU restatement of the reference (DLA-Future v0.6.0) tiled Cholesky path. The reference cannot be
// built here (pika/blaspp/lapackpp/Umpire/MPI missing), so tile arithmetic goes to
// the same class of third-party library the reference MC backend calls (LAPACK ?potrf, BLAS
// ?trsm/?syrk|?herk/?gemm): OpenBLAS 0.3.30 bundled with scipy, resolved with dlopen at run time.
// Parity is pinned on the reference's analytical known-answer tests (tests/test_oracle.py).
//
// Follows:
//   local call_L / call_U            include/dlaf/factorization/cholesky/impl.h:150-189, :316-348
//   tile op arguments                include/dlaf/factorization/cholesky/impl.h:46-146
//   random HPD generator             include/dlaf/util_matrix.h:150-178, :324-447, :499-501
//   residual (max-norm form)         miniapp/miniapp_cholesky.cpp:408-442
//   single-threaded BLAS per task    src/common/single_threaded_blas.cpp:25-36
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include <dlfcn.h>
#include <omp.h>

namespace {

using cfloat = std::complex<float>;
using cdouble = std::complex<double>;

void* g_blas = nullptr;

template <class F>
F sym(const char* name) {
  if (!g_blas) {
    std::fprintf(stderr, "oracle: BLAS library not set (call oracle_set_blas first)\n");
    std::abort();
  }
  void* p = dlsym(g_blas, name);
  if (!p) {
    std::fprintf(stderr, "oracle: missing symbol %s\n", name);
    std::abort();
  }
  return reinterpret_cast<F>(p);
}

// Fortran BLAS/LAPACK prototypes (LP64).
using potrf_t = void (*)(const char*, const int*, void*, const int*, int*);
using trsm_t = void (*)(const char*, const char*, const char*, const char*, const int*, const int*,
                        const void*, const void*, const int*, void*, const int*);
using herk_t = void (*)(const char*, const char*, const int*, const int*, const void*, const void*,
                        const int*, const void*, void*, const int*);
using gemm_t = void (*)(const char*, const char*, const int*, const int*, const int*, const void*,
                        const void*, const int*, const void*, const int*, const void*, void*,
                        const int*);

template <class T>
struct Blas;
#define ORACLE_BLAS(T, R, P, HERK, CT)                                         \
  template <>                                                                  \
  struct Blas<T> {                                                             \
    using real = R;                                                            \
    static constexpr char conj_trans = CT;                                     \
    static potrf_t potrf() { return sym<potrf_t>("scipy_" #P "potrf_"); }      \
    static trsm_t trsm() { return sym<trsm_t>("scipy_" #P "trsm_"); }          \
    static herk_t herk() { return sym<herk_t>("scipy_" #P #HERK "_"); }        \
    static gemm_t gemm() { return sym<gemm_t>("scipy_" #P "gemm_"); }          \
  };
ORACLE_BLAS(float, float, s, syrk, 'T')
ORACLE_BLAS(double, double, d, syrk, 'T')
ORACLE_BLAS(cfloat, float, c, herk, 'C')
ORACLE_BLAS(cdouble, double, z, herk, 'C')

// --- random HPD generator (util_matrix.h:150-178): mt19937_64 + uniform_real_distribution(-1,1) ---
template <class T>
struct Rand {
  std::mt19937_64 eng;
  std::uniform_real_distribution<T> dist{-1, 1};
  explicit Rand(int64_t seed) : eng(static_cast<std::size_t>(seed)) {}
  T operator()() { return dist(eng); }
};
template <class R>
struct Rand<std::complex<R>> {
  Rand<R> r;
  explicit Rand(int64_t seed) : r(seed) {}
  std::complex<R> operator()() {
    // argument evaluation order matters for the random stream: the reference evaluates
    // std::polar(abs(r()), pi * r()) with GCC, which evaluates function arguments right to left.
    // Kept as one expression so this file compiled with GCC reproduces that order.
    return std::polar<R>(std::abs(r()), static_cast<R>(M_PI) * r());
  }
};
template <class T> T conj_(T v) { return v; }
template <class R> std::complex<R> conj_(std::complex<R> v) { return std::conj(v); }
template <class T> double real_(T v) { return static_cast<double>(v); }
template <class R> double real_(std::complex<R> v) { return static_cast<double>(v.real()); }

// Fills the FULL n x n column-major matrix (both triangles), as set_random_hermitian_with_offset
// does over all tiles of a 1x1 distribution (util_matrix.h:399-447); offset = 2n (:499-501).
template <class T>
void gen_hpd(int64_t n, int64_t nb, T* a, int64_t lda) {
  const int64_t nt = (n + nb - 1) / nb;
  const int64_t offset = 2 * n;
#pragma omp parallel for collapse(2) schedule(dynamic)
  for (int64_t tj = 0; tj < nt; ++tj) {
    for (int64_t ti = 0; ti < nt; ++ti) {
      const int64_t r0 = ti * nb, c0 = tj * nb;
      const int64_t tr = std::min(nb, n - r0), tc = std::min(nb, n - c0);
      const int64_t seed = (ti >= tj) ? c0 + r0 * n : r0 + c0 * n;
      Rand<T> rnd(seed);
      T* t = a + r0 + c0 * lda;
      if (ti == tj) {
        for (int64_t j = 0; j < tc; ++j) {
          for (int64_t i = 0; i < j; ++i) {
            T v = rnd();
            t[i + j * lda] = v;
            t[j + i * lda] = conj_(v);
          }
          t[j + j * lda] = static_cast<T>(static_cast<typename Blas<T>::real>(real_(rnd())) +
                                          static_cast<typename Blas<T>::real>(offset));
        }
      }
      else {
        // values are drawn over the FULL tile size (nb x nb) in the same order for the tile and
        // its transposed twin; lower tiles store (i,j), upper tiles store conj at (j,i).
        for (int64_t j = 0; j < nb; ++j) {
          for (int64_t i = 0; i < nb; ++i) {
            T v = rnd();
            if (ti > tj) {
              if (i < tr && j < tc)
                t[i + j * lda] = v;
            }
            else {
              if (j < tr && i < tc)
                t[j + i * lda] = conj_(v);
            }
          }
        }
      }
    }
  }
}

// --- tiled right-looking Cholesky, local (impl.h:150-189 for L, :316-348 for U) ---
// One single-threaded BLAS/LAPACK call per tile task; tasks ordered by data dependencies on tiles,
// which is what the reference's per-tile async_rw_mutex pipelines enforce.
template <class T>
int cholesky_local(char uplo, int64_t n, int64_t nb, T* a, int64_t lda) {
  using B = Blas<T>;
  using R = typename B::real;
  const auto potrf = B::potrf();
  const auto trsm = B::trsm();
  const auto herk = B::herk();
  const auto gemm = B::gemm();
  const int64_t nt = (n + nb - 1) / nb;
  const int ild = static_cast<int>(lda);
  const T one = T(1), mone = T(-1);
  const R rone = R(1), rmone = R(-1);
  const char CT = B::conj_trans;
  int info_out = 0;
  std::vector<char> dep(static_cast<size_t>(nt * nt) + 1);  // dependency tokens, one per tile
  char* d = dep.data();
  auto tile = [&](int64_t i, int64_t j) { return a + i * nb + j * nb * lda; };
  auto ts = [&](int64_t i) { return static_cast<int>(std::min(nb, n - i * nb)); };
  const bool lower = (uplo == 'L' || uplo == 'l');

#pragma omp parallel
#pragma omp single
  {
    for (int64_t k = 0; k < nt; ++k) {
      const int nk = ts(k);
#pragma omp task depend(inout : d[k + k * nt]) priority(2)
      {
        int info = 0;
        potrf(lower ? "L" : "U", &nk, tile(k, k), &ild, &info);  // impl.h:46-53 / :98-105
        if (info != 0) {
#pragma omp critical
          if (info_out == 0)
            info_out = static_cast<int>(k * nb) + info;
        }
      }
      for (int64_t i = k + 1; i < nt; ++i) {
        const int ni = ts(i);
        if (lower) {
#pragma omp task depend(in : d[k + k * nt]) depend(inout : d[i + k * nt]) priority(2)
          trsm("R", "L", &CT, "N", &ni, &nk, &one, tile(k, k), &ild, tile(i, k), &ild);  // :55-67
        }
        else {
#pragma omp task depend(in : d[k + k * nt]) depend(inout : d[k + i * nt]) priority(2)
          trsm("L", "U", &CT, "N", &nk, &ni, &one, tile(k, k), &ild, tile(k, i), &ild);  // :107-119
        }
      }
      for (int64_t j = k + 1; j < nt; ++j) {
        const int nj = ts(j);
        const int prio = (j == k + 1) ? 1 : 0;  // look-ahead priority, impl.h:172-173
        if (lower) {
#pragma omp task depend(in : d[j + k * nt]) depend(inout : d[j + j * nt]) priority(prio)
          herk("L", "N", &nj, &nk, &rmone, tile(j, k), &ild, &rone, tile(j, j), &ild);  // :69-80
        }
        else {
#pragma omp task depend(in : d[k + j * nt]) depend(inout : d[j + j * nt]) priority(prio)
          herk("U", &CT, &nj, &nk, &rmone, tile(k, j), &ild, &rone, tile(j, j), &ild);  // :121-132
        }
        for (int64_t i = j + 1; i < nt; ++i) {
          const int ni = ts(i);
          if (lower) {
#pragma omp task depend(in : d[i + k * nt], d[j + k * nt]) depend(inout : d[i + j * nt]) priority(prio)
            gemm("N", &CT, &ni, &nj, &nk, &mone, tile(i, k), &ild, tile(j, k), &ild, &one, tile(i, j),
                 &ild);  // :82-94
          }
          else {
#pragma omp task depend(in : d[k + j * nt], d[k + i * nt]) depend(inout : d[j + i * nt]) priority(prio)
            gemm(&CT, "N", &nj, &ni, &nk, &mone, tile(k, j), &ild, tile(k, i), &ild, &one, tile(j, i),
                 &ild);  // :134-146
          }
        }
      }
    }
  }
  return info_out;
}

// max-norm residual of the miniapp (miniapp_cholesky.cpp:408-442), Lower only:
// returns max|A - L L^H| over the lower triangle; *norm_a = max|A| over the lower triangle.
// O(n^3): meant for n up to a few thousand.
template <class T>
double residual_lower(int64_t n, const T* a, int64_t lda, const T* l, int64_t ldl, double* norm_a) {
  using B = Blas<T>;
  std::vector<T> lz(static_cast<size_t>(n * n));
  for (int64_t j = 0; j < n; ++j)
    for (int64_t i = 0; i < n; ++i)
      lz[i + j * n] = (i >= j) ? l[i + j * ldl] : T(0);
  std::vector<T> c(static_cast<size_t>(n * n));
  for (int64_t j = 0; j < n; ++j)
    for (int64_t i = 0; i < n; ++i)
      c[i + j * n] = a[i + j * lda];
  const int in = static_cast<int>(n);
  const T one = T(1), mone = T(-1);
  const char CT = B::conj_trans;
  if (n > 0)
    B::gemm()("N", &CT, &in, &in, &in, &mone, lz.data(), &in, lz.data(), &in, &one, c.data(), &in);
  double md = 0, ma = 0;
  for (int64_t j = 0; j < n; ++j)
    for (int64_t i = j; i < n; ++i) {
      md = std::max(md, static_cast<double>(std::abs(c[i + j * n])));
      ma = std::max(ma, static_cast<double>(std::abs(a[i + j * lda])));
    }
  *norm_a = ma;
  return md;
}

}  // namespace

extern "C" {

int oracle_set_blas(const char* path, int blas_threads) {
  g_blas = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!g_blas)
    return 1;
  using setnt_t = void (*)(int);
  sym<setnt_t>("scipy_openblas_set_num_threads")(blas_threads);
  return 0;
}

void oracle_set_threads(int task_threads, int blas_threads) {
  omp_set_num_threads(task_threads);
  using setnt_t = void (*)(int);
  sym<setnt_t>("scipy_openblas_set_num_threads")(blas_threads);
}

int oracle_max_threads() { return omp_get_max_threads(); }

#define ORACLE_EXPORT(X, T)                                                                      \
  void oracle_gen_hpd_##X(int64_t n, int64_t nb, void* a, int64_t lda) {                         \
    gen_hpd<T>(n, nb, static_cast<T*>(a), lda);                                                  \
  }                                                                                              \
  int oracle_cholesky_local_##X(char uplo, int64_t n, int64_t nb, void* a, int64_t lda) {        \
    return cholesky_local<T>(uplo, n, nb, static_cast<T*>(a), lda);                              \
  }                                                                                              \
  double oracle_residual_lower_##X(int64_t n, const void* a, int64_t lda, const void* l,         \
                                   int64_t ldl, double* norm_a) {                                \
    return residual_lower<T>(n, static_cast<const T*>(a), lda, static_cast<const T*>(l), ldl,    \
                             norm_a);                                                            \
  }
ORACLE_EXPORT(s, float)
ORACLE_EXPORT(d, double)
ORACLE_EXPORT(c, cfloat)
ORACLE_EXPORT(z, cdouble)

// plain multithreaded LAPACK dpotrf on the full matrix: sanity ceiling of the CPU baseline
// (BASELINE.md §5).
int oracle_lapack_potrf_d(char uplo, int64_t n, double* a, int64_t lda) {
  int info = 0;
  const int in = static_cast<int>(n), ild = static_cast<int>(lda);
  Blas<double>::potrf()(uplo == 'L' ? "L" : "U", &in, a, &ild, &info);
  return info;
}
}
