// Element-wise helpers around the Cholesky path: transpose (Upper <-> Lower views), laset / lacpy with uplo
// (reference: src/lapack/gpu/laset.cu:66-120, src/lapack/gpu/lacpy.cu:69-209), the synthetic HPD generator used
// by timing runs, and max-norm reduction (reference: include/dlaf/auxiliary/norm/mc.h:35-95).
// All are HBM-bound: 32x32 shared-memory staging for the transpose, coalesced column walks elsewhere.
#include "tile_ops.cuh"

namespace dlafb {

std::atomic<unsigned long long> g_launches{0};

namespace {

template <class T>
__global__ void __launch_bounds__(256) transpose_kernel(int m, int n, const T* __restrict__ in, long long ldin,
                                                        T* __restrict__ out, long long ldout, int conj, int tri) {
  __shared__ T tile[32][33];
  const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = bi + tx, j = bj + ty + 8 * k;
    if (i < m && j < n)
      tile[ty + 8 * k][tx] = in[i + (long long)j * ldin];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    // out(j, i) = op(in(i, j)); out row index oj = bj + tx, out col index oi = bi + ty + 8k
    const int oj = bj + tx, oi = bi + ty + 8 * k;
    if (oj < n && oi < m) {
      if ((tri == 1 && oj < oi) || (tri == 2 && oj > oi))
        continue;
      T v = tile[tx][ty + 8 * k];
      out[oj + (long long)oi * ldout] = conj ? Tr<T>::conj(v) : v;
    }
  }
}

template <class T>
__global__ void __launch_bounds__(256) laset_kernel(int tri, int m, int n, T alpha, T beta, T* a, long long lda) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= m)
    return;
  const int j0 = blockIdx.y * 64;
  const int j1 = min(n, j0 + 64);
  for (int j = j0; j < j1; ++j) {
    if ((tri == 1 && i < j) || (tri == 2 && i > j))
      continue;
    a[i + (long long)j * lda] = (i == j) ? beta : alpha;
  }
}

template <class T>
__global__ void __launch_bounds__(256) lacpy_kernel(int tri, int m, int n, const T* __restrict__ a, long long lda,
                                                    T* __restrict__ b, long long ldb) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= m)
    return;
  const int j0 = blockIdx.y * 64;
  const int j1 = min(n, j0 + 64);
  for (int j = j0; j < j1; ++j) {
    if ((tri == 1 && i < j) || (tri == 2 && i > j))
      continue;
    b[i + (long long)j * ldb] = a[i + (long long)j * lda];
  }
}

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ double u11(unsigned long long h) {  // uniform in [-1, 1)
  return (double)(h >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}

template <class T>
__device__ __forceinline__ T hpd_value(long long gi, long long gj, long long n, unsigned long long seed) {
  using R = typename Tr<T>::Real;
  const long long hi = gi > gj ? gi : gj, lo = gi > gj ? gj : gi;
  const unsigned long long h = mix64(seed ^ mix64((unsigned long long)hi * (unsigned long long)n + lo));
  if (gi == gj)
    return Tr<T>::from_real((R)(u11(h) + 2.0 * (double)n));
  if constexpr (Tr<T>::is_complex) {
    // polar(|u1|, pi*u2) for the lower element, its conjugate above the diagonal
    const double rad = fabs(u11(h));
    const double ang = 3.14159265358979323846 * u11(mix64(h));
    T v;
    v.x = (R)(rad * cos(ang));
    v.y = (R)(rad * sin(ang));
    return gi > gj ? v : Tr<T>::conj(v);
  }
  else {
    return (T)u11(h);
  }
}

template <class T>
__global__ void __launch_bounds__(256) gen_hpd_kernel(long long n, int nb, int rows, int cols, int row_first,
                                                      int row_stride, int col_first, int col_stride, T* a,
                                                      long long lda, unsigned long long seed) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= rows)
    return;
  const long long gi = ((long long)(i / nb) * row_stride + row_first) * nb + i % nb;
  const int j0 = blockIdx.y * 64;
  const int j1 = min(cols, j0 + 64);
  for (int j = j0; j < j1; ++j) {
    const long long gj = ((long long)(j / nb) * col_stride + col_first) * nb + j % nb;
    a[i + (long long)j * lda] = hpd_value<T>(gi, gj, n, seed);
  }
}

template <class T>
__global__ void __launch_bounds__(256) max_abs_kernel(int tri, int m, int n, const T* __restrict__ a, long long lda,
                                                      double* out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  const int j0 = blockIdx.y * 64;
  const int j1 = min(n, j0 + 64);
  double v = 0.0;
  if (i < m)
    for (int j = j0; j < j1; ++j) {
      if ((tri == 1 && i < j) || (tri == 2 && i > j))
        continue;
      const double x = (double)Tr<T>::abs(a[i + (long long)j * lda]);
      v = (x > v || x != x) ? x : v;  // NaN propagates
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double w = __shfl_xor_sync(0xffffffffu, v, o);
    v = (w > v || w != w) ? w : v;
  }
  if ((threadIdx.x & 31) == 0) {
    if (v != v)
      v = __longlong_as_double(0x7ff0000000000000ll);  // report NaN as +inf: ordering of the integer max below
    atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(v));
  }
}

template <class T>
__global__ void __launch_bounds__(256) pack_tiles_kernel(int m, int n, const T* __restrict__ src, long long ld,
                                                         T* __restrict__ dst, int tile) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= m)
    return;
  T* d = dst + (long long)(i / tile) * tile * tile + i % tile;
  const int j0 = blockIdx.y * 64;
  const int j1 = min(n, j0 + 64);
  for (int j = j0; j < j1; ++j)
    d[(long long)j * tile] = src[i + (long long)j * ld];
}

inline int tri_code(char c) {
  return (c == 'L' || c == 'l') ? 1 : ((c == 'U' || c == 'u') ? 2 : 0);
}

}  // namespace

template <class T>
cudaError_t transpose(cudaStream_t stream, int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj,
                      char tri) {
  if (m <= 0 || n <= 0)
    return cudaSuccess;
  transpose_kernel<T><<<dim3(ceil_div(m, 32), ceil_div(n, 32)), 256, 0, stream>>>(m, n, in, ldin, out, ldout, conj,
                                                                                 tri_code(tri));
  ++g_launches;
  return cudaGetLastError();
}

template <class T>
cudaError_t laset(cudaStream_t stream, char uplo, int m, int n, T alpha, T beta, T* a, long long lda) {
  if (m <= 0 || n <= 0)
    return cudaSuccess;
  laset_kernel<T><<<dim3(ceil_div(m, 256), ceil_div(n, 64)), 256, 0, stream>>>(tri_code(uplo), m, n, alpha, beta, a,
                                                                              lda);
  ++g_launches;
  return cudaGetLastError();
}

template <class T>
cudaError_t lacpy(cudaStream_t stream, char uplo, int m, int n, const T* a, long long lda, T* b, long long ldb) {
  if (m <= 0 || n <= 0)
    return cudaSuccess;
  lacpy_kernel<T><<<dim3(ceil_div(m, 256), ceil_div(n, 64)), 256, 0, stream>>>(tri_code(uplo), m, n, a, lda, b, ldb);
  ++g_launches;
  return cudaGetLastError();
}

template <class T>
cudaError_t gen_hpd(cudaStream_t stream, long long n, int nb, int rows, int cols, int row_first, int row_stride,
                    int col_first, int col_stride, T* a, long long lda, unsigned long long seed) {
  if (rows <= 0 || cols <= 0)
    return cudaSuccess;
  gen_hpd_kernel<T><<<dim3(ceil_div(rows, 256), ceil_div(cols, 64)), 256, 0, stream>>>(
      n, nb, rows, cols, row_first, row_stride, col_first, col_stride, a, lda, seed);
  ++g_launches;
  return cudaGetLastError();
}

template <class T>
cudaError_t max_abs(cudaStream_t stream, char uplo, int m, int n, const T* a, long long lda, double* out) {
  if (m <= 0 || n <= 0)
    return cudaSuccess;
  max_abs_kernel<T><<<dim3(ceil_div(m, 256), ceil_div(n, 64)), 256, 0, stream>>>(tri_code(uplo), m, n, a, lda, out);
  ++g_launches;
  return cudaGetLastError();
}

template <class T>
cudaError_t pack_tiles(cudaStream_t stream, int m, int n, const T* src, long long ld, T* dst, int tile) {
  if (m <= 0 || n <= 0)
    return cudaSuccess;
  pack_tiles_kernel<T><<<dim3(ceil_div(m, 256), ceil_div(n, 64)), 256, 0, stream>>>(m, n, src, ld, dst, tile);
  ++g_launches;
  return cudaGetLastError();
}

#define DLAFB_INST_AUX(T)                                                                                          \
  template cudaError_t transpose<T>(cudaStream_t, int, int, const T*, long long, T*, long long, int, char);        \
  template cudaError_t laset<T>(cudaStream_t, char, int, int, T, T, T*, long long);                                \
  template cudaError_t lacpy<T>(cudaStream_t, char, int, int, const T*, long long, T*, long long);                 \
  template cudaError_t gen_hpd<T>(cudaStream_t, long long, int, int, int, int, int, int, int, T*, long long,       \
                                  unsigned long long);                                                             \
  template cudaError_t max_abs<T>(cudaStream_t, char, int, int, const T*, long long, double*);                      \
  template cudaError_t pack_tiles<T>(cudaStream_t, int, int, const T*, long long, T*, int);
DLAFB_INST_AUX(float)
DLAFB_INST_AUX(double)
DLAFB_INST_AUX(float2)
DLAFB_INST_AUX(double2)

}  // namespace dlafb
