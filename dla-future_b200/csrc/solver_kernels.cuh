// Element kernels of the triangular solver (reference: include/dlaf/solver/triangular/impl.h).
// The solver turns every side/uplo/op case into contractions of the C -= A B^H kernel (gemm_dmma.cu); what is left
// for this file is O(n^2) work: inverting the diagonal tiles of the triangular matrix and scaling B.
//
// Each kernel body is a __host__ __device__ function of (block index, thread index) with no shared memory, barrier
// or shuffle, so that tests/emu/ can run exactly this code on the CPU (one call per thread of the launch geometry
// below) and compare it with the oracle where no GPU is present. The __global__ wrappers in solver_kernels.cu only
// forward blockIdx / threadIdx.
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"

namespace dlafb {

constexpr int kTriInvThreads = 256;  // threads per CTA of tri_inverse: thread t owns columns t, t + 256, ...
constexpr int kScaleRows = 256;      // scale: rows per CTA (one per thread)
constexpr int kScaleCols = 64;       // scale: columns per CTA (walked by each thread)

template <class T>
__host__ __device__ inline T tr_neg(T a) {
  return Tr<T>::scale(a, typename Tr<T>::Real(-1));
}

// 1 / a
template <class T>
__host__ __device__ inline T tr_recip(T a) {
  using R = typename Tr<T>::Real;
  if constexpr (Tr<T>::is_complex) {
    // scaled to stay clear of overflow in |a|^2
    const R s = (a.x < 0 ? -a.x : a.x) + (a.y < 0 ? -a.y : a.y);
    const R xr = a.x / s, xi = a.y / s;
    const R d = (xr * xr + xi * xi) * s;
    return Tr<T>::make(xr / d, -xi / d);
  }
  else {
    return R(1) / a;
  }
}

template <class T>
__host__ __device__ inline bool tr_is_one(T a) {
  if constexpr (Tr<T>::is_complex)
    return a.x == 1 && a.y == 0;
  else
    return a == 1;
}

struct TriInvParams {
  int n;           // order of the triangular matrix
  int nb;          // tile size; CTA b inverts the diagonal tile starting at (b*nb, b*nb), of order min(nb, n - b*nb)
  long long ldt;   // leading dimension of the triangular matrix
  int upper;       // 0: lower triangular, 1: upper
  int unit;        // 1: the diagonal is not referenced and taken as 1
};

// W_b (nb x nb slot, leading dimension nb, at w + b*nb*nb) <- inv(T_bb), the other triangle of the w x w block set
// to zero (the contraction kernel reads the full block). Column-wise substitution, one column per thread at a time:
//   x_jj = 1 / t_jj,   x_ij = -(sum_{k=j}^{i-1} t_ik x_kj) / t_ii   (i > j)
// for a lower tile; an upper tile is the same recurrence on the index-reversed tile (reversing rows and columns of
// an upper triangular matrix gives a lower triangular one whose inverse is the reversed inverse).
template <class T>
__host__ __device__ inline void tri_inverse_body(const TriInvParams p, const T* __restrict__ t, T* __restrict__ w,
                                                 int block, int thread, int nthreads) {
  using R = typename Tr<T>::Real;
  const int k0 = block * p.nb;
  const int order = p.n - k0 < p.nb ? p.n - k0 : p.nb;
  if (order <= 0)
    return;
  const T* tb = t + k0 + (long long)k0 * p.ldt;
  T* wb = w + (size_t)block * p.nb * p.nb;
  const int last = order - 1;
  // position of logical index a (lower-triangular numbering) in storage
  auto at = [&](int a) { return p.upper ? last - a : a; };
  for (int j = thread; j < order; j += nthreads) {
    T* xcol = wb + (size_t)at(j) * p.nb;  // column j of the (reversed) inverse
    for (int i = 0; i < j; ++i)
      xcol[at(i)] = Tr<T>::zero();
    const T djj = p.unit ? Tr<T>::from_real(R(1)) : tr_recip<T>(tb[at(j) + (long long)at(j) * p.ldt]);
    xcol[at(j)] = djj;
    for (int i = j + 1; i < order; ++i) {
      const T* trow = tb + at(i);  // t(i, k) = trow[at(k) * ldt]
      T s0 = Tr<T>::zero(), s1 = Tr<T>::zero(), s2 = Tr<T>::zero(), s3 = Tr<T>::zero();
      int k = j;
      for (; k + 3 < i; k += 4) {
        s0 = Tr<T>::fma(trow[(long long)at(k) * p.ldt], xcol[at(k)], s0);
        s1 = Tr<T>::fma(trow[(long long)at(k + 1) * p.ldt], xcol[at(k + 1)], s1);
        s2 = Tr<T>::fma(trow[(long long)at(k + 2) * p.ldt], xcol[at(k + 2)], s2);
        s3 = Tr<T>::fma(trow[(long long)at(k + 3) * p.ldt], xcol[at(k + 3)], s3);
      }
      for (; k < i; ++k)
        s0 = Tr<T>::fma(trow[(long long)at(k) * p.ldt], xcol[at(k)], s0);
      T s = Tr<T>::add(Tr<T>::add(s0, s1), Tr<T>::add(s2, s3));
      s = tr_neg<T>(s);
      if (!p.unit)
        s = Tr<T>::mul(s, tr_recip<T>(trow[(long long)at(i) * p.ldt]));
      xcol[at(i)] = s;
    }
  }
}

// b(i, j) <- alpha * (conj ? conj(b(i, j)) : b(i, j)) on the m x n block; CTA (bx, by) covers rows
// bx*kScaleRows + thread and columns by*kScaleCols ... + kScaleCols.
template <class T>
__host__ __device__ inline void scale_body(int m, int n, T* __restrict__ b, long long ldb, T alpha, int conj,
                                           int bx, int by, int thread) {
  const int i = bx * kScaleRows + thread;
  if (i >= m)
    return;
  const int j0 = by * kScaleCols;
  const int j1 = n < j0 + kScaleCols ? n : j0 + kScaleCols;
  for (int j = j0; j < j1; ++j) {
    T v = b[i + (long long)j * ldb];
    if (conj)
      v = Tr<T>::conj(v);
    b[i + (long long)j * ldb] = Tr<T>::mul(alpha, v);
  }
}

// Launch geometry shared by the device launchers and the CPU emulation.
inline int tri_inverse_blocks(int n, int nb) { return n > 0 ? ceil_div(n, nb) : 0; }
inline int scale_blocks_x(int m) { return ceil_div(m, kScaleRows); }
inline int scale_blocks_y(int n) { return ceil_div(n, kScaleCols); }

template <class T>
cudaError_t launch_tri_inverse(cudaStream_t stream, const TriInvParams& p, const T* t, T* w);

template <class T>
cudaError_t launch_scale(cudaStream_t stream, int m, int n, T* b, long long ldb, T alpha, int conj);

}  // namespace dlafb
