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

// the host pass of these __host__ __device__ bodies does not know the pragma
#ifdef __CUDA_ARCH__
#define DLAFB_UNROLL _Pragma("unroll")
#else
#define DLAFB_UNROLL
#endif

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
  // Placement of the tiles (defaults: the diagonal tiles of one n x n matrix). CTA b inverts GLOBAL diagonal tile
  // g0 + b*gstep (its order is min(nb, n - that*nb)), stored at tile position (row0 + b*row_step, col0 + b*col_step)
  // of t: the diagonal tiles that one rank of a 2D block-cyclic grid owns form such a progression (gstep = lcm(P, Q),
  // row_step = gstep / P, col_step = gstep / Q). count >= 0: number of CTAs (else ceil(n / nb)).
  int g0 = 0, gstep = 1;
  int row0 = 0, row_step = 1;
  int col0 = 0, col_step = 1;
  int count = -1;
};

// Rows of a column that one thread keeps in registers at a time (tri_inverse_body).
template <class T>
struct TriInvChunk {
  static constexpr int value = Tr<T>::is_complex ? 16 : 32;
};

// W_b (nb x nb slot, leading dimension nb, at w + b*nb*nb) <- inv(T_bb), the other triangle of the w x w block set
// to zero (the contraction kernel reads the full block). One column of the inverse per thread at a time, by forward
// substitution on T x = e_j, R = TriInvChunk rows at a time held in registers:
//   acc_r  = (e_j)_i - sum_{k < chunk} t_ik x_k     i = chunk*R + r; t_ik for the R rows of a chunk is contiguous
//                                                   (column k of T) and the same address in every lane of a warp
//   then the R x R triangle of the chunk is solved in registers and the chunk is written once.
// (The first version formed each x_i as a dot product along ROW i of T: one 8-byte load per 32-byte sector, every
// load an L2 round trip - 25 ms for 512 x 512 tiles, 15 % of a whole solve at n = 16384; profiles/r01/
// trsolve_bench_v2_lookahead.jsonl.) An upper tile is the same recurrence on the index-reversed tile (reversing
// rows and columns of an upper triangular matrix gives a lower triangular one whose inverse is the reversed inverse).
template <class T>
__host__ __device__ inline void tri_inverse_body(const TriInvParams p, const T* __restrict__ t, T* __restrict__ w,
                                                 int block, int thread, int nthreads) {
  using R = typename Tr<T>::Real;
  constexpr int RC = TriInvChunk<T>::value;
  const int k0 = (p.g0 + block * p.gstep) * p.nb;
  const int order = p.n - k0 < p.nb ? p.n - k0 : p.nb;
  if (order <= 0)
    return;
  const T* tb = t + (long long)(p.row0 + block * p.row_step) * p.nb +
                (long long)(p.col0 + block * p.col_step) * p.nb * p.ldt;
  T* wb = w + (size_t)block * p.nb * p.nb;
  const int last = order - 1;
  const int upper = p.upper;
  // storage position of logical index a (logical = lower-triangular numbering)
  auto at = [&](int a) { return upper ? last - a : a; };
  // logical rows i0 .. i0 + RC - 1 of logical column k of T are contiguous in storage: ascending for a lower tile,
  // descending for an upper one
  const int dir = upper ? -1 : 1;
  const int nchunks = ceil_div(order, RC);
  for (int j = thread; j < order; j += nthreads) {
    T* xcol = wb + (size_t)at(j) * p.nb;  // column j of the (reversed) inverse
    const int c_first = j / RC;
    for (int i = 0; i < c_first * RC; ++i)
      xcol[at(i)] = Tr<T>::zero();
    for (int c = c_first; c < nchunks; ++c) {
      const int i0 = c * RC;
      const int rows = order - i0 < RC ? order - i0 : RC;  // rows of this chunk inside the tile
      T acc[RC];
DLAFB_UNROLL
      for (int r = 0; r < RC; ++r)
        acc[r] = (i0 + r == j) ? Tr<T>::from_real(R(1)) : Tr<T>::zero();
      // entries of this column solved in earlier chunks
      for (int k = c_first * RC; k < i0; ++k) {
        const T xk = xcol[at(k)];
        const T* tcol = tb + at(i0) + (long long)at(k) * p.ldt;  // t(i0 + r, k) = tcol[dir * r]
        if (rows == RC) {
DLAFB_UNROLL
          for (int r = 0; r < RC; ++r)
            acc[r] = Tr<T>::fnma(tcol[dir * r], xk, acc[r]);
        }
        else {
DLAFB_UNROLL
          for (int r = 0; r < RC; ++r)
            if (r < rows)
              acc[r] = Tr<T>::fnma(tcol[dir * r], xk, acc[r]);
        }
      }
      // the chunk's own triangle, in registers
DLAFB_UNROLL
      for (int r = 0; r < RC; ++r) {
        if (r < rows) {
          const T* tcol = tb + at(i0) + (long long)at(i0 + r) * p.ldt;  // t(i0 + r2, i0 + r) = tcol[dir * r2]
          if (!p.unit)
            acc[r] = Tr<T>::mul(acc[r], tr_recip<T>(tcol[dir * r]));
DLAFB_UNROLL
          for (int r2 = r + 1; r2 < RC; ++r2)
            if (r2 < rows)
              acc[r2] = Tr<T>::fnma(tcol[dir * r2], acc[r], acc[r2]);
        }
      }
DLAFB_UNROLL
      for (int r = 0; r < RC; ++r)
        if (r < rows)
          xcol[at(i0 + r)] = acc[r];
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
