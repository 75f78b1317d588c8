// Generic (any element type, any size, any alignment) contraction kernel: C <- beta*C + alpha * A * B^H.
// Used for s/c/z and for double operands that do not meet the TMA alignment rules of gemm_dmma.cu
// (odd leading dimension, block sizes that are not a multiple of 16: the reference's own test shapes,
// test/unit/factorization/test_cholesky.cpp:54-62, are mostly of that kind).
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"

namespace dlafb {

constexpr int kSimtTile = 64;  // C block per CTA (rows and cols)

template <class T>
struct SimtParams {
  T* C;
  long long ldc;
  const T* A;  // already offset to the first K column
  long long lda;
  const T* B;
  long long ldb;
  int m_total, n_total;
  int K;
  int a_row_off, b_row_off;
  typename Tr<T>::Real alpha, beta;
  const int2* list;  // LIST mode as in GemmParams, with kSimtTile-sized sub-blocks
  int sub;
  int tile;  // nb: tiles of the list are clipped to tile x tile
  int rect_row0, rect_col0;
  int lower;
  int a_tiled, b_tiled;  // operand stored as contiguous tile x tile tiles (lda / ldb = tile)
};

template <class T>
cudaError_t launch_gemm_simt(cudaStream_t stream, const SimtParams<T>& p, unsigned nblocks_x, unsigned nblocks_y);

}  // namespace dlafb
