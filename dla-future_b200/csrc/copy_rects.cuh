// Gather / scatter of a list of column-major rectangles: the one kernel of the matrix redistribution
// (src/redistribute.h; reference: matrix::copy(src, dst, grid), include/dlaf/matrix/copy.h:118-231, which posts one
// MPI send / receive per tile). HBM-bound: every element is read once and written once, 2 s bytes per element.
//
// One CTA per rectangle; thread t copies elements t, t + nthreads, ... of the rectangle in column-major order, so the
// threads of a warp read and write consecutive addresses of one column (a 64-row double tile is one 512-byte run).
// The body is a __host__ __device__ function of (block, thread) without shared memory, so tests/emu/ runs exactly this
// code on the CPU.
#pragma once
#include <cuda_runtime.h>

#include "common.cuh"

namespace dlafb {

constexpr int kCopyRectThreads = 256;

struct CopyRect {
  long long src_off, dst_off;  // element offsets into the source / destination buffers
  long long src_ld, dst_ld;    // leading dimensions (a packed rectangle has ld = rows)
  int rows, cols;
};

template <class T>
__host__ __device__ inline void copy_rects_body(const CopyRect* __restrict__ rects, const T* __restrict__ src,
                                                T* __restrict__ dst, int block, int thread, int nthreads) {
  const CopyRect r = rects[block];
  const long long total = (long long)r.rows * r.cols;
  const T* s = src + r.src_off;
  T* d = dst + r.dst_off;
  // walk the columns with an incremental (row, col) pair: no division per element
  int row = thread % r.rows;
  long long col = thread / r.rows;
  const int drow = nthreads % r.rows;
  const long long dcol = nthreads / r.rows;
  for (long long idx = thread; idx < total; idx += nthreads) {
    d[row + col * r.dst_ld] = s[row + col * r.src_ld];
    row += drow;
    col += dcol;
    if (row >= r.rows) {
      row -= r.rows;
      ++col;
    }
  }
}

// rects: DEVICE array of n rectangles. Empty rectangles are not allowed in the list (rows, cols >= 1).
template <class T>
cudaError_t launch_copy_rects(cudaStream_t stream, int n, const CopyRect* rects, const T* src, T* dst);

}  // namespace dlafb
