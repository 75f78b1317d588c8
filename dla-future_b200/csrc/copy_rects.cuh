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
  auto advance = [&]() {
    row += drow;
    col += dcol;
    if (row >= r.rows) {
      row -= r.rows;
      ++col;
    }
  };
  long long idx = thread;
  // four independent loads in flight per thread before the first store (a 64 x 64 double rectangle is 16 elements per
  // thread: one load at a time would leave the copy latency bound)
  while (idx + 3LL * nthreads < total) {
    const long long o0s = row + col * r.src_ld, o0d = row + col * r.dst_ld;
    advance();
    const long long o1s = row + col * r.src_ld, o1d = row + col * r.dst_ld;
    advance();
    const long long o2s = row + col * r.src_ld, o2d = row + col * r.dst_ld;
    advance();
    const long long o3s = row + col * r.src_ld, o3d = row + col * r.dst_ld;
    advance();
    const T v0 = s[o0s], v1 = s[o1s], v2 = s[o2s], v3 = s[o3s];
    d[o0d] = v0;
    d[o1d] = v1;
    d[o2d] = v2;
    d[o3d] = v3;
    idx += 4LL * nthreads;
  }
  while (idx < total) {
    d[row + col * r.dst_ld] = s[row + col * r.src_ld];
    advance();
    idx += nthreads;
  }
}

// rects: DEVICE array of n rectangles. Empty rectangles are not allowed in the list (rows, cols >= 1).
template <class T>
cudaError_t launch_copy_rects(cudaStream_t stream, int n, const CopyRect* rects, const T* src, T* dst);

}  // namespace dlafb
