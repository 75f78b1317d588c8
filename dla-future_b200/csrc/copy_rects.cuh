// Gather / scatter of a list of column-major rectangles: the one kernel of the matrix redistribution
// (src/redistribute.h; reference: matrix::copy(src, dst, grid), include/dlaf/matrix/copy.h:118-231, which posts one
// MPI send / receive per tile). HBM-bound: every element is read once and written once, 2 s bytes per element.
//
// One CTA per rectangle; thread t copies elements t, t + nthreads, ... of the rectangle in column-major order, so the
// threads of a warp read and write consecutive addresses of one column, 16 bytes per thread where alignment allows
// (a 64-row double tile column is one 512-byte run = one warp instruction).
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

// Column-major rectangle of `rows` x `cols` elements of V: thread t copies elements t, t + nthreads, ...
template <class V>
__host__ __device__ inline void copy_rect_loop(const V* __restrict__ s, V* __restrict__ d, int rows, long long cols,
                                               long long src_ld, long long dst_ld, int thread, int nthreads) {
  const long long total = (long long)rows * cols;
  // walk the columns with an incremental (row, col) pair: no division per element
  int row = thread % rows;
  long long col = thread / rows;
  const int drow = nthreads % rows;
  const long long dcol = nthreads / rows;
  auto advance = [&]() {
    row += drow;
    col += dcol;
    if (row >= rows) {
      row -= rows;
      ++col;
    }
  };
  long long idx = thread;
  // four independent loads in flight per thread before the first store (a 64 x 64 double rectangle is 8 vector
  // elements per thread: one load at a time would leave the copy latency bound)
  while (idx + 3LL * nthreads < total) {
    const long long o0s = row + col * src_ld, o0d = row + col * dst_ld;
    advance();
    const long long o1s = row + col * src_ld, o1d = row + col * dst_ld;
    advance();
    const long long o2s = row + col * src_ld, o2d = row + col * dst_ld;
    advance();
    const long long o3s = row + col * src_ld, o3d = row + col * dst_ld;
    advance();
    const V v0 = s[o0s], v1 = s[o1s], v2 = s[o2s], v3 = s[o3s];
    d[o0d] = v0;
    d[o1d] = v1;
    d[o2d] = v2;
    d[o3d] = v3;
    idx += 4LL * nthreads;
  }
  while (idx < total) {
    d[row + col * dst_ld] = s[row + col * src_ld];
    advance();
    idx += nthreads;
  }
}

template <class T>
__host__ __device__ inline void copy_rects_body(const CopyRect* __restrict__ rects, const T* __restrict__ src,
                                                T* __restrict__ dst, int block, int thread, int nthreads) {
  const CopyRect r = rects[block];
  const T* s = src + r.src_off;
  T* d = dst + r.dst_off;
  // 16-byte accesses whenever the rectangle allows it (rows, leading dimensions and both addresses multiples of 16
  // bytes - the usual case: tile sizes are even and buffers come from cudaMalloc): a warp then moves 512 bytes per
  // instruction, one 64-row double column per warp
  constexpr int VW = (int)(16 / sizeof(T));
  if (VW > 1 && r.rows % VW == 0 && r.src_ld % VW == 0 && r.dst_ld % VW == 0 &&
      (reinterpret_cast<uintptr_t>(s) & 15) == 0 && (reinterpret_cast<uintptr_t>(d) & 15) == 0) {
    copy_rect_loop<float4>(reinterpret_cast<const float4*>(s), reinterpret_cast<float4*>(d), r.rows / VW, r.cols,
                           r.src_ld / VW, r.dst_ld / VW, thread, nthreads);
    return;
  }
  copy_rect_loop<T>(s, d, r.rows, r.cols, r.src_ld, r.dst_ld, thread, nthreads);
}

// rects: DEVICE array of n rectangles. Empty rectangles are not allowed in the list (rows, cols >= 1).
template <class T>
cudaError_t launch_copy_rects(cudaStream_t stream, int n, const CopyRect* rects, const T* src, T* dst);

}  // namespace dlafb
