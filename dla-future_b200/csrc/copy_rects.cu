// __global__ wrapper and launcher of the rectangle-list copy (body: copy_rects.cuh).
#include "copy_rects.cuh"

namespace dlafb {
namespace {

template <class T>
__global__ void __launch_bounds__(kCopyRectThreads) copy_rects_kernel(const CopyRect* __restrict__ rects,
                                                                      const T* __restrict__ src, T* __restrict__ dst) {
  copy_rects_body<T>(rects, src, dst, blockIdx.x, threadIdx.x, kCopyRectThreads);
}

}  // namespace

template <class T>
cudaError_t launch_copy_rects(cudaStream_t stream, int n, const CopyRect* rects, const T* src, T* dst) {
  // gridDim.x limit is 2^31 - 1: one launch is always enough for the rectangles of one local block
  if (n <= 0)
    return cudaSuccess;
  copy_rects_kernel<T><<<(unsigned)n, kCopyRectThreads, 0, stream>>>(rects, src, dst);
  ++g_launches;
  return cudaGetLastError();
}

template cudaError_t launch_copy_rects<float>(cudaStream_t, int, const CopyRect*, const float*, float*);
template cudaError_t launch_copy_rects<double>(cudaStream_t, int, const CopyRect*, const double*, double*);
template cudaError_t launch_copy_rects<float2>(cudaStream_t, int, const CopyRect*, const float2*, float2*);
template cudaError_t launch_copy_rects<double2>(cudaStream_t, int, const CopyRect*, const double2*, double2*);

}  // namespace dlafb
