// __global__ wrappers and launchers of the triangular solver's element kernels (bodies: solver_kernels.cuh).
#include "solver_kernels.cuh"

namespace dlafb {
namespace {

template <class T>
__global__ void __launch_bounds__(kTriInvThreads) tri_inverse_kernel(const TriInvParams p, const T* __restrict__ t,
                                                                     T* __restrict__ w) {
  tri_inverse_body<T>(p, t, w, blockIdx.x, threadIdx.x, kTriInvThreads);
}

template <class T>
__global__ void __launch_bounds__(kScaleRows) scale_kernel(int m, int n, T* __restrict__ b, long long ldb, T alpha,
                                                           int conj) {
  scale_body<T>(m, n, b, ldb, alpha, conj, blockIdx.x, blockIdx.y, threadIdx.x);
}

}  // namespace

template <class T>
cudaError_t launch_tri_inverse(cudaStream_t stream, const TriInvParams& p, const T* t, T* w) {
  const int blocks = p.count >= 0 ? p.count : tri_inverse_blocks(p.n, p.nb);
  if (blocks <= 0)
    return cudaSuccess;
  tri_inverse_kernel<T><<<blocks, kTriInvThreads, 0, stream>>>(p, t, w);
  ++g_launches;
  return cudaGetLastError();
}

template <class T>
cudaError_t launch_scale(cudaStream_t stream, int m, int n, T* b, long long ldb, T alpha, int conj) {
  if (m <= 0 || n <= 0)
    return cudaSuccess;
  scale_kernel<T><<<dim3(scale_blocks_x(m), scale_blocks_y(n)), kScaleRows, 0, stream>>>(m, n, b, ldb, alpha, conj);
  ++g_launches;
  return cudaGetLastError();
}

#define DLAFB_INST_SOLVER(T)                                                                          \
  template cudaError_t launch_tri_inverse<T>(cudaStream_t, const TriInvParams&, const T*, T*);        \
  template cudaError_t launch_scale<T>(cudaStream_t, int, int, T*, long long, T, int);
DLAFB_INST_SOLVER(float)
DLAFB_INST_SOLVER(double)
DLAFB_INST_SOLVER(float2)
DLAFB_INST_SOLVER(double2)

}  // namespace dlafb
