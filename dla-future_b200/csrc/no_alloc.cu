// dlafb_potrf_tile_noalloc_d: the diagonal-tile Cholesky (potrf_tile<double>, tile_ops.cu) with its workspace in a
// file-scope device array instead of a caller-supplied buffer, for callers that cannot allocate device memory (code
// under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// Nothing here calls cudaMalloc / cudaFree or an allocating library routine.
//
// The array is part of the library image: every GPU that loads the library holds kScratchDoubles * 8 bytes (2.5 MiB)
// until the process exits. All calls share it and it keeps the last call's contents: the part a call uses is cleared
// on the caller's stream first, and two calls whose work overlaps on the GPU (different streams without an ordering
// between them, or several host threads) RACE on it - order them.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "../../include/dlafb_kernels.h"
#include "tile_ops.cuh"

using namespace dlafb;

namespace {

constexpr int kMaxN = DLAFB_POTRF_NOALLOC_MAX_N;  // 512
constexpr int kMinIb = 64;
// potrf_work_elems(n, ib) = ceil(n / ib) * ib * ib + n * n, largest for ib = 128 at n = 512
constexpr size_t kScratchDoubles = (size_t)(kMaxN / 128) * 128 * 128 + (size_t)kMaxN * kMaxN;
static_assert((size_t)(kMaxN / kMinIb) * kMinIb * kMinIb <= (size_t)(kMaxN / 128) * 128 * 128, "ib = 64 needs less");

static __device__ __align__(256) double g_potrf_scratch[kScratchDoubles];

// In host code `g_potrf_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes
// from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kMaxDevices = 64;
double* g_scratch_ptr[kMaxDevices] = {};
std::mutex g_scratch_mutex;

cudaError_t scratch_for_current_device(double** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return e;
  if (dev < 0 || dev >= kMaxDevices)
    return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_ptr[dev] == nullptr) {
    double* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_potrf_scratch);
    if (e != cudaSuccess)
      return e;
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return cudaSuccess;
}

}  // namespace

extern "C" int dlafb_potrf_tile_noalloc_d(void* stream, char uplo, int n, double* a, int64_t lda, int* info_dev,
                                          int info_base) {
  if (n < 0 || (n > 0 && (a == nullptr || lda < n)))
    return (int)cudaErrorInvalidValue;
  if (n > kMaxN) {
    std::fprintf(stderr,
                 "[dlaf_b200] dlafb_potrf_tile_noalloc_d: n = %d exceeds the built-in scratch (DLAFB_POTRF_NOALLOC_MAX_N"
                 " = %d); use dlafb_potrf_tile_d with a caller-supplied workspace\n",
                 n, kMaxN);
    return (int)cudaErrorInvalidValue;
  }
  if (n == 0)
    return 0;
  double* work = nullptr;
  cudaError_t e = scratch_for_current_device(&work);
  if (e != cudaSuccess)
    return (int)e;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int ib = inner_block<double>(a, lda, n);
  const size_t used = potrf_work_elems<double>(n, ib);
  if (used > kScratchDoubles)
    return (int)cudaErrorInvalidValue;
  // the array keeps the previous call's contents: clear the part this call works in
  if ((e = cudaMemsetAsync(work, 0, used * sizeof(double), s)) != cudaSuccess)
    return (int)e;
  return (int)potrf_tile<double>(s, uplo, n, a, lda, work, ib, info_dev, info_base);
}
