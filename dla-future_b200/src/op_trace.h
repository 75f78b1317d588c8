// Debugging aid of the device executors (triangular_solver.cu, gen_to_std.cu, dist_device_ops.h): with DLAFB_TRACE=1
// every operation an algorithm template issues is printed to stderr (name, dimensions, operand addresses) before it
// is enqueued; with DLAFB_TRACE=2 the stream is synchronised after every operation and the first CUDA error is reported
// at the operation that caused it (asynchronous errors otherwise surface at a later, unrelated call). Read once per
// process; costs one predictable branch per operation when off.
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

namespace dlafb {

inline int op_trace_level() {
  static const int level = [] {
    const char* e = getenv("DLAFB_TRACE");
    return e ? atoi(e) : 0;
  }();
  return level;
}

// Called by the executors as `return traced(stream, code, "name", ...)` around the launch result `code`.
inline int op_trace_finish(cudaStream_t s, int code, const char* name) {
  if (op_trace_level() >= 2 && code == 0) {
    const cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
      std::fprintf(stderr, "[dlafb trace] %s FAILED on the device: %s\n", name, cudaGetErrorString(e));
      return (int)e;
    }
  }
  if (code != 0 && op_trace_level() >= 1)
    std::fprintf(stderr, "[dlafb trace] %s returned error %d\n", name, code);
  return code;
}

#define DLAFB_TRACE_OP(...)                    \
  do {                                         \
    if (dlafb::op_trace_level() >= 1) {        \
      std::fprintf(stderr, "[dlafb trace] ");  \
      std::fprintf(stderr, __VA_ARGS__);       \
      std::fprintf(stderr, "\n");              \
    }                                          \
  } while (0)

}  // namespace dlafb
