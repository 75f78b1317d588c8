// Device executor of the distributed algorithm templates (triangular_solver_dist.h, gen_to_std_dist.h): local
// operations on the caller's stream, collectives on the grid's row / column NCCL communicators on the same stream.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include "common.cuh"
#include "grid.h"
#include "solver_kernels.cuh"
#include "tile_ops.cuh"

namespace dlafb {

template <class T>
struct NcclType;
template <>
struct NcclType<float> {
  static constexpr ncclDataType_t value = ncclFloat;
  static constexpr size_t mult = 1;
};
template <>
struct NcclType<double> {
  static constexpr ncclDataType_t value = ncclDouble;
  static constexpr size_t mult = 1;
};
template <>
struct NcclType<float2> {
  static constexpr ncclDataType_t value = ncclFloat;
  static constexpr size_t mult = 2;
};
template <>
struct NcclType<double2> {
  static constexpr ncclDataType_t value = ncclDouble;
  static constexpr size_t mult = 2;
};

inline int nccl_code(ncclResult_t r) { return r == ncclSuccess ? 0 : 1000 + (int)r; }

// Local operations on the caller's stream; collectives on the grid's row / column communicators on the same stream
// (NCCL orders them with the kernels around them; every rank issues them in the same program order).
template <class T>
struct DistDeviceOps {
  cudaStream_t s;
  dlafb_grid* g;
  using R = typename Tr<T>::Real;
  int gemm_nt(int M, int N, int K, R alpha, const T* A, long long lda, const T* B, long long ldb, R beta, T* C,
              long long ldc, int lower) {
    return (int)dlafb::gemm_nt<T>(s, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower);
  }
  int transpose(int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj, char tri) {
    return (int)dlafb::transpose<T>(s, m, n, in, ldin, out, ldout, conj, tri);
  }
  int lacpy(char uplo, int m, int n, const T* a, long long lda, T* b, long long ldb) {
    return (int)dlafb::lacpy<T>(s, uplo, m, n, a, lda, b, ldb);
  }
  int tri_inverse_local(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w, int g0, int gstep, int row0,
                        int row_step, int col0, int col_step, int count) {
    TriInvParams p{n, nb, ldt, upper, unit, g0, gstep, row0, row_step, col0, col_step, count};
    return (int)launch_tri_inverse<T>(s, p, t, w);
  }
  int scale(int m, int n, T* b, long long ldb, T alpha, int conj) {
    return (int)launch_scale<T>(s, m, n, b, ldb, alpha, conj);
  }
  int zero(int m, int n, T* a, long long lda) {
    return (int)dlafb::laset<T>(s, 'G', m, n, Tr<T>::zero(), Tr<T>::zero(), a, lda);
  }
  int bcast_row(int root, T* p, size_t count) {
    return nccl_code(ncclBroadcast(p, p, count * NcclType<T>::mult, NcclType<T>::value, root, g->row, s));
  }
  int bcast_col(int root, T* p, size_t count) {
    return nccl_code(ncclBroadcast(p, p, count * NcclType<T>::mult, NcclType<T>::value, root, g->col, s));
  }
  int reduce_row(int root, T* p, size_t count) {
    return nccl_code(ncclReduce(p, p, count * NcclType<T>::mult, NcclType<T>::value, ncclSum, root, g->row, s));
  }
  // collectives between group_begin / group_end are launched together (ncclGroupStart / ncclGroupEnd)
  int group_begin() { return nccl_code(ncclGroupStart()); }
  int group_end() { return nccl_code(ncclGroupEnd()); }
  int reduce_col(int root, T* p, size_t count) {
    return nccl_code(ncclReduce(p, p, count * NcclType<T>::mult, NcclType<T>::value, ncclSum, root, g->col, s));
  }
};

}  // namespace dlafb
