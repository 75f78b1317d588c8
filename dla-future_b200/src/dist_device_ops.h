// Device executor of the distributed algorithm templates (triangular_solver_dist.h, gen_to_std_dist.h): local
// operations on the caller's stream, collectives on the grid's row / column NCCL communicators on the same stream.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include "common.cuh"
#include "grid.h"
#include "op_trace.h"
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
    DLAFB_TRACE_OP("gemm_nt M=%d N=%d K=%d alpha=%g beta=%g lower=%d A=%p lda=%lld B=%p ldb=%lld C=%p ldc=%lld", M, N, K,
                   (double)alpha, (double)beta, lower, (const void*)A, lda, (const void*)B, ldb, (void*)C, ldc);
    return op_trace_finish(s, (int)dlafb::gemm_nt<T>(s, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower), "gemm_nt");
  }
  int transpose(int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj, char tri) {
    DLAFB_TRACE_OP("transpose m=%d n=%d conj=%d tri=%c in=%p ldin=%lld out=%p ldout=%lld", m, n, conj, tri, (const void*)in,
                   ldin, (void*)out, ldout);
    return op_trace_finish(s, (int)dlafb::transpose<T>(s, m, n, in, ldin, out, ldout, conj, tri), "transpose");
  }
  int lacpy(char uplo, int m, int n, const T* a, long long lda, T* b, long long ldb) {
    DLAFB_TRACE_OP("lacpy %c m=%d n=%d a=%p lda=%lld b=%p ldb=%lld", uplo, m, n, (const void*)a, lda, (void*)b, ldb);
    return op_trace_finish(s, (int)dlafb::lacpy<T>(s, uplo, m, n, a, lda, b, ldb), "lacpy");
  }
  int tri_inverse_local(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w, int g0, int gstep, int row0,
                        int row_step, int col0, int col_step, int count) {
    TriInvParams p{n, nb, ldt, upper, unit, g0, gstep, row0, row_step, col0, col_step, count};
    DLAFB_TRACE_OP("tri_inverse upper=%d unit=%d n=%d nb=%d t=%p ldt=%lld w=%p g0=%d gstep=%d count=%d", upper, unit, n, nb,
                   (const void*)t, ldt, (void*)w, p.g0, p.gstep, p.count);
    return op_trace_finish(s, (int)launch_tri_inverse<T>(s, p, t, w), "tri_inverse");
  }
  int scale(int m, int n, T* b, long long ldb, T alpha, int conj) {
    DLAFB_TRACE_OP("scale m=%d n=%d conj=%d b=%p ldb=%lld", m, n, conj, (void*)b, ldb);
    return op_trace_finish(s, (int)launch_scale<T>(s, m, n, b, ldb, alpha, conj), "scale");
  }
  int zero(int m, int n, T* a, long long lda) {
    DLAFB_TRACE_OP("zero m=%d n=%d a=%p lda=%lld", m, n, (void*)a, lda);
    return op_trace_finish(s, (int)dlafb::laset<T>(s, 'G', m, n, Tr<T>::zero(), Tr<T>::zero(), a, lda), "zero");
  }
  int bcast_row(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("bcast_row root=%d count=%zu p=%p", root, count, (void*)p);
    return op_trace_finish(s, nccl_code(ncclBroadcast(p, p, count * NcclType<T>::mult, NcclType<T>::value, root, g->row, s)),
                           "bcast_row");
  }
  int bcast_col(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("bcast_col root=%d count=%zu p=%p", root, count, (void*)p);
    return op_trace_finish(s, nccl_code(ncclBroadcast(p, p, count * NcclType<T>::mult, NcclType<T>::value, root, g->col, s)),
                           "bcast_col");
  }
  int reduce_row(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("reduce_row root=%d count=%zu p=%p", root, count, (void*)p);
    return op_trace_finish(s, nccl_code(ncclReduce(p, p, count * NcclType<T>::mult, NcclType<T>::value, ncclSum, root, g->row,
                                                   s)), "reduce_row");
  }
  // collectives between group_begin / group_end are launched together (ncclGroupStart / ncclGroupEnd)
  int group_begin() { return nccl_code(ncclGroupStart()); }
  int group_end() { return nccl_code(ncclGroupEnd()); }
  int reduce_col(int root, T* p, size_t count) {
    DLAFB_TRACE_OP("reduce_col root=%d count=%zu p=%p", root, count, (void*)p);
    return op_trace_finish(s, nccl_code(ncclReduce(p, p, count * NcclType<T>::mult, NcclType<T>::value, ncclSum, root, g->col,
                                                   s)), "reduce_col");
  }
};

}  // namespace dlafb
