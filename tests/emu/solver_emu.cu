// TEST INFRASTRUCTURE ONLY. CPU execution of the triangular solver for boxes without a GPU: the product's own
// algorithm template (dla-future_b200/src/triangular_solver.h) and its own kernel bodies
// (dla-future_b200/csrc/solver_kernels.cuh, run once per thread of the launch geometry) over an executor whose
// remaining operations (contraction, transpose, copy) are restated as plain loops with the device kernels'
// documented semantics. Built host-only by nvcc (no __global__ code here); never linked into the product.
//
// What this pins without a GPU: the reduction of the 24 side/uplo/op/diag cases to the two canonical solves, every
// pointer offset and leading dimension handed to the kernels, the tile-inverse recurrence (lower/upper/unit), the
// workspace sizes (workspaces are NaN-filled, so reading anything the algorithm did not write poisons the result),
// and that no contraction is asked to run with C aliasing an operand.
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "triangular_solver.h"

namespace {
using namespace dlafb;

template <class T>
T nan_value() {
  using R = typename Tr<T>::Real;
  const R q = std::numeric_limits<R>::quiet_NaN();
  if constexpr (Tr<T>::is_complex)
    return Tr<T>::make(q, q);
  else
    return q;
}

// do the element sets {p + i + j*ld : i < rows, j < cols} of two operands intersect?
template <class T>
bool overlap(const T* a, long long lda, int ra, int ca, const T* b, long long ldb, int rb, int cb) {
  for (int j = 0; j < ca; ++j) {
    const T* a0 = a + j * lda;
    const T* a1 = a0 + ra;
    for (int l = 0; l < cb; ++l) {
      const T* b0 = b + l * ldb;
      const T* b1 = b0 + rb;
      if (a0 < b1 && b0 < a1)
        return true;
    }
  }
  return false;
}

template <class T>
struct HostOps {
  using R = typename Tr<T>::Real;
  int hazards = 0;
  int gemm_nt(int M, int N, int K, R alpha, const T* A, long long lda, const T* B, long long ldb, R beta, T* C,
              long long ldc) {
    if (M <= 0 || N <= 0)
      return 0;
    if (overlap<T>(C, ldc, M, N, A, lda, M, K) || overlap<T>(C, ldc, M, N, B, ldb, N, K)) {
      ++hazards;
      return 9001;
    }
    for (int j = 0; j < N; ++j)
      for (int i = 0; i < M; ++i) {
        T acc = Tr<T>::zero();
        for (int k = 0; k < K; ++k)
          acc = Tr<T>::fma(A[i + k * lda], Tr<T>::conj(B[j + k * ldb]), acc);
        T v = Tr<T>::scale(acc, alpha);
        if (beta != R(0))
          v = Tr<T>::add(v, Tr<T>::scale(C[i + j * ldc], beta));
        C[i + j * ldc] = v;
      }
    return 0;
  }
  // aux.cu transpose_kernel: out(j, i) = op(in(i, j)); tri 'L' writes only out rows >= out cols, 'U' rows <= cols
  int transpose(int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj, char tri) {
    if (overlap<T>(out, ldout, n, m, in, ldin, m, n)) {
      ++hazards;
      return 9002;
    }
    for (int i = 0; i < m; ++i)
      for (int j = 0; j < n; ++j) {
        if ((tri == 'L' && j < i) || (tri == 'U' && j > i))
          continue;
        const T v = in[i + j * ldin];
        out[j + i * ldout] = conj ? Tr<T>::conj(v) : v;
      }
    return 0;
  }
  int lacpy(int m, int n, const T* a, long long lda, T* b, long long ldb) {
    for (int j = 0; j < n; ++j)
      for (int i = 0; i < m; ++i)
        b[i + j * ldb] = a[i + j * lda];
    return 0;
  }
  int tri_inverse(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w) {
    TriInvParams p{n, nb, ldt, upper, unit};
    const int blocks = tri_inverse_blocks(n, nb);
    for (int b = 0; b < blocks; ++b)
      for (int th = 0; th < kTriInvThreads; ++th)
        tri_inverse_body<T>(p, t, w, b, th, kTriInvThreads);
    return 0;
  }
  int scale(int m, int n, T* b, long long ldb, T alpha, int conj) {
    for (int by = 0; by < scale_blocks_y(n); ++by)
      for (int bx = 0; bx < scale_blocks_x(m); ++bx)
        for (int th = 0; th < kScaleRows; ++th)
          scale_body<T>(m, n, b, ldb, alpha, conj, bx, by, th);
    return 0;
  }
};

template <class T>
int solve(char side, char uplo, char op, char diag, int m, int n, int nb, const void* alpha, const void* a,
          long long lda, void* b, long long ldb) {
  TriSolveDesc d{side, uplo, op, diag, m, n, nb};
  std::vector<T> tcopy(tri_solve_tcopy_elems(d), nan_value<T>());
  std::vector<T> winv(tri_solve_winv_elems(d), nan_value<T>());
  std::vector<T> panel(tri_solve_panel_elems(d), nan_value<T>());
  TriSolveWork<T> w{tcopy.data(), tri_solve_ldt(d), winv.data(), panel.data(), tri_solve_ldp(d)};
  HostOps<T> ops;
  return triangular_solve<T>(ops, d, *static_cast<const T*>(alpha), static_cast<const T*>(a), lda, static_cast<T*>(b),
                             ldb, w);
}

template <class T>
int inverse(int upper, int unit, int n, int nb, const void* t, long long ldt, void* w) {
  HostOps<T> ops;
  return ops.tri_inverse(upper, unit, n, nb, static_cast<const T*>(t), ldt, static_cast<T*>(w));
}

}  // namespace

#define EMU_TYPED(X, T)                                                                                             \
  int emu_triangular_solver_##X(char side, char uplo, char op, char diag, int m, int n, int nb, const void* alpha, \
                                const void* a, long long lda, void* b, long long ldb) {                             \
    return solve<T>(side, uplo, op, diag, m, n, nb, alpha, a, lda, b, ldb);                                         \
  }                                                                                                                 \
  int emu_tri_inverse_##X(int upper, int unit, int n, int nb, const void* t, long long ldt, void* w) {              \
    return inverse<T>(upper, unit, n, nb, t, ldt, w);                                                               \
  }

extern "C" {
EMU_TYPED(s, float)
EMU_TYPED(d, double)
EMU_TYPED(c, float2)
EMU_TYPED(z, double2)
}
