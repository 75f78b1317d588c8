// Triangular solver, local (one GPU): op(A) X = alpha B (side 'L') or X op(A) = alpha B (side 'R'), X overwriting B.
// Reference: dlaf::triangular_solver, include/dlaf/solver/triangular.h:25-101 and the eight tile loops call_LLN ...
// call_RUT of include/dlaf/solver/triangular/impl.h:235-470 (tile trsm on the k-th panel of B, then tile gemms on
// the panels still to be solved).
//
// B200 design: the machine has ONE kernel that runs at 94 % of the FP64 tensor peak, the C -= A B^H contraction of
// gemm_dmma.cu (both operands with the contracted index along columns). Instead of eight loops over four product
// forms (NN, TN, NT, ...) every case is reduced, by O(n^2) layout work up front, to one of two canonical solves on a
// triangular T stored the way that kernel reads it:
//
//   canonical right   X T^H = B   (the Cholesky panel solve, impl.h:55-67, is this with T = L_kk)
//       step k:  X_k = B_k inv(T_kk)^H,      then   B_j -= X_k T_jk^H   for the column blocks j still to be solved
//   canonical left    T X = B
//       step k:  X_k = inv(T_kk) B_k,        then   B_i -= T_ik X_k     for the row blocks i still to be solved
//       (X_k is the K-major operand here: its conjugate transpose, one n x nb panel per step, is what the kernel reads)
//
//   side op   canonical   T
//   L    N    left        A
//   L    T    left        A^T   (transposed copy, uplo flips)
//   L    C    left        A^H   (copy)
//   R    C    right       A
//   R    T    right       A for real types; for complex types conj(X) A^H = conj(alpha B): B is conjugated in place
//                         before and after
//   R    N    right       A^H   (copy)
//
// The diagonal tiles are inverted once, all at the same time (tri_inverse, one CTA per tile), so that every step
// of the solve is a contraction; alpha is applied to B once at the start (the reference applies it in every tile
// trsm and compensates with beta = -1/alpha in the gemms: same result).
//
// Look-ahead (TriSolveWork::lookahead): two lanes (streams). Lane 0 solves block s and applies its update to the
// NEXT block only, lane 1 applies it to all the other remaining blocks ("bulk"), so that the solve of block s+1 -
// a chain of small dependent launches - runs under the bulk update of step s instead of after it:
//   lane 0:  solve(s) -> [bulk(s-1) done] update next(s) -> solve(s+1) -> ...
//   lane 1:  bulk(s-1) -> [solve(s) done] bulk(s) -> ...
// bulk(s-1) also writes the block that next(s) writes, hence the wait; the panel with X_s^H is double-buffered.
//
// The algorithm is a template over the executor `Ops` so that tests/emu/ can run the very same sequence of
// operations, pointer arithmetic included, on the CPU (the device executor is in triangular_solver.cu); the CPU
// executor also replays the two lanes under adversarial interleavings and checks every pair of operations that the
// events leave unordered for overlapping reads / writes.
#pragma once
#include <cstddef>

#include "common.cuh"
#include "solver_kernels.cuh"

namespace dlafb {

struct TriSolveDesc {
  char side, uplo, op, diag;  // 'L'/'R', 'L'/'U', 'N'/'T'/'C', 'N'/'U'
  int m, n;                   // B is m x n; A is m x m (side 'L') or n x n (side 'R')
  int nb;                     // tile size of A
};

template <class T>
struct TriSolveWork {
  T* tcopy;        // order x order (leading dimension ldt) when tri_solve_needs_copy(), else unused
  long long ldt;
  T* winv;         // ceil(order / nb) slots of nb x nb
  T* panel;        // ldp x nb (two of them, panel_stride apart, with look-ahead)
  long long ldp;   // >= max(m, n)
  int lookahead = 0;
  size_t panel_stride = 0;
};

// events of the look-ahead variant (Ops::signal / Ops::wait)
enum TriSolveEvent { kEvPrologue = 0, kEvSolve0 = 1, kEvSolve1 = 2, kEvBulk0 = 3, kEvBulk1 = 4, kTriSolveEvents = 5 };

inline int tri_solve_order(const TriSolveDesc& d) { return d.side == 'L' ? d.m : d.n; }

inline bool tri_solve_needs_copy(const TriSolveDesc& d) {
  return d.side == 'L' ? d.op != 'N' : d.op == 'N';
}

inline long long tri_solve_even(long long x) { return (x + 1) & ~1LL; }

// Elements of each workspace part (the caller allocates; leading dimensions are even so that double operands keep
// the 16-byte alignment the TMA path needs).
inline long long tri_solve_ldt(const TriSolveDesc& d) { return tri_solve_even(tri_solve_order(d) > 0 ? tri_solve_order(d) : 1); }
inline long long tri_solve_ldp(const TriSolveDesc& d) { return tri_solve_even(d.m > d.n ? (d.m > 0 ? d.m : 1) : (d.n > 0 ? d.n : 1)); }
inline size_t tri_solve_tcopy_elems(const TriSolveDesc& d) {
  return tri_solve_needs_copy(d) ? (size_t)tri_solve_ldt(d) * (size_t)(tri_solve_order(d) > 0 ? tri_solve_order(d) : 1) : 0;
}
inline size_t tri_solve_winv_elems(const TriSolveDesc& d) {
  const int o = tri_solve_order(d);
  return (size_t)(o > 0 ? ceil_div(o, d.nb) : 1) * d.nb * d.nb;
}
inline size_t tri_solve_panel_elems(const TriSolveDesc& d) { return (size_t)tri_solve_ldp(d) * d.nb; }

// Ops interface (all return 0 on success):
//   gemm_nt(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc)   C <- beta C + alpha A B^H (alpha, beta real)
//   transpose(m, n, in, ldin, out, ldout, conj, tri)        out (n x m) <- op(in), only the tri part of out written
//   lacpy(m, n, a, lda, b, ldb)                             b <- a
//   tri_inverse(upper, unit, n, nb, t, ldt, w)              w slots <- inverses of the diagonal tiles of t
//   scale(m, n, b, ldb, alpha, conj)                        b <- alpha * (conj ? conj(b) : b)
// and, used only with look-ahead:
//   lane(l)      following operations go to lane l (0 or 1); each lane runs its operations in order
//   signal(ev)   the current lane records event ev
//   wait(ev)     the current lane waits for the last signal of ev enqueued so far (no-op if there is none)
template <class T, class Ops>
int triangular_solve(Ops& ops, const TriSolveDesc& d, T alpha, const T* A, long long lda, T* B, long long ldb,
                     const TriSolveWork<T>& w) {
  using R = typename Tr<T>::Real;
  const bool left = d.side == 'L';
  const int order = tri_solve_order(d);
  if (d.m <= 0 || d.n <= 0)
    return 0;
  const int nb = d.nb;
  int e;

  // canonical triangular operand
  const T* t = A;
  long long ldt = lda;
  bool lower = d.uplo == 'L';
  const bool conj_b = !left && d.op == 'T' && Tr<T>::is_complex;
  if (tri_solve_needs_copy(d)) {
    const int conj = (left && d.op == 'T') ? 0 : 1;
    lower = !lower;
    if ((e = ops.transpose(order, order, A, lda, w.tcopy, w.ldt, conj, lower ? 'L' : 'U')) != 0)
      return e;
    t = w.tcopy;
    ldt = w.ldt;
  }

  if (conj_b || !tr_is_one<T>(alpha)) {
    if ((e = ops.scale(d.m, d.n, B, ldb, conj_b ? Tr<T>::conj(alpha) : alpha, conj_b ? 1 : 0)) != 0)
      return e;
  }

  if ((e = ops.tri_inverse(lower ? 0 : 1, d.diag == 'U' ? 1 : 0, order, nb, t, ldt, w.winv)) != 0)
    return e;

  const int nt = ceil_div(order, nb);
  if (w.lookahead) {
    if ((e = ops.signal(kEvPrologue)) != 0)
      return e;
    ops.lane(1);
    if ((e = ops.wait(kEvPrologue)) != 0)
      return e;
    ops.lane(0);
    bool bulk_issued[2] = {false, false};
    for (int s = 0; s < nt; ++s) {
      const int k = lower ? s : nt - 1 - s;
      const int k0 = k * nb;
      const int wk = order - k0 < nb ? order - k0 : nb;
      const T* wkk = w.winv + (size_t)k * nb * nb;
      T* pan = w.panel + (size_t)(s & 1) * w.panel_stride;
      const bool more = s + 1 < nt;
      // next block in solve order and what is left beyond it
      const int kn = lower ? k + 1 : k - 1;
      const int n0 = more ? kn * nb : 0;
      const int wn = more ? (order - n0 < nb ? order - n0 : nb) : 0;
      const int b0 = lower ? n0 + wn : 0;                            // first row / column of the bulk range
      const int bulk = more ? (lower ? order - n0 - wn : n0) : 0;    // its extent
      // ---- lane 0: solve block k
      if (!left) {
        T* bk = B + (long long)k0 * ldb;
        if ((e = ops.lacpy(d.m, wk, bk, ldb, pan, w.ldp)) != 0)
          return e;
        if ((e = ops.gemm_nt(d.m, wk, wk, R(1), pan, w.ldp, wkk, (long long)nb, R(0), bk, ldb)) != 0)
          return e;
      }
      else {
        T* bk = B + k0;
        if ((e = ops.transpose(wk, d.n, bk, ldb, pan, w.ldp, 1, 'G')) != 0)
          return e;
        if ((e = ops.gemm_nt(wk, d.n, wk, R(1), wkk, (long long)nb, pan, w.ldp, R(0), bk, ldb)) != 0)
          return e;
        if (more && (e = ops.transpose(wk, d.n, bk, ldb, pan, w.ldp, 1, 'G')) != 0)
          return e;
      }
      if (!more)
        break;
      if ((e = ops.signal(kEvSolve0 + (s & 1))) != 0)
        return e;
      // ---- lane 0: update of the next block; bulk(s-1) wrote that block too
      if (s >= 1 && bulk_issued[(s - 1) & 1] && (e = ops.wait(kEvBulk0 + ((s - 1) & 1))) != 0)
        return e;
      if (!left)
        e = ops.gemm_nt(d.m, wn, wk, R(-1), B + (long long)k0 * ldb, ldb, t + n0 + (long long)k0 * ldt, ldt, R(1),
                        B + (long long)n0 * ldb, ldb);
      else
        e = ops.gemm_nt(wn, d.n, wk, R(-1), t + n0 + (long long)k0 * ldt, ldt, pan, w.ldp, R(1), B + n0, ldb);
      if (e != 0)
        return e;
      // ---- lane 1: all the other remaining blocks
      bulk_issued[s & 1] = bulk > 0;
      if (bulk > 0) {
        ops.lane(1);
        if ((e = ops.wait(kEvSolve0 + (s & 1))) != 0)
          return e;
        if (!left)
          e = ops.gemm_nt(d.m, bulk, wk, R(-1), B + (long long)k0 * ldb, ldb, t + b0 + (long long)k0 * ldt, ldt, R(1),
                          B + (long long)b0 * ldb, ldb);
        else
          e = ops.gemm_nt(bulk, d.n, wk, R(-1), t + b0 + (long long)k0 * ldt, ldt, pan, w.ldp, R(1), B + b0, ldb);
        if (e != 0)
          return e;
        if ((e = ops.signal(kEvBulk0 + (s & 1))) != 0)
          return e;
        ops.lane(0);
      }
    }
    // join: everything lane 1 did is ordered before what follows on lane 0
    for (int q = 0; q < 2; ++q)
      if (bulk_issued[q] && (e = ops.wait(kEvBulk0 + q)) != 0)
        return e;
    if (conj_b)
      return ops.scale(d.m, d.n, B, ldb, Tr<T>::from_real(R(1)), 1);
    return 0;
  }
  for (int s = 0; s < nt; ++s) {
    const int k = lower ? s : nt - 1 - s;  // forward for a lower T, backward for an upper one
    const int k0 = k * nb;
    const int wk = order - k0 < nb ? order - k0 : nb;
    const T* wkk = w.winv + (size_t)k * nb * nb;
    // blocks still to be solved: after k (lower) or before k (upper)
    const int r0 = lower ? k0 + wk : 0;
    const int rest = lower ? order - k0 - wk : k0;
    const T* toff = t + r0 + (long long)k0 * ldt;  // T(rest blocks, k): rest x wk
    if (!left) {
      T* bk = B + (long long)k0 * ldb;  // m x wk
      if ((e = ops.lacpy(d.m, wk, bk, ldb, w.panel, w.ldp)) != 0)
        return e;
      if ((e = ops.gemm_nt(d.m, wk, wk, R(1), w.panel, w.ldp, wkk, (long long)nb, R(0), bk, ldb)) != 0)
        return e;
      if (rest > 0 &&
          (e = ops.gemm_nt(d.m, rest, wk, R(-1), bk, ldb, toff, ldt, R(1), B + (long long)r0 * ldb, ldb)) != 0)
        return e;
    }
    else {
      T* bk = B + k0;  // wk x n
      if ((e = ops.transpose(wk, d.n, bk, ldb, w.panel, w.ldp, 1, 'G')) != 0)  // panel (n x wk) = B_k^H
        return e;
      if ((e = ops.gemm_nt(wk, d.n, wk, R(1), wkk, (long long)nb, w.panel, w.ldp, R(0), bk, ldb)) != 0)
        return e;
      if (rest > 0) {
        if ((e = ops.transpose(wk, d.n, bk, ldb, w.panel, w.ldp, 1, 'G')) != 0)  // panel = X_k^H
          return e;
        if ((e = ops.gemm_nt(rest, d.n, wk, R(-1), toff, ldt, w.panel, w.ldp, R(1), B + r0, ldb)) != 0)
          return e;
      }
    }
  }

  if (conj_b)
    return ops.scale(d.m, d.n, B, ldb, Tr<T>::from_real(R(1)), 1);
  return 0;
}

}  // namespace dlafb
