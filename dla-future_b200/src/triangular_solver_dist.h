// Distributed triangular solver: op(A) X = alpha B (side 'L') or X op(A) = alpha B (side 'R'), X overwriting B, with
// A and B distributed 2D block-cyclically over a P x Q grid of GPUs (same grid, same source rank).
// Reference: dlaf::triangular_solver(grid, ...), include/dlaf/solver/triangular.h:103-178, and the eight distributed
// loops call_LLN ... call_RUT of include/dlaf/solver/triangular/impl.h:475-1195. The reference has two patterns and
// so does this file, because the pattern is dictated by where the tiles of A live, not by the runtime:
//
//   "broadcast" (side L with op N, side R with op N): right-looking. Per step k the stored part of column k (side L)
//   or row k (side R) of A travels along the grid rows (columns), block k of B is solved on the ranks that own it,
//   travels along the other axis, and every rank updates its part of the blocks still to be solved.
//   impl.h:475-566 (LLN), :658-746 (LUN), :838-926 (RLN), :1018-1107 (RUN).
//
//   "reduce" (op T / C): left-looking, because the tiles that a right-looking update would need sit on the wrong
//   ranks (A cannot be transposed without an all-to-all). Per step k every rank multiplies its part of the blocks
//   ALREADY solved with its part of column / row k of A, the partial sums are reduced onto the ranks that own block k,
//   and those solve it. impl.h:570-655 (LLT), :749-835 (LUT), :929-1015 (RLT), :1110-1195 (RUT).
//
// B200 design. Every product is the one contraction C <- beta C + alpha X Y^H that runs at 94 % of the FP64 tensor
// peak (gemm_dmma.cu; both operands with the contracted index along their columns), on operands as large as the local
// block allows - one launch per step and rank instead of one task per tile:
//   side L, broadcast:  B_rem -= A_rem,k X_k        X = column panel of A as received, Y = X_k^H (the solved block is
//                                                   transposed once, on its owners, before it travels)
//   side L, reduce:     R_k^H = B_k^H - sum X_rem^H A_rem,k   kept in the transposed domain: X = BT, the running
//                                                   conjugate transpose of the solved part of the local B, Y = the
//                                                   transposed panel; the reduced block is solved as X_k^H = R_k^H W_k
//   side R, broadcast:  B_rem -= X_k A_k,rem        X = X_k as received, Y = the transposed row panel of A
//   side R, reduce:     R_k = B_k - sum X_rem A_k,rem^H       X = the local B itself, Y = the row panel as received:
//                                                   no layout work at all (this is the Cholesky panel-solve form)
// W_k = inv(A_kk): all the diagonal tiles a rank owns are inverted up front in ONE launch (tri_inverse with a strided
// tile placement) and W_k travels inside the panel in the place of A_kk, so a block solve is a contraction too and no
// step has a dependent single-CTA kernel in it.
// alpha is applied to the local B once; op = T on complex data is op = C on conj(B) (as in triangular_solver.h).
// Collectives are issued by every rank of a communicator in the same program order with equal counts (the counts
// depend only on the coordinate shared by the communicator's ranks); a rank with an empty local block still takes
// part. Exchange = NCCL broadcast / reduce on the row and column communicators (grid.cu) on the caller's stream.
//
// Template over the executor `Ops` like the local solver: tests/emu/ runs this code for all the ranks of a simulated
// grid (one host thread per rank, collectives through shared memory) against the reference's analytical systems.
#pragma once
#include <cstddef>

#include "common.cuh"
#include "schedule.h"
#include "solver_kernels.cuh"

namespace dlafb {

struct TriDistDesc {
  char side, uplo, op, diag;
  int m, n;    // B is m x n (global); A is m x m (side 'L') or n x n (side 'R')
  int mb, nb;  // tile size of B; A has square tiles of mb (side 'L') or nb (side 'R')
  int P, Q, pr, pc, src_r, src_c;
  // 1 (side 'L', uplo 'L', op 'N', mb == nb only): B is the strictly-lower block trapezoid of a square matrix - at
  // step k only the columns of the column tiles before k take part (gen_to_std_dist.h solves all its column panels
  // side by side this way); the other tiles of B are neither read nor written.
  int trapezoid = 0;
};

struct TriDistGeom {
  Dist1D rows, cols;    // of B
  Dist1D arows, acols;  // of A
  int t;                // tile size of A
  long long lrows, lcols, larows, lacols;
  // the diagonal tiles of A that this rank owns: global g0, g0 + gstep, ... (nd of them)
  int g0, gstep, nd;
};

inline int tri_dist_gcd(int a, int b) { return b == 0 ? a : tri_dist_gcd(b, a % b); }

inline TriDistGeom tri_dist_geom(const TriDistDesc& d) {
  TriDistGeom g;
  g.rows = Dist1D{d.m, d.mb, d.P, d.pr, d.src_r};
  g.cols = Dist1D{d.n, d.nb, d.Q, d.pc, d.src_c};
  const bool left = d.side == 'L';
  g.t = left ? d.mb : d.nb;
  const long long order = left ? d.m : d.n;
  g.arows = Dist1D{order, g.t, d.P, d.pr, d.src_r};
  g.acols = Dist1D{order, g.t, d.Q, d.pc, d.src_c};
  g.lrows = g.rows.local_size();
  g.lcols = g.cols.local_size();
  g.larows = g.arows.local_size();
  g.lacols = g.acols.local_size();
  g.gstep = d.P / tri_dist_gcd(d.P, d.Q) * d.Q;
  g.g0 = -1;
  g.nd = 0;
  const long long nt = g.arows.nt();
  for (int k = 0; k < g.gstep && k < nt; ++k)
    if (g.arows.owner(k) == d.pr && g.acols.owner(k) == d.pc) {
      g.g0 = k;
      g.nd = (int)((nt - 1 - k) / g.gstep) + 1;
      break;
    }
  if (g.g0 < 0)
    g.g0 = 0;
  return g;
}

inline long long tri_dist_even(long long x) { return (x + 1) & ~1LL; }
inline long long tri_dist_ld(long long x) { return tri_dist_even(x > 0 ? x : 1); }

// Workspace parts in elements (per rank).
struct TriDistSizes {
  size_t winv, pan, pan2, xb, wkh, bt;
};

inline TriDistSizes tri_dist_sizes(const TriDistDesc& d) {
  const TriDistGeom g = tri_dist_geom(d);
  const size_t t = (size_t)g.t;
  TriDistSizes s{};
  s.winv = (size_t)(g.nd > 0 ? g.nd : 1) * t * t;
  s.wkh = (size_t)tri_dist_ld(g.t) * t;
  if (d.side == 'L') {
    s.pan = (size_t)tri_dist_ld(g.lrows) * t;   // column panel of A on the local rows
    s.pan2 = (size_t)tri_dist_ld(g.t) * (size_t)(g.lrows > 0 ? g.lrows : 1);  // its conjugate transpose (reduce pattern)
    s.xb = (size_t)tri_dist_ld(g.lcols) * t;    // transposed block of B
    s.bt = d.op == 'N' ? 0 : (size_t)tri_dist_ld(g.lcols) * (size_t)(g.lrows > 0 ? g.lrows : 1);
  }
  else {
    s.pan = (size_t)tri_dist_ld(g.t) * (size_t)(g.lcols > 0 ? g.lcols : 1);   // row panel of A on the local columns
    s.pan2 = (size_t)tri_dist_ld(g.lcols) * t;  // its conjugate transpose (broadcast pattern)
    s.xb = (size_t)tri_dist_ld(g.lrows) * t;    // block of B
    s.bt = 0;
  }
  return s;
}

template <class T>
struct TriDistWork {
  T *winv, *pan, *pan2, *xb, *wkh, *bt;
};

// C <- beta C + alpha X Y^H over K columns, issued as a part whose K is a multiple of the tile size t plus the ragged
// rest, so that one ragged tile at the end of the range does not take the whole product off the TMA + DMMA path.
template <class T, class Ops>
int tri_dist_gemm_ksplit(Ops& ops, int M, int N, int K, int t, typename Tr<T>::Real alpha, const T* X, long long ldx,
                         const T* Y, long long ldy, typename Tr<T>::Real beta, T* C, long long ldc) {
  using R = typename Tr<T>::Real;
  const int k1 = (K / t) * t;
  if (k1 == 0 || k1 == K)
    return ops.gemm_nt(M, N, K, alpha, X, ldx, Y, ldy, beta, C, ldc, 0);
  int e = ops.gemm_nt(M, N, k1, alpha, X, ldx, Y, ldy, beta, C, ldc, 0);
  if (e != 0)
    return e;
  return ops.gemm_nt(M, N, K - k1, alpha, X + (long long)k1 * ldx, ldx, Y + (long long)k1 * ldy, ldy, R(1), C, ldc, 0);
}

// Ops interface: the operations of gen_to_std.h (gemm_nt with `lower`, transpose, lacpy with uplo, tri_inverse) plus
//   tri_inverse_local(upper, unit, n, nb, t, ldt, w, g0, gstep, row0, row_step, col0, col_step, count)
//   scale(m, n, b, ldb, alpha, conj)
//   zero(m, n, a, lda)
//   bcast_row(root_col, p, count) / bcast_col(root_row, p, count)      in place, count elements of T
//   reduce_row(root_col, p, count) / reduce_col(root_row, p, count)    sum, in place; only the root's buffer is defined after
//   group_begin() / group_end()                                        collectives in between may be launched together
template <class T, class Ops>
int triangular_solve_dist(Ops& ops, const TriDistDesc& d, T alpha, const T* A, long long lda, T* B, long long ldb,
                          const TriDistWork<T>& w) {
  using R = typename Tr<T>::Real;
  if (d.m <= 0 || d.n <= 0)
    return 0;
  const TriDistGeom g = tri_dist_geom(d);
  const bool left = d.side == 'L';
  const bool lower = d.uplo == 'L';
  const bool bcast_pattern = d.op == 'N';
  const int t = g.t;
  const int lrows = (int)g.lrows, lcols = (int)g.lcols;
  const int nt = (int)g.arows.nt();
  // order of the steps: the block that depends on no other one comes first
  const bool forward = left ? (lower == bcast_pattern) : (lower != bcast_pattern);
  const bool conj_b = d.op == 'T' && Tr<T>::is_complex;
  int e;

  if (conj_b || !tr_is_one<T>(alpha)) {
    if ((e = ops.scale(lrows, lcols, B, ldb, conj_b ? Tr<T>::conj(alpha) : alpha, conj_b ? 1 : 0)) != 0)
      return e;
  }
  if (g.nd > 0) {
    const int row0 = (int)g.arows.local_of(g.g0), col0 = (int)g.acols.local_of(g.g0);
    if ((e = ops.tri_inverse_local(lower ? 0 : 1, d.diag == 'U' ? 1 : 0, left ? d.m : d.n, t, A, lda, w.winv, g.g0,
                                   g.gstep, row0, g.gstep / d.P, col0, g.gstep / d.Q, g.nd)) != 0)
      return e;
  }

  const long long ld_wkh = tri_dist_ld(t);
  for (int s = 0; s < nt; ++s) {
    const int k = forward ? s : nt - 1 - s;
    const int wk = (int)g.arows.tile_size(k);
    const int pr_k = g.arows.owner(k), qc_k = g.acols.owner(k);
    const bool own_row = d.pr == pr_k, own_col = d.pc == qc_k;
    const bool more = s + 1 < nt;
    const T* wkk = (own_row && own_col) ? w.winv + (size_t)((k - g.g0) / g.gstep) * t * t : nullptr;

    if (left) {
      // ---- column k of A on the local rows: the stored part = [diagonal tile + the tiles beyond it]
      const Dist1D& DR = g.arows;
      const int tb = lower ? (int)DR.next_local(k) : 0;
      const int te = lower ? (int)DR.nloc_tiles() : (int)DR.next_local(k + 1);
      const int r_begin = tb * t;
      const int r_end = te == (int)DR.nloc_tiles() ? lrows : te * t;
      const int prow = r_end > r_begin ? r_end - r_begin : 0;
      const long long ldap = tri_dist_ld(prow);
      // the tiles strictly beyond the diagonal: to be updated (broadcast) / already solved (reduce)
      const int mb_ = lower ? (int)DR.next_local(k + 1) : 0;
      const int me_ = lower ? (int)DR.nloc_tiles() : (int)DR.next_local(k);
      const int rem_begin = mb_ * t;
      const int rem_end = me_ == (int)DR.nloc_tiles() ? lrows : me_ * t;
      const int rem_rows = rem_end > rem_begin ? rem_end - rem_begin : 0;
      const int kl_r = own_row ? (int)DR.local_of(k) : -1;
      const long long diag_off = own_row ? (long long)kl_r * t - r_begin : 0;
      T* bk = own_row ? B + (long long)kl_r * t : nullptr;
      if (prow > 0) {
        if (own_col) {
          const long long kl_c = g.acols.local_of(k);
          if ((e = ops.lacpy('G', prow, wk, A + r_begin + kl_c * t * lda, lda, w.pan, ldap)) != 0)
            return e;
          if (own_row && (e = ops.lacpy('G', wk, wk, wkk, (long long)t, w.pan + diag_off, ldap)) != 0)
            return e;
        }
        if (d.Q > 1 && (e = ops.bcast_row(qc_k, w.pan, (size_t)ldap * wk)) != 0)
          return e;
      }
      if (bcast_pattern) {
        // columns of B that take part in this step (all of them, or the tiles before k of a block trapezoid)
        int nc = lcols;
        if (d.trapezoid) {
          const long long lim = g.cols.next_local(k) * (long long)t;
          nc = lim < lcols ? (int)lim : lcols;
        }
        if (nc <= 0)
          continue;
        const long long ldx = tri_dist_ld(nc);
        if (own_row) {
          if ((e = ops.transpose(wk, nc, bk, ldb, w.xb, ldx, 1, 'G')) != 0)  // xb = B_k^H
            return e;
          if ((e = ops.gemm_nt(wk, nc, wk, R(1), w.pan + diag_off, ldap, w.xb, ldx, R(0), bk, ldb, 0)) != 0)
            return e;
          if (more && (e = ops.transpose(wk, nc, bk, ldb, w.xb, ldx, 1, 'G')) != 0)  // xb = X_k^H
            return e;
        }
        if (!more)
          continue;
        if (d.P > 1 && (e = ops.bcast_col(pr_k, w.xb, (size_t)ldx * wk)) != 0)
          return e;
        if (rem_rows > 0 && (e = ops.gemm_nt(rem_rows, nc, wk, R(-1), w.pan + (rem_begin - r_begin), ldap, w.xb, ldx,
                                             R(1), B + rem_begin, ldb, 0)) != 0)
          return e;
        continue;
      }
      if (lcols <= 0)
        continue;
      const long long ldx = tri_dist_ld(lcols);
      {
        // reduce pattern:
        // xb (lcols x wk) = B_k^H - sum over the solved local rows of X^H A(:, k), reduced onto the owners of block k
        const long long ldpt = tri_dist_ld(t);
        if (rem_rows > 0 &&
            (e = ops.transpose(rem_rows, wk, w.pan + (rem_begin - r_begin), ldap, w.pan2, ldpt, 1, 'G')) != 0)
          return e;
        if (own_row) {
          if ((e = ops.transpose(wk, lcols, bk, ldb, w.xb, ldx, 1, 'G')) != 0)
            return e;
        }
        else if (rem_rows == 0 && d.P > 1) {
          if ((e = ops.zero(lcols, wk, w.xb, ldx)) != 0)
            return e;
        }
        if (rem_rows > 0 && (own_row || d.P > 1) &&
            (e = tri_dist_gemm_ksplit<T>(ops, lcols, wk, rem_rows, t, R(-1), w.bt + (long long)rem_begin * ldx, ldx,
                                         w.pan2, ldpt, own_row ? R(1) : R(0), w.xb, ldx)) != 0)
          return e;
        if (d.P > 1 && (e = ops.reduce_col(pr_k, w.xb, (size_t)ldx * wk)) != 0)
          return e;
        if (own_row) {
          T* btk = w.bt + (long long)kl_r * t * ldx;
          if ((e = ops.transpose(wk, wk, w.pan + diag_off, ldap, w.wkh, ld_wkh, 1, 'G')) != 0)  // W_k^H
            return e;
          if ((e = ops.gemm_nt(lcols, wk, wk, R(1), w.xb, ldx, w.wkh, ld_wkh, R(0), btk, ldx, 0)) != 0)  // X_k^H = R_k^H W_k
            return e;
          if ((e = ops.transpose(lcols, wk, btk, ldx, bk, ldb, 1, 'G')) != 0)
            return e;
        }
      }
    }
    else {
      // ---- row k of A on the local columns: the stored part = [the tiles before the diagonal + diagonal] (lower)
      const Dist1D& DC = g.acols;
      const int tb = lower ? 0 : (int)DC.next_local(k);
      const int te = lower ? (int)DC.next_local(k + 1) : (int)DC.nloc_tiles();
      const int c_begin = tb * t;
      const int c_end = te == (int)DC.nloc_tiles() ? lcols : te * t;
      const int pcol = c_end > c_begin ? c_end - c_begin : 0;
      const long long ldar = tri_dist_ld(t);
      const int mb_ = lower ? 0 : (int)DC.next_local(k + 1);
      const int me_ = lower ? (int)DC.next_local(k) : (int)DC.nloc_tiles();
      const int rem_begin = mb_ * t;
      const int rem_end = me_ == (int)DC.nloc_tiles() ? lcols : me_ * t;
      const int rem_cols = rem_end > rem_begin ? rem_end - rem_begin : 0;
      const int kl_c = own_col ? (int)DC.local_of(k) : -1;
      const long long diag_off = own_col ? ((long long)kl_c * t - c_begin) * ldar : 0;
      T* bk = own_col ? B + (long long)kl_c * t * ldb : nullptr;
      if (pcol > 0) {
        if (own_row) {
          const long long kl_r = g.arows.local_of(k);
          if ((e = ops.lacpy('G', wk, pcol, A + kl_r * t + (long long)c_begin * lda, lda, w.pan, ldar)) != 0)
            return e;
          if (own_col && (e = ops.lacpy('G', wk, wk, wkk, (long long)t, w.pan + diag_off, ldar)) != 0)
            return e;
        }
        if (d.P > 1 && (e = ops.bcast_col(pr_k, w.pan, (size_t)ldar * pcol)) != 0)
          return e;
      }
      if (lrows <= 0)
        continue;
      const long long ldx = tri_dist_ld(lrows);
      const T* arem = w.pan + (long long)(rem_begin - c_begin) * ldar;  // A(k, rem): wk x rem_cols
      if (bcast_pattern) {
        if (own_col) {
          if ((e = ops.lacpy('G', lrows, wk, bk, ldb, w.xb, ldx)) != 0)
            return e;
          if ((e = ops.transpose(wk, wk, w.pan + diag_off, ldar, w.wkh, ld_wkh, 1, 'G')) != 0)  // W_k^H
            return e;
          if ((e = ops.gemm_nt(lrows, wk, wk, R(1), w.xb, ldx, w.wkh, ld_wkh, R(0), bk, ldb, 0)) != 0)  // X_k = B_k W_k
            return e;
          if (more && (e = ops.lacpy('G', lrows, wk, bk, ldb, w.xb, ldx)) != 0)
            return e;
        }
        if (!more)
          continue;
        if (d.Q > 1 && (e = ops.bcast_row(qc_k, w.xb, (size_t)ldx * wk)) != 0)
          return e;
        if (rem_cols > 0) {
          const long long ldt2 = tri_dist_ld(lcols);
          if ((e = ops.transpose(wk, rem_cols, arem, ldar, w.pan2, ldt2, 1, 'G')) != 0)  // pan2 = A(k, rem)^H
            return e;
          if ((e = ops.gemm_nt(lrows, rem_cols, wk, R(-1), w.xb, ldx, w.pan2, ldt2, R(1), B + (long long)rem_begin * ldb,
                               ldb, 0)) != 0)
            return e;
        }
      }
      else {
        // xb (lrows x wk) = B_k - sum over the solved local columns of X A(k, :)^H, reduced onto the owners of block k
        if (own_col) {
          if ((e = ops.lacpy('G', lrows, wk, bk, ldb, w.xb, ldx)) != 0)
            return e;
        }
        else if (rem_cols == 0 && d.Q > 1) {
          if ((e = ops.zero(lrows, wk, w.xb, ldx)) != 0)
            return e;
        }
        if (rem_cols > 0 && (own_col || d.Q > 1) &&
            (e = tri_dist_gemm_ksplit<T>(ops, lrows, wk, rem_cols, t, R(-1), B + (long long)rem_begin * ldb, ldb, arem,
                                         ldar, own_col ? R(1) : R(0), w.xb, ldx)) != 0)
          return e;
        if (d.Q > 1 && (e = ops.reduce_row(qc_k, w.xb, (size_t)ldx * wk)) != 0)
          return e;
        if (own_col &&
            (e = ops.gemm_nt(lrows, wk, wk, R(1), w.xb, ldx, w.pan + diag_off, ldar, R(0), bk, ldb, 0)) != 0)  // R_k W_k^H
          return e;
      }
    }
  }
  if (conj_b)
    return ops.scale(lrows, lcols, B, ldb, Tr<T>::from_real(R(1)), 1);
  return 0;
}

}  // namespace dlafb
