// Reduction of a Hermitian definite generalized eigenproblem to standard form on one GPU:
//   uplo 'L':  A <- inv(L) A inv(L)^H      uplo 'U':  A <- inv(U)^H A inv(U)
// with L / U the Cholesky factor of B, A Hermitian, both referenced (and A written) in their uplo triangle only.
// Reference: dlaf::eigensolver::internal::generalized_to_standard, include/dlaf/eigensolver/gen_to_std.h:23-77, and
// the tile loops GenToStd::call_L / call_U of include/dlaf/eigensolver/gen_to_std/impl.h:222-283, :508-569 (LAPACK's
// blocked xHEGST, itype = 1: per block column k  hegst, trsm, hemm, her2k + gemm pairs, hemm, then a tile-by-tile
// triangular solve of the column panel with the trailing factor).
//
// B200 design: everything is expressed with the one contraction the machine runs at 94 % of the FP64 tensor peak,
// C <- beta C + alpha X Y^H (gemm_dmma.cu), on operands as large as the mathematics allows:
//
//   per block column k (W_k = inv(L_kk), inverted up front for all k in ONE launch, r = the rows below block k):
//     S      = full Hermitian copy of A_kk            (copy of the stored triangle + its conjugate transpose)
//     A_kk   = W_k S W_k^H                             two nb^3 contractions                      [tile hegst]
//     P      = A_rk W_k^H - 1/2 L_rk A_kk              two contractions into the workspace        [trsm + hemm]
//     A_rr  -= [P | L_rk] [L_rk | P]^H  (lower part)   ONE contraction with K = 2 nb over the whole trailing
//                                                      triangle instead of a her2k and two gemms per tile: C is
//                                                      read and written once per step            [her2k + gemms]
//     A_rk   = P - 1/2 L_rk A_kk                                                                  [second hemm]
//   then, once, for all the column panels together (the reference solves every panel k with the trailing factor
//   L_rr right away, tile by tile: O(nt^2) dependent steps; nothing later reads those panels, so here the solves of
//   all the panels run side by side as ONE forward substitution over the strictly-lower block trapezoid):
//     for row block j = 1 .. nt-1:   X_j = W_j A_{j,0:j}          and   A_{i>j, 0:j} -= L_ij X_j
//   i.e. nt-1 dependent steps, each one a (n - j nb - nb) x (j nb) x nb contraction.
//   Flops: 2/3 n^3 (trailing updates) + 1/3 n^3 (trapezoid) = n^3, the same as the reference's tile algorithm.
//
//   uplo 'U' is the lower algorithm on the conjugate transposes (L = U^H; the lower triangle of A is the conjugate
//   transpose of the stored upper one): three O(n^2) transposes at HBM speed around an O(n^3) computation - the same
//   choice as the Cholesky engine makes for uplo = Upper.
//
// L / U is never written (the reference may modify its diagonal tiles temporarily, gen_to_std.h:44-45).
// The algorithm is a template over the executor `Ops` (see triangular_solver.h) so that tests/emu/ runs this very
// sequence of operations, pointer arithmetic included, on the CPU with NaN-filled workspaces.
#pragma once
#include <cstddef>

#include "common.cuh"
#include "solver_kernels.cuh"

namespace dlafb {

inline long long g2s_even(long long x) { return (x + 1) & ~1LL; }
inline long long g2s_ldp(int n) { return g2s_even(n > 0 ? n : 1); }
// workspace parts, in elements
inline size_t g2s_winv_elems(int n, int nb) { return (size_t)(n > 0 ? ceil_div(n, nb) : 1) * nb * nb; }
inline size_t g2s_tile_elems(int nb) { return (size_t)g2s_even(nb) * nb; }
inline size_t g2s_panel_elems(int n, int nb) { return (size_t)g2s_ldp(n) * 2 * nb; }
inline size_t g2s_copy_elems(int n, char uplo) { return uplo == 'U' ? (size_t)g2s_ldp(n) * (size_t)(n > 0 ? n : 1) : 0; }

template <class T>
struct GenToStdWork {
  T* winv;   // ceil(n / nb) slots of nb x nb: inverses of the diagonal tiles of the factor
  T* akk;    // nb x nb (leading dimension ldk): the full Hermitian diagonal tile
  T* tmp;    // nb x nb (ldk)
  long long ldk;
  T* pa;     // ldp x 2 nb: [P | L_rk]; also the transposed panel of the trapezoid solve
  T* pb;     // ldp x 2 nb: [L_rk | P]
  long long ldp;
  T* acopy;  // uplo 'U' only: n x n (ldc) lower-triangular images of A and of the factor
  T* lcopy;
  long long ldc;
  // look-ahead variant: two lanes (streams); pa and pb are double-buffered, the second copy panel_stride elements on
  int lookahead = 0;
  size_t panel_stride = 0;
};

// events of the look-ahead variant (Ops::signal / Ops::wait); same numbering as the solver's (triangular_solver.h)
enum G2sEvent { kG2sPrologue = 0, kG2sPanel0 = 1, kG2sPanel1 = 2, kG2sBulk0 = 3, kG2sBulk1 = 4 };

// Main loop of the lower algorithm with look-ahead: lane 0 carries the dependent chain of a step (diagonal tile,
// panel P, the update of the NEXT block column only), lane 1 the update of everything beyond it ("bulk"), so that the
// chain of step k+1 runs under the bulk update of step k instead of after it:
//   lane 0:  chain(k) -> [bulk(k-1) done] next(k) -> finish column k -> chain(k+1) -> ...
//   lane 1:  bulk(k-1) -> [panels of step k built] bulk(k) -> ...
// bulk(k-1) also writes the block column that next(k) writes, hence the wait; the panels [P | L_rk], [L_rk | P] are
// double-buffered because bulk(k) reads them while lane 0 builds those of step k+1; the last correction of column k
// (P - 1/2 L_rk A_kk) goes straight into A so that the panel stays read-only once it is published.
// Additional Ops: lane(l), signal(ev), wait(ev) as in triangular_solver.h.
template <class T, class Ops>
int g2s_main_loop_lookahead(Ops& ops, int n, int nb, T* A, long long lda, const T* L, long long ldl,
                            const GenToStdWork<T>& w) {
  using R = typename Tr<T>::Real;
  int e;
  const int nt = ceil_div(n, nb);
  if ((e = ops.signal(kG2sPrologue)) != 0)
    return e;
  ops.lane(1);
  if ((e = ops.wait(kG2sPrologue)) != 0)
    return e;
  ops.lane(0);
  bool bulk_issued[2] = {false, false};
  for (int k = 0; k < nt; ++k) {
    const int k0 = k * nb;
    const int wk = n - k0 < nb ? n - k0 : nb;
    const int r0 = k0 + wk;
    const int m = n - r0;
    T* akk_a = A + k0 + (long long)k0 * lda;
    const T* wkk = w.winv + (size_t)k * nb * nb;
    T* pa = w.pa + (size_t)(k & 1) * w.panel_stride;
    T* pb = w.pb + (size_t)(k & 1) * w.panel_stride;
    // ---- lane 0: diagonal tile
    if ((e = ops.transpose(wk, wk, akk_a, lda, w.akk, w.ldk, 1, 'U')) != 0)
      return e;
    if ((e = ops.lacpy('L', wk, wk, akk_a, lda, w.akk, w.ldk)) != 0)
      return e;
    if ((e = ops.gemm_nt(wk, wk, wk, R(1), wkk, (long long)nb, w.akk, w.ldk, R(0), w.tmp, w.ldk, 0)) != 0)
      return e;
    if ((e = ops.gemm_nt(wk, wk, wk, R(1), w.tmp, w.ldk, wkk, (long long)nb, R(0), w.akk, w.ldk, 1)) != 0)
      return e;
    if ((e = ops.lacpy('L', wk, wk, w.akk, w.ldk, akk_a, lda)) != 0)
      return e;
    if (m == 0)
      break;
    if ((e = ops.transpose(wk, wk, w.akk, w.ldk, w.tmp, w.ldk, 1, 'U')) != 0)
      return e;
    if ((e = ops.lacpy('L', wk, wk, w.akk, w.ldk, w.tmp, w.ldk)) != 0)
      return e;
    const T* herm = w.tmp;
    T* ark = A + r0 + (long long)k0 * lda;
    const T* lrk = L + r0 + (long long)k0 * ldl;
    // ---- lane 0: panels of this step (their buffers were last read by bulk(k-2), which next(k-1) has waited for)
    if ((e = ops.gemm_nt(m, wk, wk, R(1), ark, lda, wkk, (long long)nb, R(0), pa, w.ldp, 0)) != 0)
      return e;
    if ((e = ops.gemm_nt(m, wk, wk, R(-0.5), lrk, ldl, herm, w.ldk, R(1), pa, w.ldp, 0)) != 0)
      return e;
    if ((e = ops.lacpy('G', m, wk, lrk, ldl, pa + (long long)wk * w.ldp, w.ldp)) != 0)
      return e;
    if ((e = ops.lacpy('G', m, wk, lrk, ldl, pb, w.ldp)) != 0)
      return e;
    if ((e = ops.lacpy('G', m, wk, pa, w.ldp, pb + (long long)wk * w.ldp, w.ldp)) != 0)
      return e;
    if ((e = ops.signal(kG2sPanel0 + (k & 1))) != 0)
      return e;
    // ---- lane 0: update of the next block column (rows r0.., columns r0 .. r0 + wn); bulk(k-1) wrote it too
    const int wn = m < nb ? m : nb;
    if (k >= 1 && bulk_issued[(k - 1) & 1] && (e = ops.wait(kG2sBulk0 + ((k - 1) & 1))) != 0)
      return e;
    T* arr = A + r0 + (long long)r0 * lda;
    if ((e = ops.gemm_nt(wn, wn, 2 * wk, R(-1), pa, w.ldp, pb, w.ldp, R(1), arr, lda, 1)) != 0)  // its diagonal tile
      return e;
    if (m > wn && (e = ops.gemm_nt(m - wn, wn, 2 * wk, R(-1), pa + wn, w.ldp, pb, w.ldp, R(1), arr + wn, lda, 0)) != 0)
      return e;
    // ---- lane 0: column k is finished directly in A (the published panel stays untouched)
    if ((e = ops.lacpy('G', m, wk, pa, w.ldp, ark, lda)) != 0)
      return e;
    if ((e = ops.gemm_nt(m, wk, wk, R(-0.5), lrk, ldl, herm, w.ldk, R(1), ark, lda, 0)) != 0)
      return e;
    // ---- lane 1: everything beyond the next block column
    const int mb = m - wn;
    bulk_issued[k & 1] = mb > 0;
    if (mb > 0) {
      ops.lane(1);
      if ((e = ops.wait(kG2sPanel0 + (k & 1))) != 0)
        return e;
      if ((e = ops.gemm_nt(mb, mb, 2 * wk, R(-1), pa + wn, w.ldp, pb + wn, w.ldp, R(1),
                           arr + wn + (long long)wn * lda, lda, 1)) != 0)
        return e;
      if ((e = ops.signal(kG2sBulk0 + (k & 1))) != 0)
        return e;
      ops.lane(0);
    }
  }
  // join: everything lane 1 did is ordered before what follows on lane 0
  for (int q = 0; q < 2; ++q)
    if (bulk_issued[q] && (e = ops.wait(kG2sBulk0 + q)) != 0)
      return e;
  return 0;
}

// Ops interface (all return 0 on success; semantics of the device launchers in tile_ops.cuh):
//   gemm_nt(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower)   C <- beta C + alpha A B^H; lower: rows >= cols only
//   transpose(m, n, in, ldin, out, ldout, conj, tri)               out (n x m) <- op(in), only the tri part written
//   lacpy(uplo, m, n, a, lda, b, ldb)                              b <- a on the uplo part ('G', 'L', 'U')
//   tri_inverse(upper, unit, n, nb, t, ldt, w)                     w slots <- inverses of the diagonal tiles of t
template <class T, class Ops>
int gen_to_std_lower(Ops& ops, int n, int nb, T* A, long long lda, const T* L, long long ldl,
                     const GenToStdWork<T>& w) {
  using R = typename Tr<T>::Real;
  int e;
  if (n <= 0)
    return 0;
  const int nt = ceil_div(n, nb);
  if ((e = ops.tri_inverse(0, 0, n, nb, L, ldl, w.winv)) != 0)
    return e;

  if (w.lookahead && (e = g2s_main_loop_lookahead<T>(ops, n, nb, A, lda, L, ldl, w)) != 0)
    return e;
  for (int k = 0; k < nt && !w.lookahead; ++k) {
    const int k0 = k * nb;
    const int wk = n - k0 < nb ? n - k0 : nb;
    const int r0 = k0 + wk;
    const int m = n - r0;
    T* akk_a = A + k0 + (long long)k0 * lda;
    const T* wkk = w.winv + (size_t)k * nb * nb;
    // full Hermitian image of the diagonal tile: stored triangle + its conjugate transpose above it
    if ((e = ops.transpose(wk, wk, akk_a, lda, w.akk, w.ldk, 1, 'U')) != 0)
      return e;
    if ((e = ops.lacpy('L', wk, wk, akk_a, lda, w.akk, w.ldk)) != 0)
      return e;
    // A_kk <- W S W^H (S Hermitian: W S^H = W S). The second product forms the stored triangle only, with a real
    // diagonal (the contraction's lower mode), as LAPACK's hegst does; its mirror image is what the panel products
    // below read - like the reference's hemm, which reads the stored triangle of A_kk only.
    if ((e = ops.gemm_nt(wk, wk, wk, R(1), wkk, (long long)nb, w.akk, w.ldk, R(0), w.tmp, w.ldk, 0)) != 0)
      return e;
    if ((e = ops.gemm_nt(wk, wk, wk, R(1), w.tmp, w.ldk, wkk, (long long)nb, R(0), w.akk, w.ldk, 1)) != 0)
      return e;
    if ((e = ops.lacpy('L', wk, wk, w.akk, w.ldk, akk_a, lda)) != 0)
      return e;
    if (m == 0)
      continue;
    if ((e = ops.transpose(wk, wk, w.akk, w.ldk, w.tmp, w.ldk, 1, 'U')) != 0)
      return e;
    if ((e = ops.lacpy('L', wk, wk, w.akk, w.ldk, w.tmp, w.ldk)) != 0)
      return e;
    const T* herm = w.tmp;  // full Hermitian A_kk
    T* ark = A + r0 + (long long)k0 * lda;
    const T* lrk = L + r0 + (long long)k0 * ldl;
    T* p = w.pa;                                   // m x wk
    T* pa_l = w.pa + (long long)wk * w.ldp;        // second half of [P | L_rk]
    T* pb_p = w.pb + (long long)wk * w.ldp;        // second half of [L_rk | P]
    // P = A_rk W^H - 1/2 L_rk A_kk
    if ((e = ops.gemm_nt(m, wk, wk, R(1), ark, lda, wkk, (long long)nb, R(0), p, w.ldp, 0)) != 0)
      return e;
    if ((e = ops.gemm_nt(m, wk, wk, R(-0.5), lrk, ldl, herm, w.ldk, R(1), p, w.ldp, 0)) != 0)
      return e;
    // A_rr -= P L_rk^H + L_rk P^H on the lower triangle, as one contraction over K = 2 wk
    if ((e = ops.lacpy('G', m, wk, lrk, ldl, pa_l, w.ldp)) != 0)
      return e;
    if ((e = ops.lacpy('G', m, wk, lrk, ldl, w.pb, w.ldp)) != 0)
      return e;
    if ((e = ops.lacpy('G', m, wk, p, w.ldp, pb_p, w.ldp)) != 0)
      return e;
    if ((e = ops.gemm_nt(m, m, 2 * wk, R(-1), w.pa, w.ldp, w.pb, w.ldp, R(1), A + r0 + (long long)r0 * lda, lda, 1)) != 0)
      return e;
    // A_rk = P - 1/2 L_rk A_kk
    if ((e = ops.gemm_nt(m, wk, wk, R(-0.5), lrk, ldl, herm, w.ldk, R(1), p, w.ldp, 0)) != 0)
      return e;
    if ((e = ops.lacpy('G', m, wk, p, w.ldp, ark, lda)) != 0)
      return e;
  }

  // all the column panels at once: forward substitution with L over the strictly-lower block trapezoid
  for (int j = 1; j < nt; ++j) {
    const int j0 = j * nb;
    const int wj = n - j0 < nb ? n - j0 : nb;
    const int r0 = j0 + wj;
    const int rest = n - r0;
    T* aj = A + j0;  // wj x j0: row block j, columns before the diagonal block
    const T* wjj = w.winv + (size_t)j * nb * nb;
    if ((e = ops.transpose(wj, j0, aj, lda, w.pa, w.ldp, 1, 'G')) != 0)  // pa (j0 x wj) = A_j^H
      return e;
    if ((e = ops.gemm_nt(wj, j0, wj, R(1), wjj, (long long)nb, w.pa, w.ldp, R(0), aj, lda, 0)) != 0)
      return e;
    if (rest > 0) {
      if ((e = ops.transpose(wj, j0, aj, lda, w.pa, w.ldp, 1, 'G')) != 0)  // pa = X_j^H
        return e;
      if ((e = ops.gemm_nt(rest, j0, wj, R(-1), L + r0 + (long long)j0 * ldl, ldl, w.pa, w.ldp, R(1), A + r0, lda, 0)) != 0)
        return e;
    }
  }
  return 0;
}

template <class T, class Ops>
int gen_to_std(Ops& ops, char uplo, int n, int nb, T* A, long long lda, const T* F, long long ldf,
               const GenToStdWork<T>& w) {
  if (n <= 0)
    return 0;
  if (uplo == 'L')
    return gen_to_std_lower<T>(ops, n, nb, A, lda, F, ldf, w);
  int e;
  if ((e = ops.transpose(n, n, A, lda, w.acopy, w.ldc, 1, 'L')) != 0)
    return e;
  if ((e = ops.transpose(n, n, F, ldf, w.lcopy, w.ldc, 1, 'L')) != 0)
    return e;
  if ((e = gen_to_std_lower<T>(ops, n, nb, w.acopy, w.ldc, w.lcopy, w.ldc, w)) != 0)
    return e;
  return ops.transpose(n, n, w.acopy, w.ldc, A, lda, 1, 'U');
}

}  // namespace dlafb
