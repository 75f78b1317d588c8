// Distributed reduction of a Hermitian definite generalized eigenproblem to standard form:
//   uplo 'L':  A <- inv(L) A inv(L)^H      uplo 'U':  A <- inv(U)^H A inv(U)
// A and the Cholesky factor distributed 2D block-cyclically over a P x Q grid of GPUs (same grid, source rank, tiles).
// Reference: dlaf::eigensolver::internal::generalized_to_standard(grid, ...), include/dlaf/eigensolver/gen_to_std.h:79-140,
// GenToStd::call_L / call_U (distributed) include/dlaf/eigensolver/gen_to_std/impl.h:286-505, :572-771.
//
// Same mathematics and the same two phases as the one-GPU algorithm (gen_to_std.h), with the Cholesky factorization's
// panel exchange around the products:
//   per block column k
//     owner of (k, k):          A_kk <- W_k S W_k^H  (S = full Hermitian image of A_kk), then [W_k | A_kk] travels down
//                               the grid column that owns block column k
//     ranks of that column:     P = A_rk W_k^H - 1/2 L_rk A_kk on their rows r > k; the row panel [P | L_rk] (2 nb wide)
//                               travels along the grid rows
//     column panel:             for every local column tile j > k the tile [L_j | P_j] travels down the grid column
//                               from the rank row that owns row tile j (the transposed-panel exchange of the Cholesky
//                               factorization, impl.h:247-251 / broadcast_panel.h:123-213)
//     every rank:               A_ij -= [P_i | L_i] [L_j | P_j]^H on its tiles with i >= j > k: per local column tile ONE
//                               contraction with K = 2 nb over all the local rows below it (+ one for a local diagonal tile,
//                               lower part only)
//     ranks of column k:        A_rk = P - 1/2 L_rk A_kk
//   then the solves of all the column panels with the trailing factor run side by side as ONE distributed forward
//   substitution over the strictly-lower block trapezoid (triangular_solver_dist.h, trapezoid mode).
// uplo 'U': the lower algorithm on the conjugate-transposed local blocks with the grid's axes swapped (the transpose of
// a P x Q block-cyclic matrix is Q x P block-cyclic and every rank already holds its block of it): two local transposes
// in, one out, no communication for them.
// The factor is never written. Template over the executor like triangular_solver_dist.h (run for simulated grids in
// tests/emu/).
#pragma once
#include <cstddef>

#include "common.cuh"
#include "schedule.h"
#include "solver_kernels.cuh"
#include "triangular_solver_dist.h"

namespace dlafb {

struct G2sDistDesc {
  char uplo;
  int n, nb;
  int P, Q, pr, pc, src_r, src_c;
};

// Workspace parts in elements (per rank); sized for both uplo values of the same problem on this rank.
struct G2sDistSizes {
  size_t winv, dk, tmp, xr, ytiles, acopy, lcopy;
};

inline G2sDistSizes g2s_dist_sizes(const G2sDistDesc& d) {
  G2sDistSizes s{};
  // the lower algorithm runs on (P, Q) for uplo L and on the swapped grid for uplo U; local sizes swap with it
  const bool up = d.uplo == 'U';
  const Dist1D R{d.n, d.nb, up ? d.Q : d.P, up ? d.pc : d.pr, up ? d.src_c : d.src_r};
  const Dist1D C{d.n, d.nb, up ? d.P : d.Q, up ? d.pr : d.pc, up ? d.src_r : d.src_c};
  const long long lrows = R.local_size(), lcols = C.local_size();
  const size_t nb = (size_t)d.nb;
  TriDistDesc td{'L', 'L', 'N', 'N', d.n, d.n, d.nb, d.nb, (int)R.P, (int)C.P, (int)R.rank, (int)C.rank, (int)R.src, (int)C.src};
  const TriDistGeom g = tri_dist_geom(td);
  s.winv = (size_t)(g.nd > 0 ? g.nd : 1) * nb * nb;
  s.dk = (size_t)tri_dist_ld(d.nb) * 2 * nb;
  s.tmp = (size_t)tri_dist_ld(d.nb) * nb;
  s.xr = (size_t)tri_dist_ld(lrows) * 2 * nb;
  const size_t slots = (size_t)(C.nloc_tiles() > 0 ? C.nloc_tiles() : 1);
  const size_t yt = slots * (size_t)tri_dist_ld(d.nb) * 2 * nb;
  const size_t xb = (size_t)tri_dist_ld(lcols) * nb;  // the trapezoid solve's transposed block lives here too
  s.ytiles = yt > xb ? yt : xb;
  s.acopy = up ? (size_t)tri_dist_ld(lrows) * (size_t)(lcols > 0 ? lcols : 1) : 0;
  s.lcopy = s.acopy;
  return s;
}

template <class T>
struct G2sDistWork {
  T *winv, *dk, *tmp, *xr, *ytiles, *acopy, *lcopy;
};

// The executor seen from the transposed problem: grid rows and columns change places.
template <class T, class Ops>
struct AxesSwappedOps {
  Ops& o;
  using R = typename Tr<T>::Real;
  int gemm_nt(int M, int N, int K, R alpha, const T* A, long long lda, const T* B, long long ldb, R beta, T* C,
              long long ldc, int lower) {
    return o.gemm_nt(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lower);
  }
  int transpose(int m, int n, const T* in, long long ldin, T* out, long long ldout, int conj, char tri) {
    return o.transpose(m, n, in, ldin, out, ldout, conj, tri);
  }
  int lacpy(char uplo, int m, int n, const T* a, long long lda, T* b, long long ldb) {
    return o.lacpy(uplo, m, n, a, lda, b, ldb);
  }
  int tri_inverse_local(int upper, int unit, int n, int nb, const T* t, long long ldt, T* w, int g0, int gstep, int row0,
                        int row_step, int col0, int col_step, int count) {
    return o.tri_inverse_local(upper, unit, n, nb, t, ldt, w, g0, gstep, row0, row_step, col0, col_step, count);
  }
  int scale(int m, int n, T* b, long long ldb, T alpha, int conj) { return o.scale(m, n, b, ldb, alpha, conj); }
  int zero(int m, int n, T* a, long long lda) { return o.zero(m, n, a, lda); }
  int bcast_row(int root, T* p, size_t count) { return o.bcast_col(root, p, count); }
  int bcast_col(int root, T* p, size_t count) { return o.bcast_row(root, p, count); }
  int reduce_row(int root, T* p, size_t count) { return o.reduce_col(root, p, count); }
  int reduce_col(int root, T* p, size_t count) { return o.reduce_row(root, p, count); }
  int group_begin() { return o.group_begin(); }
  int group_end() { return o.group_end(); }
};

// Lower algorithm on the grid (P, Q, pr, pc, src_r, src_c) as given.
template <class T, class Ops>
int gen_to_std_dist_lower(Ops& ops, int n, int nb, int P, int Q, int pr, int pc, int src_r, int src_c, T* A,
                          long long lda, const T* L, long long ldl, const G2sDistWork<T>& w) {
  using R = typename Tr<T>::Real;
  int e;
  const Dist1D DR{n, nb, P, pr, src_r}, DC{n, nb, Q, pc, src_c};
  const int nt = (int)DR.nt();
  const int lrows = (int)DR.local_size();
  const int nlc = (int)DC.nloc_tiles();
  TriDistDesc td{'L', 'L', 'N', 'N', n, n, nb, nb, P, Q, pr, pc, src_r, src_c};
  td.trapezoid = 1;
  const TriDistGeom g = tri_dist_geom(td);
  if (g.nd > 0) {
    const int row0 = (int)DR.local_of(g.g0), col0 = (int)DC.local_of(g.g0);
    if ((e = ops.tri_inverse_local(0, 0, n, nb, L, ldl, w.winv, g.g0, g.gstep, row0, g.gstep / P, col0, g.gstep / Q,
                                   g.nd)) != 0)
      return e;
  }
  const long long ldk = tri_dist_ld(nb);
  T* dk_w = w.dk;                         // W_k
  T* dk_a = w.dk + (long long)nb * ldk;   // A_kk, full Hermitian
  const size_t slot_elems = (size_t)ldk * 2 * nb;

  for (int k = 0; k < nt; ++k) {
    const int wk = (int)DR.tile_size(k);
    const int pr_k = DR.owner(k), qc_k = DC.owner(k);
    const bool own_row = pr == pr_k, own_col = pc == qc_k;
    const bool more = k + 1 < nt;
    if (own_row && own_col) {
      T* akk = A + DR.local_of(k) * nb + DC.local_of(k) * (long long)nb * lda;
      const T* wkk = w.winv + (size_t)((k - g.g0) / g.gstep) * nb * nb;
      if ((e = ops.transpose(wk, wk, akk, lda, dk_a, ldk, 1, 'U')) != 0)
        return e;
      if ((e = ops.lacpy('L', wk, wk, akk, lda, dk_a, ldk)) != 0)
        return e;
      if ((e = ops.gemm_nt(wk, wk, wk, R(1), wkk, (long long)nb, dk_a, ldk, R(0), w.tmp, ldk, 0)) != 0)
        return e;
      // the stored triangle only, with a real diagonal (see gen_to_std.h); its mirror image travels
      if ((e = ops.gemm_nt(wk, wk, wk, R(1), w.tmp, ldk, wkk, (long long)nb, R(0), dk_a, ldk, 1)) != 0)
        return e;
      if ((e = ops.lacpy('L', wk, wk, dk_a, ldk, akk, lda)) != 0)
        return e;
      if (more) {
        if ((e = ops.transpose(wk, wk, dk_a, ldk, w.tmp, ldk, 1, 'U')) != 0)
          return e;
        if ((e = ops.lacpy('U', wk, wk, w.tmp, ldk, dk_a, ldk)) != 0)  // the diagonal is real: both copies agree on it
          return e;
        if ((e = ops.lacpy('G', wk, wk, wkk, (long long)nb, dk_w, ldk)) != 0)
          return e;
      }
    }
    if (!more)
      break;
    // from here on wk == nb (k is not the last tile)
    if (own_col && P > 1 && (e = ops.bcast_col(pr_k, w.dk, slot_elems)) != 0)
      return e;
    // local rows beyond tile k
    const int il1 = (int)DR.next_local(k + 1);
    const int r_begin = il1 * nb;
    const int prow = lrows > r_begin ? lrows - r_begin : 0;
    const long long ldxr = tri_dist_ld(prow);
    T* xp = w.xr;                          // P
    T* xl = w.xr + (long long)nb * ldxr;   // L_rk
    const long long klc = own_col ? DC.local_of(k) : 0;
    if (prow > 0) {
      if (own_col) {
        const T* ark = A + r_begin + klc * nb * lda;
        const T* lrk = L + r_begin + klc * nb * ldl;
        if ((e = ops.gemm_nt(prow, nb, nb, R(1), ark, lda, dk_w, ldk, R(0), xp, ldxr, 0)) != 0)
          return e;
        if ((e = ops.gemm_nt(prow, nb, nb, R(-0.5), lrk, ldl, dk_a, ldk, R(1), xp, ldxr, 0)) != 0)
          return e;
        if ((e = ops.lacpy('G', prow, nb, lrk, ldl, xl, ldxr)) != 0)
          return e;
      }
      if (Q > 1 && (e = ops.bcast_row(qc_k, w.xr, (size_t)ldxr * 2 * nb)) != 0)
        return e;
    }
    // column panel: tile [L_j | P_j] for every local column tile j > k, from the rank row that owns row tile j
    const int jl1 = (int)DC.next_local(k + 1);
    for (int jl = jl1; jl < nlc; ++jl) {
      const int j = (int)DC.global_of(jl);
      if (pr != DR.owner(j))
        continue;
      const int wj = (int)DC.tile_size(j);
      T* slot = w.ytiles + (size_t)(jl - jl1) * slot_elems;
      const long long rloc = DR.local_of(j) * nb - r_begin;
      if ((e = ops.lacpy('G', wj, nb, xl + rloc, ldxr, slot, ldk)) != 0)
        return e;
      if ((e = ops.lacpy('G', wj, nb, xp + rloc, ldxr, slot + (long long)nb * ldk, ldk)) != 0)
        return e;
    }
    if (P > 1 && jl1 < nlc) {
      // one grouped launch for the whole column panel (the tiles have different roots)
      if ((e = ops.group_begin()) != 0)
        return e;
      for (int jl = jl1; jl < nlc; ++jl) {
        const int root = DR.owner(DC.global_of(jl));
        if ((e = ops.bcast_col(root, w.ytiles + (size_t)(jl - jl1) * slot_elems, slot_elems)) != 0)
          return e;
      }
      if ((e = ops.group_end()) != 0)
        return e;
    }
    // trailing update of the local tiles (i, j), i >= j > k
    if (prow > 0) {
      for (int jl = jl1; jl < nlc; ++jl) {
        const int j = (int)DC.global_of(jl);
        const int wj = (int)DC.tile_size(j);
        const T* slot = w.ytiles + (size_t)(jl - jl1) * slot_elems;
        const int il0 = (int)DR.next_local(j);
        int below = il0 * nb;  // first local row strictly below the diagonal tile of column j
        T* cj = A + (long long)jl * nb * lda;
        if (DR.owner(j) == pr) {
          if ((e = ops.gemm_nt(wj, wj, 2 * nb, R(-1), w.xr + (below - r_begin), ldxr, slot, ldk, R(1), cj + below, lda, 1)) != 0)
            return e;
          below += nb;
        }
        const int mrows = lrows - below;
        if (mrows > 0 && (e = ops.gemm_nt(mrows, wj, 2 * nb, R(-1), w.xr + (below - r_begin), ldxr, slot, ldk, R(1),
                                          cj + below, lda, 0)) != 0)
          return e;
      }
      if (own_col) {
        const T* lrk = L + r_begin + klc * nb * ldl;
        if ((e = ops.gemm_nt(prow, nb, nb, R(-0.5), lrk, ldl, dk_a, ldk, R(1), xp, ldxr, 0)) != 0)
          return e;
        if ((e = ops.lacpy('G', prow, nb, xp, ldxr, A + r_begin + klc * nb * lda, lda)) != 0)
          return e;
      }
    }
  }
  // all the column panels at once: L X = (strictly-lower block trapezoid of A)
  TriDistWork<T> tw{w.winv, w.xr, nullptr, w.ytiles, nullptr, nullptr};
  // the solver inverts the diagonal tiles again into the same slots (same values): one launch, keeps its interface closed
  return triangular_solve_dist<T>(ops, td, Tr<T>::from_real(R(1)), L, ldl, A, lda, tw);
}

template <class T, class Ops>
int gen_to_std_dist(Ops& ops, const G2sDistDesc& d, T* A, long long lda, const T* F, long long ldf,
                    const G2sDistWork<T>& w) {
  if (d.n <= 0)
    return 0;
  if (d.uplo == 'L')
    return gen_to_std_dist_lower<T>(ops, d.n, d.nb, d.P, d.Q, d.pr, d.pc, d.src_r, d.src_c, A, lda, F, ldf, w);
  const Dist1D DR{d.n, d.nb, d.P, d.pr, d.src_r}, DC{d.n, d.nb, d.Q, d.pc, d.src_c};
  const int lrows = (int)DR.local_size(), lcols = (int)DC.local_size();
  const long long ldc = tri_dist_ld(lcols);
  int e;
  if ((e = ops.transpose(lrows, lcols, A, lda, w.acopy, ldc, 1, 'G')) != 0)
    return e;
  if ((e = ops.transpose(lrows, lcols, F, ldf, w.lcopy, ldc, 1, 'G')) != 0)
    return e;
  AxesSwappedOps<T, Ops> sw{ops};
  if ((e = gen_to_std_dist_lower<T>(sw, d.n, d.nb, d.Q, d.P, d.pc, d.pr, d.src_c, d.src_r, w.acopy, ldc, w.lcopy, ldc,
                                    w)) != 0)
    return e;
  return ops.transpose(lcols, lrows, w.acopy, ldc, A, lda, 1, 'G');
}

}  // namespace dlafb
