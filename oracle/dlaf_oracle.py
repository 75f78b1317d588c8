"""ORACLE — TEST INFRASTRUCTURE ONLY (not the product; never imported by the product package).

CPU restatement (numpy + scipy LAPACK/BLAS for tile arithmetic) of the reference Cholesky path of
DLA-Future v0.6.0. The reference itself cannot be built here; parity is pinned on
the reference's own analytical known-answer tests, which `tests/test_oracle.py` replays against
this file. Citations are relative to /root/reference.

  Distribution index algebra   include/dlaf/matrix/util_distribution.h:27-200, distribution.h
  analytical setters           test/include/dlaf_test/matrix/util_generic_lapack.h:27-68
  local call_L / call_U        include/dlaf/factorization/cholesky/impl.h:150-189, :316-348
  distributed call_L / call_U  include/dlaf/factorization/cholesky/impl.h:191-313, :350-452
  panel broadcast routing      include/dlaf/communication/broadcast_panel.h:59-213
  residual                     miniapp/miniapp_cholesky.cpp:408-442
"""
import ctypes
import glob
import os
import subprocess

import numpy as np
import scipy.linalg.blas as sblas
import scipy.linalg.lapack as slapack

HERE = os.path.dirname(os.path.abspath(__file__))

DTYPES = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}


def type_error(dtype):
    """TypeUtilities<T>::error: 2*eps real, 8*eps complex (test/include/dlaf_test/util_types.h:40,:62)."""
    dtype = np.dtype(dtype)
    eps = np.finfo(dtype).eps
    return (8 if dtype.kind == "c" else 2) * float(eps)


# ----------------------------------------------------------------------------------------------
# Distribution (2D block-cyclic index algebra), one tile per block, no element offset.
# ----------------------------------------------------------------------------------------------
class Distribution:
    """matrix::Distribution restricted to what the Cholesky path uses.

    rank_global_tile: util_distribution.h:85-95; local_tile_from_global_tile: :110-125;
    next_local_tile_from_global_tile: :150-175; local_nr_tiles via next_local(nr_tiles).
    """

    def __init__(self, n, nb, grid=(1, 1), rank=(0, 0), src=(0, 0)):
        self.n, self.nb = int(n), int(nb)
        self.grid, self.rank, self.src = tuple(grid), tuple(rank), tuple(src)
        self.nt = -(-self.n // self.nb) if self.n > 0 else 0

    def rank_global_tile(self, g, axis):
        return (g + self.src[axis]) % self.grid[axis]

    def local_tile_from_global(self, g, axis):
        return g // self.grid[axis]

    def next_local_tile_from_global(self, g, axis, rank=None):
        r = self.rank[axis] if rank is None else rank
        P = self.grid[axis]
        rank_to_src = (r + P - self.src[axis]) % P
        owner_to_src = g % P  # == (rank_global_tile + P - src) % P
        local = g // P
        return local + 1 if rank_to_src < owner_to_src else local

    def local_nr_tiles(self, axis, rank=None):
        return self.next_local_tile_from_global(self.nt, axis, rank)

    def global_tile_from_local(self, l, axis, rank=None):
        r = self.rank[axis] if rank is None else rank
        P = self.grid[axis]
        return l * P + (r + P - self.src[axis]) % P

    def tile_size(self, g):
        return min(self.nb, self.n - g * self.nb)

    def local_size(self, axis, rank=None):
        return sum(self.tile_size(self.global_tile_from_local(l, axis, rank))
                   for l in range(self.local_nr_tiles(axis, rank)))


# ----------------------------------------------------------------------------------------------
# Analytical known-answer pair (getCholeskySetters).
# ----------------------------------------------------------------------------------------------
def cholesky_setters(n, uplo, dtype, rows=None, cols=None):
    """Returns (A, T) restricted to global index arrays rows x cols (default: full n x n).

    t_ij = 2^-|i-j| e^{I(j-i)},  a_ij = (4^(min(i,j)+1)-1)/(3*2^(i+j)) e^{I(j-i)}; the triangle
    that is not referenced holds -9.9 (util_generic_lapack.h:40-68).
    """
    dtype = np.dtype(dtype)
    rows = np.arange(n) if rows is None else np.asarray(rows)
    cols = np.arange(n) if cols is None else np.asarray(cols)
    i = rows[:, None].astype(np.float64)
    j = cols[None, :].astype(np.float64)
    mn = np.minimum(i, j)
    with np.errstate(over="ignore", invalid="ignore"):
        mag_a = np.exp2(-(i + j)) / 3 * (np.exp2(2 * (mn + 1)) - 1)
    # The reference expression overflows double for min(i,j) >= 511 (its own tests stop at n = 128); beyond that
    # the same value is formed as (2^(2min+2-i-j) - 2^-(i+j)) / 3, which has no large intermediate.
    mag_a = np.where(2 * (mn + 1) < 1000, mag_a, (np.exp2(2 * mn + 2 - i - j) - np.exp2(-(i + j))) / 3)
    mag_t = np.exp2(-np.abs(i - j))
    if dtype.kind == "c":
        ph = np.exp(1j * (j - i))
        a, t = mag_a * ph, mag_t * ph
    else:
        a, t = mag_a + 0 * j, mag_t + 0 * j
    unref = (i < j) if uplo == "L" else (i > j)
    a = np.where(unref, -9.9, a)
    t = np.where(unref, -9.9, t)
    return np.asfortranarray(a.astype(dtype)), np.asfortranarray(t.astype(dtype))


def check_near(actual, expected, abs_tol, rel_tol=None):
    """CHECK_MATRIX_NEAR: passes if abs diff <= abs_tol OR rel diff <= rel_tol
    (test/include/dlaf_test/matrix/util_matrix.h:255-283). Returns the number of failing entries."""
    rel_tol = abs_tol if rel_tol is None else rel_tol
    diff = np.abs(actual - expected)
    with np.errstate(divide="ignore", invalid="ignore"):
        rel = diff / np.abs(expected)
    ok = (diff <= abs_tol) | (rel <= rel_tol)
    return int(np.count_nonzero(~ok))


# ----------------------------------------------------------------------------------------------
# Tile kernels (what the reference MC backend calls through lapackpp / blaspp).
# ----------------------------------------------------------------------------------------------
def _f(name, dtype):
    pre = {np.float32: "s", np.float64: "d", np.complex64: "c", np.complex128: "z"}[np.dtype(dtype).type]
    mod = slapack if name == "potrf" else sblas
    if name == "herk" and pre in "sd":
        name = "syrk"
    return getattr(mod, pre + name)


def tile_potrf(uplo, a):
    """lapack/tile.h:363-378. Returns info; a is overwritten in the uplo triangle only."""
    if a.shape[0] == 0:
        return 0
    c, info = _f("potrf", a.dtype)(a, lower=(uplo == "L"), clean=False, overwrite_a=False)
    tri = np.tril if uplo == "L" else np.triu
    mask = tri(np.ones(a.shape, dtype=bool))
    if info == 0:
        a[mask] = c[mask]
    return info


def tile_trsm_L(lkk, b):
    """b <- b * lkk^-H (Right, Lower, ConjTrans, NonUnit, alpha=1), impl.h:55-67."""
    if b.size:
        b[...] = _f("trsm", b.dtype)(1.0, lkk, b, side=1, lower=1, trans_a=2, diag=0)


def tile_trsm_U(ukk, b):
    """b <- ukk^-H * b (Left, Upper, ConjTrans, NonUnit, alpha=1), impl.h:107-119."""
    if b.size:
        b[...] = _f("trsm", b.dtype)(1.0, ukk, b, side=0, lower=0, trans_a=2, diag=0)


def tile_herk_L(a, c):
    """c <- c - a a^H on the lower triangle (impl.h:69-80)."""
    if c.size and a.size:
        r = _f("herk", c.dtype)(-1.0, a, beta=1.0, c=c, trans=0, lower=1)
        m = np.tril(np.ones(c.shape, dtype=bool))
        c[m] = r[m]


def tile_herk_U(a, c):
    """c <- c - a^H a on the upper triangle (impl.h:121-132)."""
    if c.size and a.size:
        tr = 2 if np.dtype(c.dtype).kind == "c" else 1
        r = _f("herk", c.dtype)(-1.0, a, beta=1.0, c=c, trans=tr, lower=0)
        m = np.triu(np.ones(c.shape, dtype=bool))
        c[m] = r[m]


def tile_gemm_L(a, b, c):
    """c <- c - a b^H (impl.h:82-94)."""
    if c.size and a.size:
        c[...] = _f("gemm", c.dtype)(-1.0, a, b, beta=1.0, c=c, trans_a=0, trans_b=2)


def tile_gemm_U(a, b, c):
    """c <- c - a^H b (impl.h:134-146)."""
    if c.size and a.size:
        c[...] = _f("gemm", c.dtype)(-1.0, a, b, beta=1.0, c=c, trans_a=2, trans_b=0)


# ----------------------------------------------------------------------------------------------
# Local tiled algorithm (impl.h:150-189, :316-348).
# ----------------------------------------------------------------------------------------------
def cholesky_local(uplo, a, nb):
    """In-place tiled right-looking Cholesky of the column-major array `a` (n x n). Returns info
    (0, or k*nb + tile info of the first failing diagonal tile: ScaLAPACK-style global column)."""
    n = a.shape[0]
    nt = -(-n // nb) if n else 0
    t = lambda i, j: a[i * nb:min(n, (i + 1) * nb), j * nb:min(n, (j + 1) * nb)]
    for k in range(nt):
        info = tile_potrf(uplo, t(k, k))
        if info:
            return k * nb + info
        for i in range(k + 1, nt):
            if uplo == "L":
                tile_trsm_L(t(k, k), t(i, k))
            else:
                tile_trsm_U(t(k, k), t(k, i))
        for j in range(k + 1, nt):
            if uplo == "L":
                tile_herk_L(t(j, k), t(j, j))
                for i in range(j + 1, nt):
                    tile_gemm_L(t(i, k), t(j, k), t(i, j))
            else:
                tile_herk_U(t(k, j), t(j, j))
                for i in range(j + 1, nt):
                    tile_gemm_U(t(k, j), t(k, i), t(j, i))
    return 0


# ----------------------------------------------------------------------------------------------
# Distributed algorithm with simulated ranks (impl.h:191-313, :350-452; broadcast_panel.h).
# ----------------------------------------------------------------------------------------------
def scatter(a_global, nb, grid, src=(0, 0)):
    """Global matrix -> {rank: dict[(gi,gj)] -> tile copy} following rank_global_tile."""
    n = a_global.shape[0]
    d0 = Distribution(n, nb, grid, (0, 0), src)
    local = {(r, c): {} for r in range(grid[0]) for c in range(grid[1])}
    for gi in range(d0.nt):
        for gj in range(d0.nt):
            owner = (d0.rank_global_tile(gi, 0), d0.rank_global_tile(gj, 1))
            local[owner][(gi, gj)] = np.array(
                a_global[gi * nb:min(n, (gi + 1) * nb), gj * nb:min(n, (gj + 1) * nb)], order="F")
    return local


def gather(local, n, nb, dtype):
    out = np.zeros((n, n), dtype=dtype, order="F")
    for tiles in local.values():
        for (gi, gj), t in tiles.items():
            out[gi * nb:gi * nb + t.shape[0], gj * nb:gj * nb + t.shape[1]] = t
    return out


def panel_broadcast_plan(dist, k, axis="col"):
    """Tile routing of one step of `broadcast(root, panel, panelT, row_chain, col_chain)`
    (broadcast_panel.h:123-213) for the column panel k of a Lower factorization (axis='col') or
    the row panel k of an Upper factorization (axis='row'), seen from rank `dist.rank`.

    Returns (step1, step2):
      step1: list of (global_tile_index_along_panel, root_rank_in_orthogonal_axis) this rank holds
             in `panel` after the first broadcast (root = rank owning grid column/row of k);
      step2: list of (global_tile_index, root_rank_along_panel_axis) for every slot of `panelT`
             this rank fills; the LAST global tile is not transposed (:186-191) unless this rank
             does not own it.
    """
    kt = k + 1
    ax_p = 0 if axis == "col" else 1      # axis along which panel tiles are indexed
    ax_t = 1 - ax_p                       # axis along which panelT tiles are indexed
    root1 = dist.rank_global_tile(k, ax_t)
    step1 = []
    for l in range(dist.next_local_tile_from_global(kt, ax_p), dist.local_nr_tiles(ax_p)):
        step1.append((dist.global_tile_from_local(l, ax_p), root1))
    step2 = []
    last = max(kt, dist.nt - 1)
    lo = dist.next_local_tile_from_global(kt, ax_t)
    hi = dist.local_nr_tiles(ax_t)
    if dist.rank[ax_t] == dist.rank_global_tile(last, ax_t):
        hi -= 1
    for l in range(lo, hi):
        g = dist.global_tile_from_local(l, ax_t)
        step2.append((g, dist.rank_global_tile(g, ax_p)))
    return step1, step2


def cholesky_distributed(uplo, local, n, nb, grid, src=(0, 0)):
    """Runs call_L / call_U over simulated ranks. `local` is the output of scatter() and is
    updated in place. Each rank only ever touches its own tiles plus what a broadcast delivers;
    the asserts check the routing tables of panel_broadcast_plan against tile ownership."""
    P, Q = grid
    ranks = [(r, c) for r in range(P) for c in range(Q)]
    dists = {rc: Distribution(n, nb, grid, rc, src) for rc in ranks}
    d0 = dists[(0, 0)]
    nt = d0.nt
    lower = uplo == "L"
    for k in range(nt):
        kk_rank = (d0.rank_global_tile(k, 0), d0.rank_global_tile(k, 1))
        info = tile_potrf(uplo, local[kk_rank][(k, k)])
        if info:
            return k * nb + info
        kt = k + 1
        if kt == nt:
            continue
        lkk = local[kk_rank][(k, k)]
        # diagonal tile broadcast along the column (L) / row (U) communicator, then panel update
        for rc in ranks:
            if lower and rc[1] == kk_rank[1]:
                for (gi, gj), t in local[rc].items():
                    if gj == k and gi > k:
                        tile_trsm_L(lkk, t)
            if not lower and rc[0] == kk_rank[0]:
                for (gi, gj), t in local[rc].items():
                    if gi == k and gj > k:
                        tile_trsm_U(lkk, t)
        # panel / panelT broadcast (routing per broadcast_panel.h)
        panels, panelsT = {}, {}
        for rc in ranks:
            d = dists[rc]
            step1, step2 = panel_broadcast_plan(d, k, "col" if lower else "row")
            pan, panT = {}, {}
            for g, root in step1:
                owner = (rc[0], root) if lower else (root, rc[1])
                key = (g, k) if lower else (k, g)
                assert key in local[owner], (rc, key, owner)
                pan[g] = local[owner][key]
            for g, root in step2:
                # the tile comes from the rank (in my column / row) that holds it in ITS panel
                holder = (root, rc[1]) if lower else (rc[0], root)
                s1, _ = panel_broadcast_plan(dists[holder], k, "col" if lower else "row")
                assert g in [x for x, _ in s1], (rc, g, holder)
                src_owner = (holder[0], s1[0][1]) if lower else (s1[0][1], holder[1])
                key = (g, k) if lower else (k, g)
                panT[g] = local[src_owner][key]
            panels[rc], panelsT[rc] = pan, panT
        # trailing matrix
        for rc in ranks:
            pan, panT = panels[rc], panelsT[rc]
            for (gi, gj), t in local[rc].items():
                if lower and gj >= kt and gi >= gj:
                    if gi == gj:
                        tile_herk_L(pan[gi], t)
                    else:
                        tile_gemm_L(pan[gi], panT[gj], t)
                if not lower and gi >= kt and gj >= gi:
                    if gi == gj:
                        tile_herk_U(pan[gj], t)
                    else:
                        tile_gemm_U(panT[gi], pan[gj], t)
    return 0


# ----------------------------------------------------------------------------------------------
# Residuals.
# ----------------------------------------------------------------------------------------------
def residual_max(a, l, uplo="L"):
    """max|A - L L^H| / max|A| over the referenced triangle (miniapp_cholesky.cpp:408-442).
    Returns (ratio, eps*n) so that callers can apply the miniapp thresholds."""
    n = a.shape[0]
    if n == 0:
        return 0.0, 0.0
    if uplo == "L":
        f = np.tril(l)
        d = np.tril(a - f @ f.conj().T)
        na = np.abs(np.tril(a)).max()
    else:
        f = np.triu(l)
        d = np.triu(a - f.conj().T @ f)
        na = np.abs(np.triu(a)).max()
    return float(np.abs(d).max() / na), float(np.finfo(a.dtype).eps * n)


def residual_inf(a_herm, l, uplo="L"):
    """north_star form: ||A - L L^H||_inf / (||A||_inf * n * eps), A given as full Hermitian."""
    n = a_herm.shape[0]
    f = np.tril(l) if uplo == "L" else np.triu(l).conj().T
    d = a_herm - f @ f.conj().T
    return float(np.abs(d).sum(axis=1).max() /
                 (np.abs(a_herm).sum(axis=1).max() * n * np.finfo(a_herm.dtype).eps))


def residual_probe(a_full_herm_matvec, l, x):
    """Size-independent check for big n: || A x - L (L^H x) || / (||A x||) with O(n^2) work.
    `a_full_herm_matvec(x)` returns A @ x."""
    f = np.tril(l)
    ax = a_full_herm_matvec(x)
    r = ax - f @ (f.conj().T @ x)
    return float(np.linalg.norm(r) / np.linalg.norm(ax))


# ----------------------------------------------------------------------------------------------
# C++ oracle (threaded tile tasks + exact mt19937_64 generator): build + ctypes binding.
# ----------------------------------------------------------------------------------------------
_LIB = None


def blas_path():
    import scipy
    c = sorted(glob.glob(os.path.join(os.path.dirname(os.path.dirname(scipy.__file__)), "scipy.libs",
                                      "libscipy_openblas*.so")))
    if not c:
        raise RuntimeError("scipy-bundled OpenBLAS not found")
    return c[0]


def build(force=False):
    so = os.path.join(HERE, "_build", "liboracle.so")
    src = os.path.join(HERE, "chol_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(so), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-fopenmp", "-shared", "-fPIC", "-o", so, src, "-ldl"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(HERE, "_build", "liboracle.so")
        if not os.path.exists(so):
            build()
        L = ctypes.CDLL(so)
        if L.oracle_set_blas(blas_path().encode(), 1) != 0:
            raise RuntimeError("oracle: cannot dlopen OpenBLAS")
        i64, vp = ctypes.c_int64, ctypes.c_void_p
        for x in "sdcz":
            getattr(L, f"oracle_gen_hpd_{x}").argtypes = [i64, i64, vp, i64]
            getattr(L, f"oracle_gen_hpd_{x}").restype = None
            getattr(L, f"oracle_cholesky_local_{x}").argtypes = [ctypes.c_char, i64, i64, vp, i64]
            getattr(L, f"oracle_residual_lower_{x}").argtypes = [i64, vp, i64, vp, i64,
                                                                 ctypes.POINTER(ctypes.c_double)]
            getattr(L, f"oracle_residual_lower_{x}").restype = ctypes.c_double
        L.oracle_lapack_potrf_d.argtypes = [ctypes.c_char, i64, vp, i64]
        L.oracle_set_threads.argtypes = [ctypes.c_int, ctypes.c_int]
        L.oracle_set_threads.restype = None
        _LIB = L
    return _LIB


def _x(dtype):
    return {np.float32: "s", np.float64: "d", np.complex64: "c", np.complex128: "z"}[np.dtype(dtype).type]


def gen_hpd(n, nb, dtype):
    """set_random_hermitian_positive_definite on a 1x1 grid: full Hermitian n x n, column-major."""
    a = np.zeros((n, n), dtype=dtype, order="F")
    if n:
        getattr(lib(), f"oracle_gen_hpd_{_x(dtype)}")(n, nb, a.ctypes.data, a.strides[1] // a.itemsize)
    return a


def cholesky_local_threaded(uplo, a, nb, task_threads=None, blas_threads=1):
    """C++ tile-task Cholesky (one single-threaded BLAS call per tile task) — the CPU baseline."""
    L = lib()
    if task_threads is None:
        task_threads = os.cpu_count() or 1
    L.oracle_set_threads(task_threads, blas_threads)
    assert a.flags.f_contiguous
    n = a.shape[0]
    return getattr(L, f"oracle_cholesky_local_{_x(a.dtype)}")(uplo.encode(), n, nb, a.ctypes.data,
                                                              max(1, a.strides[1] // a.itemsize))


# ----------------------------------------------------------------------------------------------
# Triangular solver: the consumer of the Cholesky factor.
#   analytical system   test/include/dlaf_test/matrix/util_generic_blas.h:235-371
#   local algorithms    include/dlaf/solver/triangular/impl.h:235-470 (call_LLN ... call_RUT)
#   test sizes/tolerance  test/unit/solver/test_triangular.cpp:57-106 (40 (m+1) err, alpha = (-1.2, .7))
# ----------------------------------------------------------------------------------------------
TRIANGULAR_TEST_SIZES = [  # (m, n, mb, nb), test_triangular.cpp:57-63
    (0, 0, 1, 1), (0, 2, 1, 2), (7, 0, 2, 1), (2, 2, 5, 5), (10, 10, 2, 3), (7, 7, 3, 2), (3, 2, 7, 7), (12, 3, 5, 5),
    (7, 6, 3, 2), (15, 7, 3, 5), (2, 3, 7, 7), (4, 13, 5, 5), (7, 8, 2, 9), (19, 25, 6, 5),
]


def triangular_test_alpha(dtype):
    """TypeUtilities<T>::element(-1.2, .7): the imaginary part only exists for complex types."""
    dtype = np.dtype(dtype)
    return dtype.type(-1.2 + 0.7j) if dtype.kind == "c" else dtype.type(-1.2)


def _polar(r, theta, dtype):
    dtype = np.dtype(dtype)
    return (r * np.exp(1j * theta)).astype(dtype) if dtype.kind == "c" else (r + 0 * theta).astype(dtype)


def triangular_system(side, uplo, op, diag, alpha, m, n, dtype):
    """getTriangularSystem: returns (A, B, X) with op(A) X = alpha B (side 'L') or X op(A) = alpha B ('R').

    A is what the reference's `set(mat_a, el_op_a, op)` stores: A = el_op_a for 'N', its transpose for 'T', its
    conjugate transpose for 'C'; elements outside the uplo triangle, and the diagonal when diag == 'U', are -9.9.
    """
    dtype = np.dtype(dtype)
    op_a_lower = (uplo == "L") == (op == "N")
    na = m if side == "L" else n
    r = np.arange(na, dtype=np.float64)[:, None]
    c = np.arange(na, dtype=np.float64)[None, :]
    i = np.arange(m, dtype=np.float64)[:, None]
    j = np.arange(n, dtype=np.float64)[None, :]
    if side == "L":
        op_a = _polar((r + 1) / (c + .5), 2 * r - c, dtype)
        x = _polar((i + .5) / (j + 2), i + j, dtype)
        kk = (i + 1) if op_a_lower else (m - i)
        gamma = _polar((i + 1) / (j + 2), 2 * i + j, dtype)
    else:
        op_a = _polar((c + 1) / (r + .5), 2 * c - r, dtype)
        x = _polar((j + .5) / (i + 2), i + j, dtype)
        kk = (n - j) if op_a_lower else (j + 1)
        gamma = _polar((j + 1) / (i + 2), i + 2 * j, dtype)
    unref = (r < c) if op_a_lower else (r > c)
    if diag == "U":
        unref = unref | (r == c)
    op_a = np.where(unref, dtype.type(-9.9), op_a)
    kk = (kk + 0 * (i + j)).astype(dtype)
    if diag == "U":
        b = ((kk - 1) * gamma + x) / alpha
    else:
        b = kk * gamma / alpha
    a = op_a if op == "N" else op_a.T if op == "T" else op_a.conj().T
    return np.asfortranarray(a.astype(dtype)), np.asfortranarray(b.astype(dtype)), np.asfortranarray(x.astype(dtype))


def _tile_trsm(side, uplo, op, diag, alpha, a, b):
    """tile::trsm (blas/tile.h:359-366): b <- alpha op(a)^-1 b or alpha b op(a)^-1."""
    if b.size:
        b[...] = _f("trsm", b.dtype)(alpha, a, b, side=int(side == "R"), lower=int(uplo == "L"),
                                     trans_a={"N": 0, "T": 1, "C": 2}[op], diag=int(diag == "U"))


def _opm(a, op):
    return a if op == "N" else a.T if op == "T" else a.conj().T


def triangular_solver_local(side, uplo, op, diag, alpha, a, b, mb, nb):
    """In-place tiled solve of op(A) X = alpha B / X op(A) = alpha B following the eight local loops of
    solver/triangular/impl.h:235-470: per step a tile trsm with `alpha` on the k-th panel of B, then tile gemms with
    beta = -1/alpha on the tiles still to be solved. A has (mb x mb) tiles for side 'L', (nb x nb) for 'R'."""
    m, n = b.shape
    mt = -(-m // mb) if m else 0
    nt = -(-n // nb) if n else 0
    alpha = b.dtype.type(alpha)
    beta = b.dtype.type(-1.0) / alpha
    ab = mb if side == "L" else nb
    ta = lambda i, j: a[i * ab:(i + 1) * ab, j * ab:(j + 1) * ab]
    tb = lambda i, j: b[i * mb:(i + 1) * mb, j * nb:(j + 1) * nb]
    forward = {("L", "L", "N"): True, ("L", "L", "T"): False, ("L", "U", "N"): False, ("L", "U", "T"): True,
               ("R", "L", "N"): False, ("R", "L", "T"): True, ("R", "U", "N"): True, ("R", "U", "T"): False}[
        (side, uplo, "N" if op == "N" else "T")]
    if side == "L":
        order = range(mt) if forward else range(mt - 1, -1, -1)
        for k in order:
            rest = range(k + 1, mt) if forward else range(k - 1, -1, -1)
            for j in range(nt):
                _tile_trsm(side, uplo, op, diag, alpha, ta(k, k), tb(k, j))
                for i in rest:
                    aik = ta(i, k) if op == "N" else _opm(ta(k, i), op)
                    tb(i, j)[...] += beta * (aik @ tb(k, j))
    else:
        order = range(nt) if forward else range(nt - 1, -1, -1)
        for k in order:
            rest = range(k + 1, nt) if forward else range(k - 1, -1, -1)
            for i in range(mt):
                _tile_trsm(side, uplo, op, diag, alpha, ta(k, k), tb(i, k))
                for j in rest:
                    akj = ta(k, j) if op == "N" else _opm(ta(j, k), op)
                    tb(i, j)[...] += beta * (tb(i, k) @ akj)
    return b


# ----------------------------------------------------------------------------------------------
# Reduction of the generalized Hermitian problem to standard form (local):
#   include/dlaf/eigensolver/gen_to_std.h:23-77 (front end), gen_to_std/impl.h:222-283 (call_L), :508-569 (call_U)
#   analytical test: test/unit/eigensolver/test_gen_to_std.cpp:53-83,
#                    test/include/dlaf_test/matrix/util_generic_lapack.h:70-149 (getGenToStdElementSetters)
# ----------------------------------------------------------------------------------------------
GEN_TO_STD_TEST_SIZES = [(0, 2), (5, 8), (34, 34), (4, 3), (16, 10), (34, 13), (32, 5)]  # test_gen_to_std.cpp:53-57


def gen_to_std_setters(n, uplo, dtype, alpha=-2.0, beta=1.5, gamma=0.95):
    """getGenToStdElementSetters with itype = 1: returns (T, A, RES); T is the Cholesky factor, A the Hermitian
    matrix, RES = inv(T) A inv(T)^H (uplo 'L') or inv(T)^H A inv(T) (uplo 'U'); the triangle that is not referenced
    holds -9.9 in all three."""
    dtype = np.dtype(dtype)
    i = np.arange(n, dtype=np.float64)[:, None]
    j = np.arange(n, dtype=np.float64)[None, :]
    t = _polar(beta / np.exp2(np.abs(i - j)), alpha * (i - j), dtype)
    a = _polar((i + 1) * (j + 1) * (beta * beta * gamma) / np.exp2(i + j), alpha * (i - j), dtype)
    r = _polar(gamma / np.exp2(i + j), alpha * (i - j), dtype)
    unref = (i < j) if uplo == "L" else (i > j)
    out = []
    for x in (t, a, r):
        out.append(np.asfortranarray(np.where(unref, dtype.type(-9.9), x).astype(dtype)))
    return tuple(out)


def _herm_from(uplo, a):
    """Full Hermitian tile from its referenced triangle (what hemm / hegst read)."""
    tri = np.tril(a) if uplo == "L" else np.triu(a)
    full = tri + tri.conj().T
    full[np.diag_indices_from(full)] = np.diag(a).real if a.dtype.kind == "c" else np.diag(a)
    return full


def _tile_hegst(uplo, a, b):
    """tile::hegst(itype = 1, uplo, a, b) (lapack/tile.h:253-263 CPU, :550-575 GPU): LAPACK ?sygst / ?hegst."""
    pre = {np.float32: "s", np.float64: "d", np.complex64: "c", np.complex128: "z"}[a.dtype.type]
    fn = getattr(slapack, pre + ("hegst" if pre in "cz" else "sygst"))
    res, info = fn(a, b, itype=1, lower=int(uplo == "L"))
    assert info == 0
    keep = np.tril(np.ones(a.shape, bool)) if uplo == "L" else np.triu(np.ones(a.shape, bool))
    a[keep] = res[keep]


def _tile_her2k(uplo, x, y, c):
    """c (uplo triangle) <- c - x y^H - y x^H."""
    upd = c - x @ y.conj().T - y @ x.conj().T
    keep = np.tril(np.ones(c.shape, bool)) if uplo == "L" else np.triu(np.ones(c.shape, bool))
    c[keep] = upd[keep]


def gen_to_std_local(uplo, a, t, nb):
    """In-place tile loops of GenToStd::call_L / call_U (local). `a` holds the Hermitian matrix in its uplo triangle,
    `t` the Cholesky factor; only the uplo triangles are referenced, only the uplo triangle of `a` is written."""
    n = a.shape[0]
    nt = -(-n // nb) if n else 0
    ta = lambda i, j: a[i * nb:(i + 1) * nb, j * nb:(j + 1) * nb]
    tt = lambda i, j: t[i * nb:(i + 1) * nb, j * nb:(j + 1) * nb]
    half = a.dtype.type(0.5)
    for k in range(nt):
        _tile_hegst(uplo, ta(k, k), tt(k, k))
        if k == nt - 1:
            continue
        akk = _herm_from(uplo, ta(k, k))
        if uplo == "L":
            for i in range(k + 1, nt):
                _tile_trsm("R", "L", "C", "N", 1, tt(k, k), ta(i, k))
                ta(i, k)[...] -= half * (tt(i, k) @ akk)
            for j in range(k + 1, nt):
                _tile_her2k("L", ta(j, k), tt(j, k), ta(j, j))
                for i in range(j + 1, nt):
                    ta(i, j)[...] -= ta(i, k) @ tt(j, k).conj().T
                    ta(i, j)[...] -= tt(i, k) @ ta(j, k).conj().T
            for i in range(k + 1, nt):
                ta(i, k)[...] -= half * (tt(i, k) @ akk)
            for j in range(k + 1, nt):
                _tile_trsm("L", "L", "N", "N", 1, tt(j, j), ta(j, k))
                for i in range(j + 1, nt):
                    ta(i, k)[...] -= tt(i, j) @ ta(j, k)
        else:
            for i in range(k + 1, nt):
                _tile_trsm("L", "U", "C", "N", 1, tt(k, k), ta(k, i))
                ta(k, i)[...] -= half * (akk @ tt(k, i))
            for i in range(k + 1, nt):
                # her2k(Upper, ConjTrans): c <- c - x^H y - y^H x
                _tile_her2k("U", ta(k, i).conj().T, tt(k, i).conj().T, ta(i, i))
                for j in range(i + 1, nt):
                    ta(i, j)[...] -= ta(k, i).conj().T @ tt(k, j)
                    ta(i, j)[...] -= tt(k, i).conj().T @ ta(k, j)
            for i in range(k + 1, nt):
                ta(k, i)[...] -= half * (akk @ tt(k, i))
            for i in range(k + 1, nt):
                _tile_trsm("R", "U", "N", "N", 1, tt(i, i), ta(k, i))
                for j in range(i + 1, nt):
                    ta(k, j)[...] -= ta(k, i) @ tt(i, j)
    return a


# ----------------------------------------------------------------------------------------------
# Local blocks of a rectangular 2D block-cyclic matrix (matrix::Distribution with (mb x nb) tiles,
# include/dlaf/matrix/distribution.h, util_distribution.h:85-175): index lists per rank.
# ----------------------------------------------------------------------------------------------
def local_indices(n, nb, nprocs, rank, src):
    """Global element indices, in local order, of the part of an axis of n elements (tiles of nb) that `rank` of
    `nprocs` owns when tile 0 belongs to `src` (rank_global_tile / global_tile_from_local_tile)."""
    idx = []
    nt = -(-n // nb) if n else 0
    for g in range(nt):
        if (g + src) % nprocs == rank:
            idx.extend(range(g * nb, min(n, (g + 1) * nb)))
    return np.asarray(idx, dtype=np.int64)


def local_block(a_global, mb, nb, grid, rank, src=(0, 0)):
    """The local block of `rank` = (pr, pc) as a Fortran-ordered copy."""
    m, n = a_global.shape
    ri = local_indices(m, mb, grid[0], rank[0], src[0])
    ci = local_indices(n, nb, grid[1], rank[1], src[1])
    return np.asfortranarray(a_global[np.ix_(ri, ci)]), ri, ci
