"""-m gpu parity tests of the tile kernels through the C ABI (include/dlafb_kernels.h) against the oracle.

Shapes, padding and tolerances follow the reference's tile tests:
  potrf  test/unit/test_lapack_tile.cpp:142-205, test_lapack_tile/test_potrf.h:33-77   tol 4(n+1)*err
  trsm   test/unit/test_blas_tile.cpp:232-237, test_blas_tile/test_trsm.h:36-66         tol 10(m+1)*err
  herk   test/unit/test_blas_tile.cpp:136-141, test_blas_tile/test_herk.h:33-89         tol (k+1)*err
  gemm   test/unit/test_blas_tile.cpp:49-56,   test_blas_tile/test_gemm.h:33-67         tol 2(k+1)*err
with err = TypeUtilities<T>::error (2 eps real, 8 eps complex).
"""
import ctypes

import numpy as np
import pytest

import dlaf_oracle as orc
from gpu_util import NP, DevMat, stream_ptr, sync

pytestmark = pytest.mark.gpu

TYPES = "sdcz"


def rnd(rng, shape, dtype):
    a = rng.uniform(-1, 1, shape)
    if np.dtype(dtype).kind == "c":
        a = a + 1j * rng.uniform(-1, 1, shape)
    return a.astype(dtype)


def near(actual, expected, tol):
    return orc.check_near(actual, expected, tol, tol) == 0


# ------------------------------------------------------------------------------------------- potrf
POTRF_SIZES = [(0, 0), (1, 0), (1, 3), (11, 0), (12, 1), (17, 2), (128, 0), (128, 3),  # reference table
               (64, 0), (65, 1), (200, 0), (256, 0), (300, 4), (512, 0)]              # blocked path, nb-sized tile


@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n,extra", POTRF_SIZES)
def test_potrf_tile_analytic(L, x, uplo, n, extra):
    dt = NP[x]
    a, t = orc.cholesky_setters(n, uplo, dt)
    A = DevMat(a, ld=max(1, n) + extra)
    ib = L.dlafb_inner_block(x.encode(), A.ptr, A.ld, n)
    work = DevMat(np.zeros((L.dlafb_potrf_work_elems(x.encode(), n, ib), 1), dtype=dt))
    info = DevMat(np.zeros((1, 1), dtype=np.int32))
    err = getattr(L, f"dlafb_potrf_tile_{x}")(stream_ptr(), uplo.encode(), n, A.ptr, A.ld, work.ptr, ib, info.ptr, 0)
    sync()
    assert err == 0
    assert int(info.get()[0, 0]) == 0
    tol = 4 * (n + 1) * orc.type_error(dt)
    got = A.get()
    assert near(got, t, tol)                       # includes the -9.9 poison of the unreferenced triangle
    unref = np.triu(np.ones((n, n), bool), 1) if uplo == "L" else np.tril(np.ones((n, n), bool), -1)
    assert np.all(got[unref] == dt(-9.9))
    assert np.all(A.pad() == dt(-7.7))


@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n", [1, 17, 128, 200])
def test_potrf_tile_not_positive_definite(L, x, uplo, n):
    """Zero matrix => info == 1 (test_potrf.h:59-77); a late failure reports its 1-based column."""
    dt = NP[x]
    A = DevMat(np.zeros((n, n), dtype=dt))
    ib = L.dlafb_inner_block(x.encode(), A.ptr, A.ld, n)
    work = DevMat(np.zeros((L.dlafb_potrf_work_elems(x.encode(), n, ib), 1), dtype=dt))
    info = DevMat(np.zeros((1, 1), dtype=np.int32))
    getattr(L, f"dlafb_potrf_tile_{x}")(stream_ptr(), uplo.encode(), n, A.ptr, A.ld, work.ptr, ib, info.ptr, 0)
    sync()
    assert int(info.get()[0, 0]) == 1
    if n >= 17:
        a = np.eye(n, dtype=dt) * 4
        bad = n - 3
        a[bad, bad] = -1
        A = DevMat(a)
        info = DevMat(np.zeros((1, 1), dtype=np.int32))
        getattr(L, f"dlafb_potrf_tile_{x}")(stream_ptr(), uplo.encode(), n, A.ptr, A.ld, work.ptr, ib, info.ptr, 100)
        sync()
        assert int(info.get()[0, 0]) == 100 + bad + 1


@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("n", [128, 512])
def test_potrf_tile_inverse_blocks(L, x, n):
    """The workspace holds inv(L_jj) of every inner diagonal block (what the panel solve multiplies by)."""
    dt = NP[x]
    rng = np.random.default_rng(5)
    g = rnd(rng, (n, n), dt)
    a = g @ g.conj().T + n * np.eye(n, dtype=dt)
    A = DevMat(a)
    ib = L.dlafb_inner_block(x.encode(), A.ptr, A.ld, n)
    work = DevMat(np.zeros((L.dlafb_potrf_work_elems(x.encode(), n, ib), 1), dtype=dt))
    info = DevMat(np.zeros((1, 1), dtype=np.int32))
    assert getattr(L, f"dlafb_potrf_tile_{x}")(stream_ptr(), b"L", n, A.ptr, A.ld, work.ptr, ib, info.ptr, 0) == 0
    sync()
    l = np.tril(A.get())
    ref = np.linalg.cholesky(a.astype(np.complex128 if np.dtype(dt).kind == "c" else np.float64))
    assert near(l, ref.astype(dt), 4 * (n + 1) * orc.type_error(dt) * np.abs(ref).max())
    w = work.get()[:, 0]
    for j in range(n // ib):
        wj = w[j * ib * ib:(j + 1) * ib * ib].reshape(ib, ib).T
        ljj = l[j * ib:(j + 1) * ib, j * ib:(j + 1) * ib]
        assert np.abs(wj @ ljj - np.eye(ib)).max() < 50 * ib * orc.type_error(dt)
        assert np.all(np.triu(wj, 1) == 0)


# -------------------------------------------------------------------------------------------- gemm
GEMM_SIZES = [(0, 0, 0, 0), (7, 0, 2, 3), (0, 5, 2, 1), (1, 1, 0, 0), (3, 2, 7, 1), (12, 20, 11, 3), (8, 8, 11, 0),
              (128, 128, 128, 5),  # reference: up to 128^3 with padded ld
              (8, 24, 16, 0), (10, 128, 128, 0), (256, 256, 128, 0), (384, 128, 128, 0), (200, 136, 64, 0), (130, 270, 48, 2), (512, 512, 512, 0)]


@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("m,n,k,extra", GEMM_SIZES)
def test_gemm_nt(L, x, m, n, k, extra):
    """C <- C - A B^H (impl.h:82-94: NoTrans x ConjTrans, alpha = -1, beta = 1)."""
    dt = NP[x]
    rng = np.random.default_rng(m * 10007 + n * 101 + k)
    a, b, c = rnd(rng, (m, k), dt), rnd(rng, (n, k), dt), rnd(rng, (m, n), dt)
    A, B, C = DevMat(a, max(1, m) + extra), DevMat(b, max(1, n) + extra), DevMat(c, max(1, m) + extra)
    err = getattr(L, f"dlafb_gemm_nt_{x}")(stream_ptr(), m, n, k, -1.0, A.ptr, A.ld, B.ptr, B.ld, 1.0, C.ptr, C.ld, 0)
    sync()
    assert err == 0
    ref = c.copy(order="F")
    orc.tile_gemm_L(np.asfortranarray(a), np.asfortranarray(b), ref)
    tol = 2 * (k + 1) * orc.type_error(dt)
    assert near(C.get(), ref, tol)
    assert np.all(C.pad() == dt(-7.7))


@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("n,k,extra", [(0, 0, 0), (3, 0, 1), (5, 7, 2), (17, 11, 0), (128, 128, 3), (256, 128, 0),
                                       (200, 64, 0), (512, 512, 0)])
def test_herk_lower(L, x, n, k, extra):
    """C <- C - A A^H on the lower triangle, upper triangle untouched (impl.h:69-80)."""
    dt = NP[x]
    rng = np.random.default_rng(n * 31 + k)
    a, c = rnd(rng, (n, k), dt), rnd(rng, (n, n), dt)
    c[np.diag_indices(n)] = c[np.diag_indices(n)].real
    A, C = DevMat(a, max(1, n) + extra), DevMat(c, max(1, n) + extra)
    err = getattr(L, f"dlafb_gemm_nt_{x}")(stream_ptr(), n, n, k, -1.0, A.ptr, A.ld, A.ptr, A.ld, 1.0, C.ptr, C.ld, 1)
    sync()
    assert err == 0
    ref = c.copy(order="F")
    orc.tile_herk_L(np.asfortranarray(a), ref)
    tol = max(1, k + 1) * orc.type_error(dt) * 2   # |a_ik a_jk| <= 2 for complex entries in [-1,1]^2
    got = C.get()
    assert near(np.tril(got), np.tril(ref), tol)
    up = np.triu(np.ones((n, n), bool), 1)
    assert np.all(got[up] == c[up])


# -------------------------------------------------------------------------------------------- trsm
@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("m,nk,extra", [(0, 1, 0), (1, 1, 0), (3, 7, 1), (10, 128, 0), (13, 32, 2), (50, 64, 0), (64, 128, 0),
                                        (300, 128, 0), (1000, 256, 0), (384, 512, 0), (100, 200, 3),
                                        # the one-kernel solve (double, nk % 128 == 0): ragged row blocks, several waves,
                                        # 16- and 32-row CTAs (m > 4736), odd leading dimension of B
                                        (2501, 512, 0), (4805, 384, 0), (5000, 256, 1), (17, 512, 0), (9700, 512, 0)])
def test_trsm_panel(L, x, m, nk, extra):
    """B <- B L^-H (Right, Lower, ConjTrans, NonUnit, alpha 1; impl.h:55-67) after potrf of the diagonal tile."""
    dt = NP[x]
    rng = np.random.default_rng(m * 7 + nk)
    g = rnd(rng, (nk, nk), dt)
    akk = g @ g.conj().T + nk * np.eye(nk, dtype=dt)
    b = rnd(rng, (m, nk), dt)
    Akk, B = DevMat(akk, nk + extra), DevMat(b, max(1, m) + extra)
    ib = L.dlafb_inner_block(x.encode(), Akk.ptr, Akk.ld, nk)
    ib = min(ib, L.dlafb_inner_block(x.encode(), B.ptr, B.ld, nk))
    work = DevMat(np.zeros((L.dlafb_potrf_work_elems(x.encode(), nk, ib), 1), dtype=dt))
    info = DevMat(np.zeros((1, 1), dtype=np.int32))
    s = stream_ptr()
    assert getattr(L, f"dlafb_potrf_tile_{x}")(s, b"L", nk, Akk.ptr, Akk.ld, work.ptr, ib, info.ptr, 0) == 0
    assert getattr(L, f"dlafb_trsm_panel_{x}")(s, nk, Akk.ptr, Akk.ld, work.ptr, ib, B.ptr, B.ld, m) == 0
    sync()
    assert int(info.get()[0, 0]) == 0
    lkk = np.asfortranarray(np.tril(Akk.get()))
    ref = np.asfortranarray(b.copy())
    orc.tile_trsm_L(lkk, ref)
    tol = 10 * (m + 1) * orc.type_error(dt)
    assert near(B.get(), ref, tol)
    assert np.all(B.pad() == dt(-7.7))


@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("m,nk,tile", [(700, 256, 256), (1030, 512, 512), (300, 128, 256), (40, 24, 32), (2600, 512, 512)])
def test_trsm_panel_packed(L, x, m, nk, tile):
    """Solve + copy into the row panel's tile layout (one kernel for double): the packed copy must equal the solved B
    bit for bit, slots beyond the rows stay untouched."""
    dt = NP[x]
    rng = np.random.default_rng(m + nk)
    g = rnd(rng, (nk, nk), dt)
    akk = g @ g.conj().T + nk * np.eye(nk, dtype=dt)
    b = rnd(rng, (m, nk), dt)
    Akk, B = DevMat(akk), DevMat(b, m + 2)
    ib = min(L.dlafb_inner_block(x.encode(), Akk.ptr, Akk.ld, nk), L.dlafb_inner_block(x.encode(), B.ptr, B.ld, nk))
    work = DevMat(np.zeros((L.dlafb_potrf_work_elems(x.encode(), nk, ib), 1), dtype=dt))
    info = DevMat(np.zeros((1, 1), dtype=np.int32))
    nslots = -(-m // tile)
    P = DevMat(np.full((tile * tile * nslots, 1), -3.3, dtype=dt))
    s = stream_ptr()
    assert getattr(L, f"dlafb_potrf_tile_{x}")(s, b"L", nk, Akk.ptr, Akk.ld, work.ptr, ib, info.ptr, 0) == 0
    assert getattr(L, f"dlafb_trsm_panel_packed_{x}")(s, nk, Akk.ptr, Akk.ld, work.ptr, ib, B.ptr, B.ld, m, P.ptr,
                                                      tile) == 0
    sync()
    ref = np.asfortranarray(b.copy())
    orc.tile_trsm_L(np.asfortranarray(np.tril(Akk.get())), ref)
    got = B.get()
    assert near(got, ref, 10 * (m + 1) * orc.type_error(dt))
    assert np.all(B.pad() == dt(-7.7))
    packed = P.get()[:, 0].reshape(nslots, tile, tile)          # [slot][col][row]
    for sl in range(nslots):
        rows = min(tile, m - sl * tile)
        assert np.array_equal(packed[sl, :nk, :rows].T, got[sl * tile: sl * tile + rows, :])
        assert np.all(packed[sl, nk:, :] == dt(-3.3)) and np.all(packed[sl, :, rows:] == dt(-3.3))


# --------------------------------------------------------------------------------------------- aux
@pytest.mark.parametrize("x", TYPES)
@pytest.mark.parametrize("m,n", [(1, 1), (33, 70), (128, 64), (100, 100)])
def test_transpose_laset_lacpy_maxabs(L, x, m, n):
    dt = NP[x]
    rng = np.random.default_rng(m + n)
    a = rnd(rng, (m, n), dt)
    A, O = DevMat(a, m + 1), DevMat(np.zeros((n, m), dtype=dt), n + 2)
    s = stream_ptr()
    assert getattr(L, f"dlafb_transpose_{x}")(s, m, n, A.ptr, A.ld, O.ptr, O.ld, 1, b"G") == 0
    sync()
    assert np.array_equal(O.get(), a.conj().T)
    O2 = DevMat(np.zeros((n, m), dtype=dt))
    assert getattr(L, f"dlafb_transpose_{x}")(s, m, n, A.ptr, A.ld, O2.ptr, O2.ld, 0, b"L") == 0
    sync()
    assert np.array_equal(O2.get(), np.tril(a.T))
    # laset (src/lapack/gpu/laset.cu): off-diagonal alpha, diagonal beta on the uplo part
    for uplo, mask in (("G", np.ones((m, n), bool)), ("L", np.tril(np.ones((m, n), bool))),
                       ("U", np.triu(np.ones((m, n), bool)))):
        B = DevMat(a, m + 1)
        assert getattr(L, f"dlafb_laset_{x}")(s, uplo.encode(), m, n, 0.5, 0.0, 2.0, 0.0, B.ptr, B.ld) == 0
        sync()
        ref = a.copy()
        ref[mask] = 0.5
        for d in range(min(m, n)):
            ref[d, d] = 2.0
        assert np.array_equal(B.get(), ref)
        Cm = DevMat(np.zeros((m, n), dtype=dt), m + 3)
        assert getattr(L, f"dlafb_lacpy_{x}")(s, uplo.encode(), m, n, A.ptr, A.ld, Cm.ptr, Cm.ld) == 0
        out = DevMat(np.zeros((1, 1), dtype=np.float64))
        assert getattr(L, f"dlafb_max_abs_{x}")(s, uplo.encode(), m, n, A.ptr, A.ld, out.ptr) == 0
        sync()
        assert np.array_equal(Cm.get(), np.where(mask, a, 0))
        assert out.get()[0, 0] == pytest.approx(np.abs(np.where(mask, a, 0)).astype(np.float64).max(), rel=1e-6)


@pytest.mark.parametrize("x", TYPES)
def test_gen_hpd_is_hermitian_and_distribution_independent(L, x):
    dt = NP[x]
    n, nb = 96, 16
    full = DevMat(np.zeros((n, n), dtype=dt))
    s = stream_ptr()
    assert getattr(L, f"dlafb_gen_hpd_{x}")(s, n, nb, n, n, 0, 1, 0, 1, full.ptr, full.ld, 42) == 0
    # local block of rank (1, 0) on a 2 x 3 grid: global row tiles 1,3,5 and column tiles 0,3
    loc = DevMat(np.zeros((48, 32), dtype=dt))
    assert getattr(L, f"dlafb_gen_hpd_{x}")(s, n, nb, 48, 32, 1, 2, 0, 3, loc.ptr, loc.ld, 42) == 0
    sync()
    a = full.get()
    assert np.array_equal(a, a.conj().T)
    assert np.all(np.abs(a - np.diag(np.diag(a))) <= 1.0)
    assert np.all(np.abs(np.diag(a).real - 2 * n) <= 1.0) and np.all(np.diag(a).imag == 0)
    rows = np.concatenate([np.arange(t * nb, (t + 1) * nb) for t in (1, 3, 5)])
    cols = np.concatenate([np.arange(t * nb, (t + 1) * nb) for t in (0, 3)])
    assert np.array_equal(loc.get(), a[np.ix_(rows, cols)])
    np.linalg.cholesky(a.astype(np.complex128))
