"""-m gpu tests of the reference-facing C API on one GPU (1x1 grid): dlaf_pdpotrf / dlaf_cholesky_factorization_*
with HOST pointers (mirrored like src/c_api/factorization/cholesky.h:48-55) and with device pointers (in place).

Mirrors test/unit/c_api/factorization/test_cholesky_c_api.cpp: analytical input / factor from getCholeskySetters,
CHECK_MATRIX_NEAR with 4(m+1)*err on the referenced triangle; in addition the unreferenced triangle and the
padding rows of the caller's buffer must come back untouched (only the referenced triangle travels).
"""
import ctypes

import numpy as np
import pytest

import dlaf_oracle as orc
from gpu_util import NP, DevMat, sync

pytestmark = pytest.mark.gpu

CTX = 41


@pytest.fixture(scope="module")
def capi(pkg, L):
    L.dlaf_initialize(0, None, 0, None)
    pkg.check(L.dlafb_mpi_init(0, 1, None), "mpi_init")
    L.dlaf_create_grid_from_blacs_info(CTX, L.dlafb_mpi_comm_world(), 1, 1, b"R")
    yield L
    L.dlaf_free_grid(CTX)
    L.dlaf_finalize()


def host_colmajor(a, ld, fill):
    n = a.shape[0]
    h = np.full((max(1, n), ld), fill, dtype=a.dtype)  # h[j, i] = a[i, j]
    h[:n, :n] = a.T
    return h


def referenced(n, uplo):
    return np.tril(np.ones((n, n), bool)) if uplo == "L" else np.triu(np.ones((n, n), bool))


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n,nb,extra", [(34, 13, 0), (300, 128, 5), (1000, 128, 24), (1024, 256, 0), (1536, 512, 0),
                                          (2304, 128, 8)])   # >= 16 tile columns: upload overlapped with the first steps
def test_pxpotrf_host_pointer(pkg, capi, x, uplo, n, nb, extra):
    L = capi
    dt = NP[x]
    a, t = orc.cholesky_setters(n, uplo, dt)
    ld = n + extra
    ref = referenced(n, uplo)
    a_in = np.where(ref, a, dt(-9.9))  # poison the unreferenced triangle
    h = host_colmajor(a_in, ld, dt(-7.7))
    desca = pkg.DESC9(1, CTX, n, n, nb, nb, 0, 0, ld)
    info = ctypes.c_int(-1)
    getattr(L, f"dlaf_p{x}potrf")(uplo.encode(), n, h.ctypes.data, 1, 1, desca, ctypes.byref(info))
    assert info.value == 0
    got = np.ascontiguousarray(h[:n, :n].T)
    tol = 4 * (n + 1) * orc.type_error(dt)
    assert orc.check_near(np.where(ref, got, 0), np.where(ref, t, 0), tol, tol) == 0
    assert np.all(got[~ref] == dt(-9.9))
    assert np.all(h[:n, n:] == dt(-7.7))


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("uplo", "LU")
def test_cholesky_factorization_host_pointer_random(pkg, capi, x, uplo):
    """Seeded random HPD matrix (oracle port of the reference generator) against the oracle factor + residual."""
    L = capi
    dt = NP[x]
    n, nb = 1280, 256
    a = orc.gen_hpd(n, nb, dt)
    want = a.copy(order="F")
    assert orc.cholesky_local(uplo, want, nb) == 0
    h = host_colmajor(a, n, dt(0))
    d = pkg.DLAFDescriptor(n, n, nb, nb, 0, 0, 0, 0, n)
    r = getattr(L, f"dlaf_cholesky_factorization_{x}")(CTX, uplo.encode(), h.ctypes.data, d)
    assert r == 0
    got = np.ascontiguousarray(h[:n, :n].T)
    ref = referenced(n, uplo)
    assert np.abs(np.where(ref, got - want, 0)).max() < 1e-9
    ratio, thresh = orc.residual_max(a, got, uplo)
    assert ratio <= thresh
    assert np.array_equal(got[~ref], a[~ref])  # untouched


@pytest.mark.parametrize("uplo", "LU")
def test_pdpotrf_device_pointer_in_place(pkg, capi, uplo):
    L = capi
    n, nb = 768, 256
    a, t = orc.cholesky_setters(n, uplo, np.float64)
    A = DevMat(a, ld=n + 64)
    desca = pkg.DESC9(1, CTX, n, n, nb, nb, 0, 0, A.ld)
    info = ctypes.c_int(-1)
    L.dlaf_pdpotrf(uplo.encode(), n, A.ptr, 1, 1, desca, ctypes.byref(info))
    sync()
    assert info.value == 0
    tol = 4 * (n + 1) * orc.type_error(np.float64)
    ref = referenced(n, uplo)
    assert orc.check_near(np.where(ref, A.get(), 0), np.where(ref, t, 0), tol, tol) == 0
    assert np.all(A.pad() == -7.7)


def test_pdpotrf_not_positive_definite_info(pkg, capi):
    """ScaLAPACK convention: info = 1-based global column of the first non-positive pivot."""
    L = capi
    n, nb = 512, 128
    a = orc.gen_hpd(n, nb, np.float64)
    a[300, 300] = -1.0
    h = host_colmajor(a, n, 0.0)
    desca = pkg.DESC9(1, CTX, n, n, nb, nb, 0, 0, n)
    info = ctypes.c_int(0)
    L.dlaf_pdpotrf(b"L", n, h.ctypes.data, 1, 1, desca, ctypes.byref(info))
    assert info.value == 301


def test_make_dlaf_descriptor_is_one_based_like_the_reference(pkg, capi):
    """A caller ported from the reference: make_dlaf_descriptor(n, n, 1, 1, desca) (ScaLAPACK ia = ja = 1, stored
    0-based: src/c_api/utils.cpp:25-33) followed by dlaf_cholesky_factorization_d."""
    L = capi
    n, nb = 384, 128
    a, t = orc.cholesky_setters(n, "L", np.float64)
    h = host_colmajor(a, n, 0.0)
    desca = pkg.DESC9(1, CTX, n, n, nb, nb, 0, 0, n)
    d = L.make_dlaf_descriptor(n, n, 1, 1, desca)
    assert (d.m, d.n, d.mb, d.nb, d.isrc, d.jsrc, d.i, d.j, d.ld) == (n, n, nb, nb, 0, 0, 0, 0, n)
    assert L.dlaf_cholesky_factorization_d(CTX, b"L", h.ctypes.data, d) == 0
    got = np.ascontiguousarray(h[:n, :n].T)
    tol = 4 * (n + 1) * orc.type_error(np.float64)
    assert orc.check_near(np.tril(got), np.tril(t), tol, tol) == 0
    # tile-aligned sub-matrix through the same 1-based arguments
    e = L.make_dlaf_descriptor(n - nb, n - nb, nb + 1, nb + 1, desca)
    assert (e.i, e.j) == (nb, nb)


def test_plan_cache_of_the_c_api_is_bounded(pkg, capi):
    """A caller sweeping over sizes must not accumulate plans and n^2-byte device mirrors (LRU, DLAFB_PLAN_CACHE = 4)."""
    L = capi
    for n in (128, 192, 256, 320, 384, 448, 512):
        a, t = orc.cholesky_setters(n, "L", np.float64)
        h = host_colmajor(a, n, 0.0)
        desca = pkg.DESC9(1, CTX, n, n, 64, 64, 0, 0, n)
        info = ctypes.c_int(-1)
        L.dlaf_pdpotrf(b"L", n, h.ctypes.data, 1, 1, desca, ctypes.byref(info))
        assert info.value == 0
        tol = 4 * (n + 1) * orc.type_error(np.float64)
        assert orc.check_near(np.tril(np.ascontiguousarray(h[:n, :n].T)), np.tril(t), tol, tol) == 0
        assert L.dlafb_capi_cached_plans() <= 4
    # an evicted size is simply rebuilt
    n = 128
    a, t = orc.cholesky_setters(n, "L", np.float64)
    h = host_colmajor(a, n, 0.0)
    desca = pkg.DESC9(1, CTX, n, n, 64, 64, 0, 0, n)
    info = ctypes.c_int(-1)
    L.dlaf_pdpotrf(b"L", n, h.ctypes.data, 1, 1, desca, ctypes.byref(info))
    assert info.value == 0 and L.dlafb_capi_cached_plans() <= 4


def test_cached_plans_of_the_engine(pkg, L):
    """dlafb_cholesky_plan_acquire / release (what dlaf::cholesky_factorization<> uses): the same plan comes back for
    the same key, idle plans beyond the limit go, a grid's plans go with the grid."""
    grid = pkg.Grid(L)
    p1, p2 = ctypes.c_void_p(), ctypes.c_void_p()
    base = L.dlafb_cholesky_plan_cache_size()
    pkg.check(L.dlafb_cholesky_plan_acquire(ctypes.byref(p1), grid.h, b"d", b"L", 256, 64, 0, 0), "acquire")
    pkg.check(L.dlafb_cholesky_plan_release(p1), "release")
    pkg.check(L.dlafb_cholesky_plan_acquire(ctypes.byref(p2), grid.h, b"d", b"L", 256, 64, 0, 0), "acquire")
    assert p1.value == p2.value and L.dlafb_cholesky_plan_cache_size() == base + 1
    a, t = orc.cholesky_setters(256, "L", np.float64)
    A = DevMat(a)
    pkg.check(L.dlafb_cholesky_plan_run(p2, A.ptr, A.ld, None), "run")
    info = ctypes.c_int(-1)
    pkg.check(L.dlafb_cholesky_plan_info(p2, ctypes.byref(info)), "info")
    sync()
    assert info.value == 0
    tol = 4 * 257 * orc.type_error(np.float64)
    assert orc.check_near(np.tril(A.get()), np.tril(t), tol, tol) == 0
    pkg.check(L.dlafb_cholesky_plan_release(p2), "release")
    for n in (64, 128, 192, 320, 384, 448):
        q = ctypes.c_void_p()
        pkg.check(L.dlafb_cholesky_plan_acquire(ctypes.byref(q), grid.h, b"d", b"L", n, 64, 0, 0), "acquire")
        pkg.check(L.dlafb_cholesky_plan_release(q), "release")
        assert L.dlafb_cholesky_plan_cache_size() <= 4
    grid.close()
    assert L.dlafb_cholesky_plan_cache_size() <= base
