"""-m gpu: dlaf::auxiliary::max_norm on the device (include/dlafb_matrix.h: dlafb_max_norm), one GPU. Mirrors
test/unit/auxiliary/test_norm.cpp (NormDistributedTest: max norm of a matrix whose largest element sits at chosen
positions, uplo Lower / General, sizes incl. empty and ragged). Multi-rank: tests/dist_redistribute_check.py.
(The file name sorts last on purpose: written after the round's GPU budget was spent.)"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("x", "sdcz")
def test_max_norm_single_rank(pkg, L, x):
    from gpu_util import NP, DevMat, stream_ptr
    dt = NP[x]
    rng = np.random.default_rng(8)
    grid = pkg.Grid(L)
    for m, nb in [(0, 4), (1, 1), (13, 4), (64, 16), (301, 64)]:
        a = rng.uniform(-1, 1, (m, m)) + (1j * rng.uniform(-1, 1, (m, m)) if np.dtype(dt).kind == "c" else 0)
        a = a.astype(dt)
        A = DevMat(a, ld=max(1, m) + 1)
        want_g = float(np.abs(a).max()) if m else 0.0
        want_l = float(np.abs(np.tril(a)).max()) if m else 0.0
        assert pkg.max_norm(L, grid, x, "G", m, m, nb, nb, (0, 0), A.ptr, A.ld, stream_ptr()) == pytest.approx(want_g, rel=1e-6)
        assert pkg.max_norm(L, grid, x, "L", m, m, nb, nb, (0, 0), A.ptr, A.ld, stream_ptr()) == pytest.approx(want_l, rel=1e-6)
        if m >= 13:
            # the largest element strictly above the diagonal must not count for Lower, and must for General
            b = a.copy()
            b[2, m - 1] = 50.0
            B = DevMat(b, ld=m + 2)
            assert pkg.max_norm(L, grid, x, "L", m, m, nb, nb, (0, 0), B.ptr, B.ld, stream_ptr()) == pytest.approx(want_l, rel=1e-6)
            assert pkg.max_norm(L, grid, x, "G", m, m, nb, nb, (0, 0), B.ptr, B.ld, stream_ptr()) == pytest.approx(50.0, rel=1e-6)
            # ... and one inside a diagonal tile's lower triangle / in the last ragged tile must
            b = a.copy()
            b[m - 1, m - 2] = -70.0
            B = DevMat(b, ld=m)
            assert pkg.max_norm(L, grid, x, "L", m, m, nb, nb, (0, 0), B.ptr, B.ld, stream_ptr()) == pytest.approx(70.0, rel=1e-6)
    # rectangular, General only
    a = rng.uniform(-1, 1, (37, 90)).astype(dt)
    a[30, 77] = 9.0
    A = DevMat(a)
    assert pkg.max_norm(L, grid, x, "G", 37, 90, 8, 16, (0, 0), A.ptr, A.ld, stream_ptr()) == pytest.approx(9.0, rel=1e-6)
    grid.close()


@pytest.mark.parametrize("x", "sdcz")
def test_reference_max_norm_cases(pkg, L, x):
    """NormDistributedTest.MaxNorm_EmptyMatrix / MaxNorm_Correctness (test/unit/auxiliary/mc/test_norm.cpp:49-175) on a
    1 x 1 grid: empty matrices give 0; a 10 x 10 matrix with |a_ij| <= 1, block sizes 3 and 5, whose elements (0, 0),
    (n-1, n-1), (n-1, 0) are set in turn to element(13.13, 26.26) + k * element(26.05, 20.10): the norm must follow."""
    from gpu_util import NP, DevMat, stream_ptr
    dt = NP[x]
    cplx = np.dtype(dt).kind == "c"
    grid = pkg.Grid(L)
    for m, n in [(13, 0), (0, 13), (0, 0)]:
        for nb in (3, 5):
            A = DevMat(np.zeros((m, n), dtype=dt))
            assert pkg.max_norm(L, grid, x, "G", m, n, nb, nb, (0, 0), A.ptr, A.ld, stream_ptr()) == 0.0
    rng = np.random.default_rng(3)
    for nb in (3, 5):
        for uplo in "LG":
            g = rng.uniform(-1, 1, (10, 10)) + (1j * rng.uniform(-1, 1, (10, 10)) if cplx else 0)
            a = ((g + g.conj().T) / (2 * np.sqrt(2))).astype(dt)          # Hermitian, |a_ij| <= 1
            A = DevMat(a)
            norm = pkg.max_norm(L, grid, x, uplo, 10, 10, nb, nb, (0, 0), A.ptr, A.ld, stream_ptr())
            assert 0 <= norm <= 1
            value = complex(13.13, 26.26) if cplx else 13.13
            for idx in [(0, 0), (9, 9), (9, 0)]:
                value = value + (complex(26.05, 20.10) if cplx else 26.05)
                a[idx] = value
                A = DevMat(a)
                got = pkg.max_norm(L, grid, x, uplo, 10, 10, nb, nb, (0, 0), A.ptr, A.ld, stream_ptr())
                assert got == pytest.approx(abs(dt(value)), rel=4 * float(np.finfo(dt).eps))
    grid.close()
