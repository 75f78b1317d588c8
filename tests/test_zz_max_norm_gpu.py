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
