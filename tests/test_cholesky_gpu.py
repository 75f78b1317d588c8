"""-m gpu parity tests of the factorization through the engine C ABI (dlafb_cholesky_plan_*), single GPU.

Cases and tolerance follow test/unit/factorization/test_cholesky.cpp:54-139 (local part): analytical input and
factor from getCholeskySetters, element-wise CHECK_MATRIX_NEAR with 4(m+1)*err, both uplo, four element types,
default (padded, ld multiple of 64) and tight layouts. Larger seeded cases use the reference's random HPD
generator (oracle port of util_matrix.h:499-501) and are checked against the oracle's factor and by the
miniapp's residual (miniapp_cholesky.cpp:408-442).
"""
import numpy as np
import pytest

import dlaf_oracle as orc
from gpu_util import NP, DevMat, stream_ptr, sync

pytestmark = pytest.mark.gpu

REF_SIZES = [(0, 2), (5, 8), (34, 34), (4, 3), (16, 10), (34, 13), (32, 5)]


def factor(pkg, L, x, uplo, a, nb, ld=None):
    n = a.shape[0]
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, x, uplo, n, nb)
    assert (plan.local_rows, plan.local_cols) == (n, n)
    A = DevMat(a, ld=ld if ld is not None else max(1, -(-n // 64) * 64))
    plan.run(A.ptr, A.ld, stream_ptr())
    info = plan.info()
    sync()
    out = A.get()
    assert np.all(A.pad() == a.dtype.type(-7.7))
    plan.close()
    grid.close()
    return out, info


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("m,mb", REF_SIZES)
def test_local_analytic(pkg, L, x, uplo, m, mb):
    dt = NP[x]
    a, t = orc.cholesky_setters(m, uplo, dt)
    got, info = factor(pkg, L, x, uplo, a, mb)
    assert info == 0
    tol = 4 * (m + 1) * orc.type_error(dt)
    assert orc.check_near(got, t, tol, tol) == 0


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("m,mb,tight", [(256, 128, False), (384, 128, True), (300, 128, False), (512, 64, False),
                                        (200, 256, False), (1024, 256, False), (650, 128, True)])
def test_local_analytic_fast_path_shapes(pkg, L, x, uplo, m, mb, tight):
    """Block sizes that take the TMA + DMMA kernels for double (nb % 128 == 0), incl. ragged last tiles and an
    odd leading dimension (mirror path)."""
    dt = NP[x]
    a, t = orc.cholesky_setters(m, uplo, dt)
    got, info = factor(pkg, L, x, uplo, a, mb, ld=(m + 1 if tight else None))
    assert info == 0
    tol = 4 * (m + 1) * orc.type_error(dt)
    assert orc.check_near(got, t, tol, tol) == 0


@pytest.mark.parametrize("x,n,nb", [("d", 2048, 512), ("d", 1536, 256), ("d", 1100, 128), ("z", 768, 128),
                                    ("s", 1024, 256), ("c", 640, 128)])
def test_local_random_hpd_vs_oracle(pkg, L, x, n, nb):
    dt = NP[x]
    a = orc.gen_hpd(n, nb, dt)
    ref = a.copy(order="F")
    assert orc.cholesky_local("L", ref, nb) == 0
    got, info = factor(pkg, L, x, "L", a, nb)
    assert info == 0
    eps = np.finfo(dt).eps
    # same algorithm, different summation order: compare factors relative to the diagonal scale sqrt(2n)
    scale = np.sqrt(2.0 * n)
    assert np.abs(np.tril(got) - np.tril(ref)).max() <= 50 * n * eps * scale / np.sqrt(n)
    assert np.array_equal(np.triu(got, 1), np.triu(a, 1))   # unreferenced triangle untouched
    ratio, thresh = orc.residual_max(a, got, "L")
    assert ratio <= thresh                                   # miniapp: no "Warning" line
    assert orc.residual_inf(a, got, "L") <= 1.0


@pytest.mark.parametrize("x", "dz")
def test_not_positive_definite_reports_global_column(pkg, L, x):
    dt = NP[x]
    n, nb = 640, 128
    a = orc.gen_hpd(n, nb, dt)
    a[300, 300] = -5.0
    got, info = factor(pkg, L, x, "L", a, nb)
    assert info == 301


def test_repeated_runs_same_plan(pkg, L):
    n, nb = 1024, 256
    a = orc.gen_hpd(n, nb, np.float64)
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
    outs = []
    for _ in range(3):
        A = DevMat(a, ld=1024)
        plan.run(A.ptr, A.ld, stream_ptr())
        assert plan.info() == 0
        outs.append(A.get())
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[1], outs[2])   # deterministic
    plan.close()
    grid.close()
