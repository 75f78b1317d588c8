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


def test_profiling_timeline_and_exchange_mode(pkg, L):
    """The profiling hooks bench.py relies on: per-step marks are monotonic on each stream and the bulk-update
    timing adds up; a single-rank plan reports the 'local' exchange mode."""
    n, nb = 2048, 256
    a = orc.gen_hpd(n, nb, np.float64)
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
    assert plan.exchange_mode() == "local"
    plan.set_profiling(True)
    A = DevMat(a, ld=n)
    plan.run(A.ptr, A.ld, stream_ptr())
    assert plan.info() == 0
    ms, flops, launches = plan.profile()
    tl = plan.timeline()
    nt = n // nb
    assert len(tl) == nt and launches == nt - 2 and ms > 0 and flops > 0
    for k in range(nt - 1):
        hp = tl[k][:6]
        assert all(x >= 0 for x in hp) and all(hp[i] <= hp[i + 1] + 1e-3 for i in range(5))
        if k < nt - 2:
            assert 0 <= tl[k][6] <= tl[k][7]
    assert all(x == -1.0 for x in tl[nt - 1])
    plan.set_profiling(False)
    want = a.copy(order="F")
    assert orc.cholesky_local("L", want, nb) == 0
    assert np.abs(np.tril(A.get()) - np.tril(want)).max() < 1e-10
    plan.close()
    grid.close()


@pytest.mark.parametrize("n,nb", [(1536, 256), (4096, 256), (2560, 128)])
@pytest.mark.parametrize("uplo", "LU")
def test_run_host_moves_only_the_referenced_triangle(pkg, L, uplo, n, nb):
    """dlafb_cholesky_plan_run_host: device mirror supplied by the caller, pinned or pageable host block; the
    unreferenced triangle of the host block stays untouched and the mirror's one is never read."""
    import torch
    a = orc.gen_hpd(n, nb, np.float64)
    want = a.copy(order="F")
    assert orc.cholesky_local(uplo, want, nb) == 0
    ref = np.tril(np.ones((n, n), bool)) if uplo == "L" else np.triu(np.ones((n, n), bool))
    host = np.where(ref, a, -9.9).T.copy()           # column-major n x n, ld = n
    dev = torch.full((n * n + 64,), float("nan"), dtype=torch.float64, device="cuda")   # poisoned mirror
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, "d", uplo, n, nb)
    pkg.check(L.dlafb_cholesky_plan_run_host(plan.h, host.ctypes.data, n, dev.data_ptr(), n, stream_ptr()), "run_host")
    assert plan.info() == 0
    sync()
    got = host.T
    assert np.abs(np.where(ref, got - want, 0)).max() < 1e-10
    assert np.all(got[~ref] == -9.9)
    # same plan again (the upload-group events and streams are reused), pinned host block this time
    pinned = torch.from_numpy(np.where(ref, a, -9.9).T.copy()).pin_memory()
    pkg.check(L.dlafb_cholesky_plan_run_host(plan.h, pinned.data_ptr(), n, dev.data_ptr(), n, stream_ptr()), "run_host")
    assert plan.info() == 0
    sync()
    assert np.array_equal(pinned.numpy().T, got)
    plan.close()
    grid.close()


@pytest.mark.parametrize("x,n,nb", [("d", 1536, 256), ("d", 1100, 128), ("d", 650, 128), ("z", 768, 128), ("s", 640, 64),
                                    ("c", 300, 128), ("d", 34, 13)])
def test_device_residual_matches_numpy(pkg, L, x, n, nb):
    """dlafb_cholesky_plan_residual (the miniapp's acceptance check, on the device) against numpy: exact max|A| over
    the lower triangle, max|A - L L^H| equal up to rounding, and sensitivity to a wrong factor."""
    dt = NP[x]
    a = orc.gen_hpd(n, nb, dt)
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, x, "L", n, nb)
    ld = -(-n // 64) * 64
    A = DevMat(a, ld=ld)
    plan.run(A.ptr, A.ld, stream_ptr())
    assert plan.info() == 0
    got = A.get()
    A0 = DevMat(a, ld=ld)
    amax, dmax = plan.residual(A0.ptr, A0.ld, A.ptr, A.ld, stream_ptr())
    low = np.tril(np.ones((n, n), bool))
    lf = np.tril(got).astype(np.complex128 if x in "cz" else np.float64)
    want = np.abs((a.astype(lf.dtype) - lf @ lf.conj().T)[low]).max()
    eps = np.finfo(dt).eps
    assert amax == pytest.approx(np.abs(a[low]).max(), rel=1e-6)
    assert dmax <= n * eps * amax                      # the reference's acceptance threshold
    assert dmax == pytest.approx(want, rel=0.5, abs=4 * eps * amax)
    # the residual block itself: R = A - L L^H on the referenced triangle
    r = A0.get()
    assert np.abs(r[low] - (a.astype(lf.dtype) - lf @ lf.conj().T)[low]).max() <= 16 * eps * amax
    assert np.all(A0.pad() == dt(-7.7))
    # a wrong factor is seen
    bad = got.copy()
    bad[n // 2, n // 3] += 1
    B, A1 = DevMat(bad, ld=ld), DevMat(a, ld=ld)
    _, dbad = plan.residual(A1.ptr, A1.ld, B.ptr, B.ld, stream_ptr())
    assert dbad > 0.5
    plan.close()
    grid.close()
