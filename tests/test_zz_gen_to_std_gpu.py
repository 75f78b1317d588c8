"""-m gpu tests of the generalized-to-standard reduction (include/dlafb_gen_to_std.h) on one GPU.

Mirrors test/unit/eigensolver/test_gen_to_std.cpp (EigensolverGenToStdTestGPU, CorrectnessLocal): the analytical case
of test/include/dlaf_test/matrix/util_generic_lapack.h:70-149 for uplo L/U, the reference's sizes, four element types,
CHECK_MATRIX_NEAR(res_a, mat_a, 0, 10 (m+1) error) - which also demands that the unreferenced triangle of A keeps its
-9.9; then larger random problems (TMA + DMMA contraction path for d and z) against the oracle's restatement of the
reference tile loops, with the unreferenced triangles poisoned with NaN.
The same algorithm template and kernel bodies run on the CPU in tests/test_gen_to_std_emu.py.
(The file name sorts last on purpose: this row was written after the round's GPU budget was spent, so its first run
on a device must not stand in front of the tests that have already passed there.)
"""
import numpy as np
import pytest

import dlaf_oracle as orc
from gpu_util import NP, DevMat, stream_ptr, sync

pytestmark = pytest.mark.gpu


def reduce_(pkg, L, x, uplo, a, f, nb, pad_a=0, pad_f=0):
    n = a.shape[0]
    A = DevMat(a, ld=max(1, n) + pad_a)
    F = DevMat(f, ld=max(1, n) + pad_f)
    plan = pkg.GenToStdPlan(L, x, uplo, n, nb)
    plan.run(A.ptr, A.ld, F.ptr, F.ld, stream_ptr())
    sync()
    plan.close()
    assert np.all(A.pad() == NP[x](-7.7))
    assert np.array_equal(F.get(), f, equal_nan=True)  # the factor is read-only
    return A.get()


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
def test_reference_known_answer(pkg, L, x, uplo):
    dt = NP[x]
    for m, mb in orc.GEN_TO_STD_TEST_SIZES:
        t, a, res = orc.gen_to_std_setters(m, uplo, dt)
        got = reduce_(pkg, L, x, uplo, a, t, mb, pad_a=m % 3, pad_f=(m + 1) % 3)
        tol = 10 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(got, res, tol, 0.0) == 0, (m, mb)


def random_problem(rng, dt, uplo, n):
    cplx = np.dtype(dt).kind == "c"
    g = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a_full = (g + g.conj().T).astype(dt)
    a_full[np.diag_indices(n)] = a_full.diagonal().real
    h = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    lfac = np.linalg.cholesky((h @ h.conj().T / n + 2 * np.eye(n)).astype(dt))
    keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
    fac = lfac if uplo == "L" else lfac.conj().T
    return a_full, fac, keep


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("uplo", "LU")
def test_random_tma_path(pkg, L, x, uplo):
    """Aligned operands, tile sizes that are multiples of 16: every contraction runs on the TMA + DMMA kernel."""
    dt = NP[x]
    rng = np.random.default_rng(31)
    n, nb = (1280, 256) if x == "d" else (640, 128)
    a_full, fac, keep = random_problem(rng, dt, uplo, n)
    want = orc.gen_to_std_local(uplo, np.asfortranarray(np.where(keep, a_full, 0).astype(dt)),
                                np.asfortranarray(np.where(keep, fac, 0).astype(dt)), nb)
    got = reduce_(pkg, L, x, uplo, np.where(keep, a_full, np.nan).astype(dt), np.where(keep, fac, np.nan).astype(dt), nb)
    assert np.all(np.isnan(got[~keep]))
    assert np.abs(got[keep] - want[keep]).max() <= 1e-11 * n * np.abs(want[keep]).max()


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
def test_random_ragged(pkg, L, x, uplo):
    """Odd size and leading dimensions, ragged last tile (generic contraction kernel for every type)."""
    dt = NP[x]
    rng = np.random.default_rng(9)
    n, nb = 301, 64
    a_full, fac, keep = random_problem(rng, dt, uplo, n)
    wide = np.complex128 if np.dtype(dt).kind == "c" else np.float64
    want = orc.gen_to_std_local(uplo, np.asfortranarray(np.where(keep, a_full, 0).astype(wide)),
                                np.asfortranarray(np.where(keep, fac, 0).astype(wide)), nb)
    got = reduce_(pkg, L, x, uplo, np.where(keep, a_full, np.nan).astype(dt), np.where(keep, fac, np.nan).astype(dt), nb,
                  pad_a=1, pad_f=3)
    assert np.all(np.isnan(got[~keep]))
    assert np.abs(got[keep] - want[keep]).max() <= 200 * np.finfo(dt).eps * n * np.abs(want[keep]).max()


def test_generalized_eigenvalues_survive(pkg, L):
    """What the reduction is for: with B = L L^H factored on the device by the Cholesky engine, the eigenvalues of
    inv(L) A inv(L)^H are the generalized eigenvalues of (A, B)."""
    import scipy.linalg as sla
    n, nb = 768, 256
    b = orc.gen_hpd(n, nb, np.float64)
    rng = np.random.default_rng(4)
    g = rng.standard_normal((n, n))
    a_full = g + g.T
    Bd = DevMat(b)
    Ad = DevMat(np.asfortranarray(np.tril(a_full)))
    grid = pkg.Grid(L)
    chol = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
    chol.run(Bd.ptr, Bd.ld, stream_ptr())
    assert chol.info() == 0
    plan = pkg.GenToStdPlan(L, "d", "L", n, nb)
    plan.run(Ad.ptr, Ad.ld, Bd.ptr, Bd.ld, stream_ptr())
    sync()
    plan.close()
    chol.close()
    grid.close()
    c = np.tril(Ad.get())
    c = c + np.tril(c, -1).T
    b_full = np.tril(b) + np.tril(b, -1).T
    want = sla.eigh(a_full, b_full, eigvals_only=True)
    got = np.linalg.eigvalsh(c)
    assert np.abs(got - want).max() <= 1e-10 * np.abs(want).max() * n


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
def test_single_rank_grid_matches_reference_known_answer(pkg, L, x, uplo):
    """A 1 x 1 grid runs the DISTRIBUTED algorithm (row / column panels, per-column-tile contractions, trapezoid solve,
    axes-swapped executor for uplo U) without NCCL."""
    dt = NP[x]
    grid = pkg.Grid(L)
    for m, mb in orc.GEN_TO_STD_TEST_SIZES + [(300, 64)]:
        t, a, res = orc.gen_to_std_setters(m, uplo, dt)
        if m > 100:  # the analytical entries underflow far from the diagonal: compare with the local plan instead
            rng = np.random.default_rng(3)
            a_full, fac, keep = random_problem(rng, dt, uplo, m)
            a = np.where(keep, a_full, np.nan).astype(dt)
            t = np.where(keep, fac, np.nan).astype(dt)
            res = reduce_(pkg, L, x, uplo, a, t, mb)
        A = DevMat(a, ld=max(1, m) + 1)
        F = DevMat(t, ld=max(1, m) + 2)
        plan = pkg.GenToStdDistPlan(L, grid, x, uplo, m, mb)
        plan.run(A.ptr, A.ld, F.ptr, F.ld, stream_ptr())
        sync()
        plan.close()
        got = A.get()
        tol = 10 * (m + 1) * orc.type_error(dt) * (1 if m <= 100 else np.nanmax(np.abs(res)))
        if m > 100:
            inside = ~np.isnan(res)
            assert np.all(np.isnan(got[~inside]))
            assert np.abs(got[inside] - res[inside]).max() <= tol
        else:
            assert orc.check_near(got, res, tol, 0.0) == 0, (m, mb)
        assert np.array_equal(F.get(), t, equal_nan=True)
    grid.close()


@pytest.mark.parametrize("world", [2, 4, 6, 8])
def test_distributed_parity(world):
    import os
    import subprocess
    import sys
    import torch
    env = os.environ.get("DLAFB_DIST_WORLDS")
    worlds = [int(w) for w in env.split(",")] if env else [2, 4, 6, 8]
    if world not in worlds or torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    e = dict(os.environ)
    if world > 2:
        e.setdefault("DLAFB_DIST_QUICK", "1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29800 + world), os.path.join(root, "tests", "dist_g2s_check.py")]
    r = subprocess.run(cmd, cwd=root, capture_output=True, text=True, timeout=2400, env=e)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "total_failures=0" in r.stdout, tail


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("uplo", "LU")
def test_lookahead_variant_matches_single_stream(pkg, L, monkeypatch, x, uplo):
    """DLAFB_G2S_LOOKAHEAD=1 (read when the plan is created): two streams, the chain of step k+1 under the bulk update of
    step k. Same operations on the same data as the single-stream plan (CPU replay with race detection:
    tests/test_gen_to_std_emu.py), so the results must agree to rounding of the split launches."""
    dt = NP[x]
    rng = np.random.default_rng(37)
    n, nb = (1536, 256) if x == "d" else (768, 128)
    a_full, fac, keep = random_problem(rng, dt, uplo, n)
    a = np.where(keep, a_full, np.nan).astype(dt)
    f = np.where(keep, fac, np.nan).astype(dt)
    monkeypatch.setenv("DLAFB_G2S_LOOKAHEAD", "0")
    base = reduce_(pkg, L, x, uplo, a, f, nb)
    monkeypatch.setenv("DLAFB_G2S_LOOKAHEAD", "1")
    for _ in range(3):   # a race would not show every time
        got = reduce_(pkg, L, x, uplo, a, f, nb)
        assert np.all(np.isnan(got[~keep]))
        assert np.abs(got[keep] - base[keep]).max() <= 1e-12 * n * np.abs(base[keep]).max()
