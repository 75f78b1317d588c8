"""-m gpu tests of the local triangular solver (include/dlafb_solver.h) on one GPU.

Mirrors test/unit/solver/test_triangular.cpp (TriangularSolverTestGPU, CorrectnessLocal): the analytical system of
test/include/dlaf_test/matrix/util_generic_blas.h:235-371 for side x uplo x op x diag, the reference's sizes, four
element types, alpha = element(-1.2, .7), CHECK_MATRIX_NEAR with 40 (m+1) error; then larger random systems (TMA +
DMMA contraction path for d and z) against LAPACK, with the unreferenced triangle of A poisoned.
The same algorithm and kernel bodies run on the CPU in tests/test_triangular_emu.py.
"""
import numpy as np
import pytest
import scipy.linalg as sla

import dlaf_oracle as orc
from gpu_util import NP, DevMat, stream_ptr, sync

pytestmark = pytest.mark.gpu


def solve(pkg, L, x, side, uplo, op, diag, alpha, a, b, nb, pad_a=0, pad_b=0):
    m, n = b.shape
    na = a.shape[0]
    A = DevMat(a, ld=max(1, na) + pad_a)
    B = DevMat(b, ld=max(1, m) + pad_b)
    plan = pkg.TriangularSolverPlan(L, x, side, uplo, op, diag, m, n, nb)
    plan.run(alpha, A.ptr, A.ld, B.ptr, B.ld, stream_ptr())
    sync()
    plan.close()
    assert np.all(B.pad() == NP[x](-7.7))
    assert np.array_equal(A.get(), a, equal_nan=True)  # A is read-only
    return B.get()


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
@pytest.mark.parametrize("diag", "NU")
def test_reference_known_answer(pkg, L, x, side, uplo, op, diag):
    dt = NP[x]
    alpha = orc.triangular_test_alpha(dt)
    for m, n, mb, nb in orc.TRIANGULAR_TEST_SIZES:
        a, b, want = orc.triangular_system(side, uplo, op, diag, alpha, m, n, dt)
        got = solve(pkg, L, x, side, uplo, op, diag, alpha, a, b, mb if side == "L" else nb, pad_a=m % 3, pad_b=n % 2)
        tol = 40 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(got, want, tol, tol) == 0, (m, n, mb, nb)


def random_system(rng, dt, side, uplo, diag, m, n):
    na = m if side == "L" else n
    cplx = np.dtype(dt).kind == "c"
    a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if cplx else 0)
    a = a / np.sqrt(na) + 2 * np.eye(na) * np.sign(rng.standard_normal(na))
    b = rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cplx else 0)
    a = a.astype(dt)
    tri = np.tril(a) if uplo == "L" else np.triu(a)
    if diag == "U":
        np.fill_diagonal(tri, 1)
    unref = np.triu(np.ones(a.shape, bool), 1) if uplo == "L" else np.tril(np.ones(a.shape, bool), -1)
    if diag == "U":
        unref |= np.eye(na, dtype=bool)
    a[unref] = np.nan
    return np.asfortranarray(a), np.asfortranarray(b.astype(dt)), tri


def lapack_solve(side, op, alpha, tri, b):
    opa = tri if op == "N" else tri.T if op == "T" else tri.conj().T
    if side == "L":
        return sla.solve(opa, alpha * b)
    return sla.solve(opa.T, (alpha * b).T).T


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
def test_random_tma_path(pkg, L, x, side, uplo, op):
    """Aligned operands, tile sizes that are multiples of 16: every contraction runs on the TMA + DMMA kernel."""
    dt = NP[x]
    rng = np.random.default_rng(21)
    m, n, nb = (768, 640, 256) if x == "d" else (512, 384, 128)
    for diag, alpha in (("N", orc.triangular_test_alpha(dt)), ("U", 1.0)):
        a, b, tri = random_system(rng, dt, side, uplo, diag, m, n)
        want = lapack_solve(side, op, alpha, tri, b)
        got = solve(pkg, L, x, side, uplo, op, diag, alpha, a, b, nb)
        assert np.abs(got - want).max() <= 1e-11 * np.abs(want).max() * max(m, n)


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("side,uplo,op", [("L", "L", "N"), ("L", "U", "C"), ("R", "L", "C"), ("R", "U", "N"),
                                          ("R", "L", "T"), ("L", "L", "T")])
def test_random_ragged(pkg, L, x, side, uplo, op):
    """Odd sizes and leading dimensions, ragged last tile (generic contraction kernel for every type)."""
    dt = NP[x]
    rng = np.random.default_rng(8)
    m, n, nb = 301, 215, 64
    a, b, tri = random_system(rng, dt, side, uplo, "N", m, n)
    alpha = orc.triangular_test_alpha(dt)
    want = lapack_solve(side, op, alpha, tri.astype(np.complex128 if np.dtype(dt).kind == "c" else np.float64),
                        b.astype(np.complex128 if np.dtype(dt).kind == "c" else np.float64))
    got = solve(pkg, L, x, side, uplo, op, "N", alpha, a, b, nb, pad_a=1, pad_b=3)
    eps = np.finfo(dt).eps
    assert np.abs(got - want).max() <= 200 * eps * np.abs(want).max() * max(m, n)


def test_solves_with_the_cholesky_factor(pkg, L):
    """A x = b through the factorization and two solves (what the reference's users do with the factor)."""
    import torch
    n, nb, nrhs = 1024, 256, 256
    a = orc.gen_hpd(n, nb, np.float64)
    rng = np.random.default_rng(2)
    xs = rng.standard_normal((n, nrhs))
    b = np.asfortranarray(a @ xs)
    A = DevMat(a)
    B = DevMat(b)
    grid = pkg.Grid(L)
    chol = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
    chol.run(A.ptr, A.ld, stream_ptr())
    assert chol.info() == 0
    for op in "NC":
        plan = pkg.TriangularSolverPlan(L, "d", "L", "L", op, "N", n, nrhs, nb)
        plan.run(1.0, A.ptr, A.ld, B.ptr, B.ld, stream_ptr())
        sync()
        plan.close()
    chol.close()
    grid.close()
    got = B.get()
    assert np.abs(got - xs).max() <= 1e-9 * np.abs(xs).max()
    del torch
