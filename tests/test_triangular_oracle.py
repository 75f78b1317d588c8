"""Pins the triangular-solver oracle (oracle/dlaf_oracle.py: triangular_system, triangular_solver_local) on the
reference's own known-answer test:

  test/unit/solver/test_triangular.cpp:57-106   sizes (m, n, mb, nb), side x uplo x op x diag, four element types,
                                                alpha = element(-1.2, .7), CHECK_MATRIX_NEAR with 40 (m+1) error
  test/include/dlaf_test/matrix/util_generic_blas.h:235-371   the analytical system op(A) X = alpha B / X op(A) = alpha B
"""
import numpy as np
import pytest
import scipy.linalg as sla

import dlaf_oracle as orc

TYPES = [np.float32, np.float64, np.complex64, np.complex128]


@pytest.mark.parametrize("dt", TYPES)
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
@pytest.mark.parametrize("diag", "NU")
def test_local_matches_reference_known_answer(dt, side, uplo, op, diag):
    alpha = orc.triangular_test_alpha(dt)
    for m, n, mb, nb in orc.TRIANGULAR_TEST_SIZES:
        a, b, x = orc.triangular_system(side, uplo, op, diag, alpha, m, n, dt)
        orc.triangular_solver_local(side, uplo, op, diag, alpha, a, b, mb, nb)
        tol = 40 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(b, x, tol, tol) == 0, (m, n, mb, nb)


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
def test_analytical_system_is_consistent(dt, side, uplo, op):
    """The generator itself: op(A) X = alpha B with the referenced triangle only (poisoned entries masked)."""
    m, n = 13, 9
    alpha = orc.triangular_test_alpha(dt)
    for diag in "NU":
        a, b, x = orc.triangular_system(side, uplo, op, diag, alpha, m, n, dt)
        tri = np.tril(a) if uplo == "L" else np.triu(a)
        if diag == "U":
            np.fill_diagonal(tri, 1)
        opa = tri if op == "N" else tri.T if op == "T" else tri.conj().T
        lhs = opa @ x if side == "L" else x @ opa
        assert np.allclose(lhs, alpha * b, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
def test_local_random_against_lapack(dt):
    """Random well-conditioned systems (not the analytical one) against scipy's whole-matrix solve."""
    rng = np.random.default_rng(5)
    m, n, mb, nb = 37, 29, 8, 6
    for side in "LR":
        na = m if side == "L" else n
        for uplo in "LU":
            for op in "NTC":
                a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if np.dtype(dt).kind == "c" else 0)
                a = np.asfortranarray((a + 2 * na * np.eye(na)).astype(dt))
                b = rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if np.dtype(dt).kind == "c" else 0)
                b = np.asfortranarray(b.astype(dt))
                alpha = orc.triangular_test_alpha(dt)
                tri = np.tril(a) if uplo == "L" else np.triu(a)
                opa = tri if op == "N" else tri.T if op == "T" else tri.conj().T
                want = sla.solve(opa, alpha * b) if side == "L" else sla.solve(opa.T, (alpha * b).T).T
                got = orc.triangular_solver_local(side, uplo, op, "N", alpha, a, b.copy(order="F"), mb, nb)
                assert np.allclose(got, want, rtol=1e-10, atol=1e-10)
