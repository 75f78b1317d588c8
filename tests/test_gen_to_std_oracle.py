"""Pins the generalized-to-standard oracle (oracle/dlaf_oracle.py: gen_to_std_setters, gen_to_std_local) on the
reference's own known-answer test:

  test/unit/eigensolver/test_gen_to_std.cpp:53-83   sizes (m, mb), uplo L/U, four element types,
                                                    CHECK_MATRIX_NEAR(res_a, mat_a, 0, 10 (m+1) error)
  test/include/dlaf_test/matrix/util_generic_lapack.h:70-149   getGenToStdElementSetters(itype = 1, -2, 1.5, .95)
"""
import numpy as np
import pytest
import scipy.linalg as sla

import dlaf_oracle as orc

TYPES = [np.float32, np.float64, np.complex64, np.complex128]


@pytest.mark.parametrize("dt", TYPES)
@pytest.mark.parametrize("uplo", "LU")
def test_local_matches_reference_known_answer(dt, uplo):
    for m, mb in orc.GEN_TO_STD_TEST_SIZES:
        t, a, res = orc.gen_to_std_setters(m, uplo, dt)
        t0 = t.copy()
        orc.gen_to_std_local(uplo, a, t, mb)
        tol = 10 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(a, res, tol, 0.0) == 0, (m, mb)  # the -9.9 of the other triangle must survive too
        assert np.array_equal(t, t0)


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("uplo", "LU")
def test_setters_are_consistent(dt, uplo):
    """The generator itself: RES = inv(T) A inv(T)^H (L) / inv(T)^H A inv(T) (U) on the referenced triangles."""
    n = 12
    t, a, res = orc.gen_to_std_setters(n, uplo, dt)
    tri = np.tril if uplo == "L" else np.triu
    herm = lambda x: tri(x) + tri(x, -1 if uplo == "L" else 1).conj().T
    tt = tri(t)
    want = sla.solve(tt, sla.solve(tt, herm(a)).conj().T) if uplo == "L" else \
        sla.solve(tt.conj().T, sla.solve(tt.conj().T, herm(a)).conj().T)
    assert np.allclose(tri(want.conj().T), tri(res), rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("uplo", "LU")
def test_local_random_against_dense(dt, uplo):
    rng = np.random.default_rng(11)
    n, nb = 45, 8
    cplx = np.dtype(dt).kind == "c"
    g = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a_full = (g + g.conj().T).astype(dt)
    h = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    b_full = (h @ h.conj().T + n * np.eye(n)).astype(dt)
    lfac = np.linalg.cholesky(b_full)
    tri = np.tril if uplo == "L" else np.triu
    fac = lfac if uplo == "L" else lfac.conj().T
    a = np.asfortranarray(np.where(tri(np.ones((n, n))) > 0, a_full, -9.9).astype(dt))
    f = np.asfortranarray(np.where(tri(np.ones((n, n))) > 0, fac, -9.9).astype(dt))
    got = orc.gen_to_std_local(uplo, a, f, nb)
    want = sla.solve(lfac, sla.solve(lfac, a_full).conj().T).conj().T
    assert np.allclose(tri(got), tri(want), rtol=1e-10, atol=1e-10)
    other = ~(tri(np.ones((n, n))) > 0)
    assert np.all(got[other] == dt(-9.9))
