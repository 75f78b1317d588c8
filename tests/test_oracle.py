"""Pins the CPU oracle (oracle/dlaf_oracle.py, oracle/chol_oracle.cpp) on the reference's own known-answer tests.

  algorithm, local + distributed   test/unit/factorization/test_cholesky.cpp:54-139
      sizes (m, mb), both uplo, four element types, grids of the 6-rank fixture
      (test/include/dlaf_test/comm_grids/grids_6_ranks.h:26-88), source rank (max(0,P-1), min(1,Q-1)),
      element-wise CHECK_MATRIX_NEAR with 4(m+1)*TypeUtilities<T>::error
  tile potrf                       test/unit/test_lapack_tile.cpp:142-205, test_lapack_tile/test_potrf.h:33-77
  index algebra                    test/unit/matrix/test_util_distribution.cpp (spot values restated below)
  input generator / residual       include/dlaf/util_matrix.h:499-501, miniapp/miniapp_cholesky.cpp:408-442
"""
import numpy as np
import pytest

import dlaf_oracle as orc

TYPES = [np.float32, np.float64, np.complex64, np.complex128]
REF_SIZES = [(0, 2), (5, 8), (34, 34), (4, 3), (16, 10), (34, 13), (32, 5)]
GRIDS = [(3, 2), (2, 3), (3, 1), (1, 2), (1, 1)]


@pytest.mark.parametrize("dt", TYPES)
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("m,mb", REF_SIZES)
def test_local_matches_reference_known_answer(dt, uplo, m, mb):
    a, t = orc.cholesky_setters(m, uplo, dt)
    assert orc.cholesky_local(uplo, a, mb) == 0
    tol = 4 * (m + 1) * orc.type_error(dt)
    assert orc.check_near(a, t, tol, tol) == 0


@pytest.mark.parametrize("dt", TYPES)
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("grid", GRIDS)
@pytest.mark.parametrize("m,mb", REF_SIZES)
def test_distributed_matches_reference_known_answer(dt, uplo, grid, m, mb):
    P, Q = grid
    src = (max(0, P - 1), min(1, Q - 1))
    a, t = orc.cholesky_setters(m, uplo, dt)
    local = orc.scatter(a, mb, grid, src)
    assert orc.cholesky_distributed(uplo, local, m, mb, grid, src) == 0
    got = orc.gather(local, m, mb, dt)
    tol = 4 * (m + 1) * orc.type_error(dt)
    assert orc.check_near(got, t, tol, tol) == 0


@pytest.mark.parametrize("dt", TYPES)
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n", [0, 1, 11, 12, 17, 128])
def test_tile_potrf_known_answer(dt, uplo, n):
    a, t = orc.cholesky_setters(n, uplo, dt)
    assert orc.tile_potrf(uplo, a) == 0
    tol = 4 * (n + 1) * orc.type_error(dt)
    assert orc.check_near(a, t, tol, tol) == 0


@pytest.mark.parametrize("dt", TYPES)
def test_tile_potrf_not_positive_definite(dt):
    a = np.zeros((4, 4), dtype=dt, order="F")
    assert orc.tile_potrf("L", a) == 1  # test_potrf.h:59-77


def test_distribution_known_answers():
    # util_distribution.h:85-175, values as in test/unit/matrix/test_util_distribution.cpp style tables
    d = orc.Distribution(100, 10, grid=(3, 2), rank=(1, 0), src=(2, 1))
    assert d.nt == 10
    assert [d.rank_global_tile(g, 0) for g in range(6)] == [2, 0, 1, 2, 0, 1]
    assert [d.rank_global_tile(g, 1) for g in range(4)] == [1, 0, 1, 0]
    assert [d.local_tile_from_global(g, 0) for g in (0, 2, 3, 5, 9)] == [0, 0, 1, 1, 3]
    # rank row 1 with source 2 owns global row tiles 2, 5, 8
    assert [d.global_tile_from_local(l, 0) for l in range(3)] == [2, 5, 8]
    assert [d.next_local_tile_from_global(g, 0) for g in range(11)] == [0, 0, 0, 1, 1, 1, 2, 2, 2, 3, 3]
    assert d.local_nr_tiles(0) == 3 and d.local_nr_tiles(1) == 5
    assert d.local_size(0) == 30 and d.local_size(1) == 50
    e = orc.Distribution(95, 10, grid=(3, 2), rank=(2, 1), src=(0, 0))
    assert e.nt == 10 and e.tile_size(9) == 5
    assert e.local_size(0) == 30 and e.local_size(1) == 45


@pytest.mark.parametrize("dt", TYPES)
def test_random_hpd_generator_properties(dt):
    n, nb = 70, 16
    a = orc.gen_hpd(n, nb, dt)
    assert np.array_equal(a, a.conj().T)
    off = a - np.diag(np.diag(a))
    assert np.abs(off).max() <= 1.0 + 1e-6
    assert np.all(np.abs(np.diag(a).real - 2 * n) <= 1.0) and np.all(np.diag(a).imag == 0)
    # the stream is per tile: a different tile size gives a different (still valid) matrix, the same one repeats
    assert np.array_equal(a, orc.gen_hpd(n, nb, dt))
    b = a.copy(order="F")
    assert orc.cholesky_local("L", b, nb) == 0
    ratio, thresh = orc.residual_max(a, b, "L")
    assert ratio <= thresh
    assert orc.residual_inf(a, b, "L") <= 1.0


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
def test_cpp_oracle_agrees_with_python_oracle(dt):
    n, nb = 200, 32
    a = orc.gen_hpd(n, nb, dt)
    p = a.copy(order="F")
    c = a.copy(order="F")
    assert orc.cholesky_local("L", p, nb) == 0
    assert orc.cholesky_local_threaded("L", c, nb, task_threads=4) == 0
    assert np.abs(np.tril(p) - np.tril(c)).max() <= 1e-10
    assert np.array_equal(np.triu(c, 1), np.triu(a, 1))


def test_not_positive_definite_info_is_global_column():
    n, nb = 40, 8
    a = orc.gen_hpd(n, nb, np.float64)
    a[19, 19] = -3.0
    assert orc.cholesky_local("L", a.copy(order="F"), nb) == 20
    local = orc.scatter(a, nb, (2, 3))
    assert orc.cholesky_distributed("L", local, n, nb, (2, 3)) == 20
