"""Property tests (hypothesis) of the distributed rows on simulated grids: random grid shapes, rank orderings, source
ranks, sizes and tile sizes, far beyond the reference's fixed test tables - empty local blocks, more ranks than tiles,
ragged edges, 1-wide tiles. The product's algorithm templates, index algebra and kernel bodies run on the CPU
(tests/emu/); properties: redistribution is the identity on the global matrix (bit exact) and composes to the identity;
the distributed solve satisfies op(A) X = alpha B; the distributed reduction agrees with the dense formula."""
import numpy as np
import pytest
import scipy.linalg as sla
from hypothesis import HealthCheck, given, settings, strategies as st

import dlaf_oracle as orc
from test_gen_to_std_dist_emu import emu as g2s_emu_fixture  # noqa: F401  (module-scoped fixtures, reused)
from test_gen_to_std_dist_emu import run as g2s_run
from test_redistribute_emu import emu as redist_emu_fixture  # noqa: F401
from test_redistribute_emu import linear, redistribute
from test_triangular_dist_emu import emu as tri_emu_fixture  # noqa: F401
from test_triangular_dist_emu import run as tri_run

import os

N_EXAMPLES = int(os.environ.get("DLAFB_PROPERTY_EXAMPLES", "60"))   # raise for a deeper run
COMMON = dict(deadline=None, max_examples=N_EXAMPLES, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
grids = st.tuples(st.integers(1, 4), st.integers(1, 4)).filter(lambda g: g[0] * g[1] <= 8)


@settings(**COMMON)
@given(grid=grids, ordering=st.sampled_from("RC"), m=st.integers(0, 40), n=st.integers(0, 40),
       base=st.tuples(st.integers(1, 6), st.integers(1, 6)), fs=st.tuples(st.integers(1, 4), st.integers(1, 4)),
       fd=st.tuples(st.integers(1, 4), st.integers(1, 4)), seed=st.integers(0, 10 ** 6))
def test_redistribution_is_the_identity_on_the_global_matrix(redist_emu_fixture, grid, ordering, m, n, base, fs, fd, seed):  # noqa: F811
    rng = np.random.default_rng(seed)
    # in each dimension one block size must divide the other: one of the two factors is 1
    pick = rng.integers(0, 2, size=2)
    src_tile = (base[0] * (fs[0] if pick[0] else 1), base[1] * (fs[1] if pick[1] else 1))
    dst_tile = (base[0] * (1 if pick[0] else fd[0]), base[1] * (1 if pick[1] else fd[1]))
    sr = (int(rng.integers(0, grid[0])), int(rng.integers(0, grid[1])))
    dr = (int(rng.integers(0, grid[0])), int(rng.integers(0, grid[1])))
    a = rng.standard_normal((m, n))
    out = redistribute(redist_emu_fixture, "d", grid, ordering, a, src_tile, sr, dst_tile, dr, pad=seed % 3)
    for got, want in out.values():
        assert np.array_equal(got, want)
    # gather and check the global matrix
    b = np.full((m, n), np.nan)
    for pr in range(grid[0]):
        for pc in range(grid[1]):
            _, ri, ci = orc.local_block(a, dst_tile[0], dst_tile[1], grid, (pr, pc), dr)
            if ri.size and ci.size:
                b[np.ix_(ri, ci)] = out[linear(grid, ordering, pr, pc)][0]
    assert np.array_equal(a, b)


@settings(**COMMON)
@given(grid=grids, side=st.sampled_from("LR"), uplo=st.sampled_from("LU"), op=st.sampled_from("NTC"),
       diag=st.sampled_from("NU"), m=st.integers(1, 30), n=st.integers(1, 30), mb=st.integers(1, 9), nb=st.integers(1, 9),
       x=st.sampled_from("dz"), seed=st.integers(0, 10 ** 6))
def test_distributed_solve_satisfies_the_equation(tri_emu_fixture, grid, side, uplo, op, diag, m, n, mb, nb, x, seed):  # noqa: F811
    dt = {"d": np.float64, "z": np.complex128}[x]
    rng = np.random.default_rng(seed)
    na = m if side == "L" else n
    cplx = x == "z"
    a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if cplx else 0)
    a = (a / max(1.0, np.sqrt(na)) + 2 * np.eye(na) * np.sign(rng.standard_normal(na))).astype(dt)
    b = (rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cplx else 0)).astype(dt)
    tri = np.tril(a) if uplo == "L" else np.triu(a)
    if diag == "U":
        np.fill_diagonal(tri, 1)
    unref = np.triu(np.ones(a.shape, bool), 1) if uplo == "L" else np.tril(np.ones(a.shape, bool), -1)
    if diag == "U":
        unref |= np.eye(na, dtype=bool)
    a[unref] = np.nan
    alpha = orc.triangular_test_alpha(dt)
    src = (int(rng.integers(0, grid[0])), int(rng.integers(0, grid[1])))
    got = tri_run(tri_emu_fixture, x, grid, src, side, uplo, op, diag, alpha, a, b, mb, nb, pad=seed % 3)
    opa = tri if op == "N" else tri.T if op == "T" else tri.conj().T
    lhs = opa @ got if side == "L" else got @ opa
    scale = np.abs(opa).max() * max(1.0, np.abs(got).max()) * max(m, n)
    assert np.abs(lhs - alpha * b).max() <= 1e-12 * scale * np.linalg.cond(opa)


@settings(**{**COMMON, "max_examples": max(1, 2 * N_EXAMPLES // 3)})
@given(grid=grids, uplo=st.sampled_from("LU"), n=st.integers(1, 36), nb=st.integers(1, 9), x=st.sampled_from("dz"),
       seed=st.integers(0, 10 ** 6))
def test_distributed_reduction_matches_the_dense_formula(g2s_emu_fixture, grid, uplo, n, nb, x, seed):  # noqa: F811
    dt = {"d": np.float64, "z": np.complex128}[x]
    rng = np.random.default_rng(seed)
    cplx = x == "z"
    g = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a_full = (g + g.conj().T).astype(dt)
    a_full[np.diag_indices(n)] = a_full.diagonal().real
    h = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    lfac = np.linalg.cholesky((h @ h.conj().T + n * np.eye(n)).astype(dt))
    keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
    fac = lfac if uplo == "L" else lfac.conj().T
    src = (int(rng.integers(0, grid[0])), int(rng.integers(0, grid[1])))
    got = g2s_run(g2s_emu_fixture, x, grid, src, uplo, np.where(keep, a_full, np.nan).astype(dt),
                  np.where(keep, fac, np.nan).astype(dt), nb, pad=seed % 2)
    want = sla.solve(lfac, sla.solve(lfac, a_full).conj().T).conj().T     # L^-1 A L^-H
    assert np.all(np.isnan(got[~keep]))
    assert np.abs(got[keep] - want[keep]).max() <= 1e-11 * n * max(1.0, np.abs(want).max())


# ---- the one-GPU templates (single lane and the two-lane look-ahead variants with race detection) --------------------
from test_gen_to_std_emu import emu as g2s_local_emu_fixture  # noqa: E402,F401
from test_gen_to_std_emu import run as g2s_local_run  # noqa: E402
from test_triangular_emu import emu as tri_local_emu_fixture  # noqa: E402,F401
from test_triangular_emu import run as tri_local_run  # noqa: E402


@settings(**COMMON)
@given(side=st.sampled_from("LR"), uplo=st.sampled_from("LU"), op=st.sampled_from("NTC"), diag=st.sampled_from("NU"),
       m=st.integers(1, 40), n=st.integers(1, 40), nb=st.integers(1, 11), x=st.sampled_from("sdcz"),
       lookahead=st.sampled_from([None, 0, 1]), seed=st.integers(0, 10 ** 6))
def test_local_solve_satisfies_the_equation(tri_local_emu_fixture, side, uplo, op, diag, m, n, nb, x, lookahead, seed):  # noqa: F811
    dt = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}[x]
    rng = np.random.default_rng(seed)
    na = m if side == "L" else n
    cplx = np.dtype(dt).kind == "c"
    a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if cplx else 0)
    a = (a / max(1.0, np.sqrt(na)) + 2 * np.eye(na) * np.sign(rng.standard_normal(na))).astype(dt)
    b = (rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cplx else 0)).astype(dt)
    tri = np.tril(a) if uplo == "L" else np.triu(a)
    if diag == "U":
        np.fill_diagonal(tri, 1)
    unref = np.triu(np.ones(a.shape, bool), 1) if uplo == "L" else np.tril(np.ones(a.shape, bool), -1)
    if diag == "U":
        unref |= np.eye(na, dtype=bool)
    a[unref] = np.nan
    alpha = orc.triangular_test_alpha(dt)
    got = tri_local_run(tri_local_emu_fixture, x, side, uplo, op, diag, alpha, a, b, nb, pad_a=seed % 2, pad_b=seed % 3,
                        lookahead=lookahead)
    wide = np.complex128 if cplx else np.float64
    opa = (tri if op == "N" else tri.T if op == "T" else tri.conj().T).astype(wide)
    lhs = opa @ got.astype(wide) if side == "L" else got.astype(wide) @ opa
    scale = np.abs(opa).max() * max(1.0, np.abs(got).max()) * max(m, n)
    assert np.abs(lhs - wide(alpha) * b.astype(wide)).max() <= 20 * np.finfo(dt).eps * scale * np.linalg.cond(opa)


@settings(**COMMON)
@given(uplo=st.sampled_from("LU"), n=st.integers(1, 48), nb=st.integers(1, 11), x=st.sampled_from("dz"),
       lookahead=st.sampled_from([None, 0, 1]), seed=st.integers(0, 10 ** 6))
def test_local_reduction_matches_the_dense_formula(g2s_local_emu_fixture, uplo, n, nb, x, lookahead, seed):  # noqa: F811
    dt = {"d": np.float64, "z": np.complex128}[x]
    rng = np.random.default_rng(seed)
    cplx = x == "z"
    g = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a_full = (g + g.conj().T).astype(dt)
    a_full[np.diag_indices(n)] = a_full.diagonal().real
    h = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    lfac = np.linalg.cholesky((h @ h.conj().T + n * np.eye(n)).astype(dt))
    keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
    fac = lfac if uplo == "L" else lfac.conj().T
    got = g2s_local_run(g2s_local_emu_fixture, x, uplo, np.where(keep, a_full, np.nan).astype(dt),
                        np.where(keep, fac, np.nan).astype(dt), nb, pad_a=seed % 2, lookahead=lookahead)
    want = sla.solve(lfac, sla.solve(lfac, a_full).conj().T).conj().T
    assert np.all(np.isnan(got[~keep]))
    assert np.abs(got[keep] - want[keep]).max() <= 1e-11 * n * max(1.0, np.abs(want).max())
    if cplx:
        assert np.all(got[np.diag_indices(n)].imag == 0)   # the reduced matrix keeps a real diagonal, as hegst does
