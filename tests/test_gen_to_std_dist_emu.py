"""CPU tests of the DISTRIBUTED generalized-to-standard reduction's product code without a GPU: tests/emu/solver_emu.cu
runs the algorithm template (dla-future_b200/src/gen_to_std_dist.h, which ends in the trapezoid mode of
src/triangular_solver_dist.h) for every rank of a simulated P x Q grid - one host thread per rank, collectives through
shared memory with NCCL's semantics, workspaces NaN-filled, contractions checked for aliasing.

Mirrors test/unit/eigensolver/test_gen_to_std.cpp:85-125 (CorrectnessDistributed): the analytical case for uplo L/U, the
reference's sizes, four element types, the grids of its 6-rank environment with source rank (max(0, P-1), min(1, Q-1)),
CHECK_MATRIX_NEAR(res_a, mat_a, 0, 10 (m+1) error) on the whole matrix (the unreferenced triangle must keep its -9.9)
and the factor unchanged. The GPU run of the same code: tests/test_zz_gen_to_std_gpu.py -> tests/dist_g2s_check.py.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import dlaf_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TYPES = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
GRIDS = [(3, 2), (2, 3), (3, 1), (1, 2), (1, 1)]


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so")
    deps = [os.path.join(ROOT, "tests", "emu", "solver_emu.cu")] + [
        os.path.join(ROOT, "dla-future_b200", d, f) for d, f in
        (("src", "gen_to_std_dist.h"), ("src", "triangular_solver_dist.h"), ("src", "schedule.h"), ("src", "gen_to_std.h"),
         ("src", "triangular_solver.h"), ("csrc", "solver_kernels.cuh"), ("csrc", "common.cuh"))]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    lib = ctypes.CDLL(so)
    for x in TYPES:
        f = getattr(lib, f"emu_gen_to_std_dist_{x}")
        f.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char] + [ctypes.c_int] * 2 + [
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong),
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong)]
        f.restype = ctypes.c_int
    return lib


def colmajor(a, ld, fill):
    m, n = a.shape
    h = np.full((max(n, 1), ld), fill, dtype=a.dtype)
    h[:n, :m] = a.T
    return h


def run(emu, x, grid, src, uplo, a, f, nb, pad=0):
    dt = TYPES[x]
    P, Q = grid
    n = a.shape[0]
    nr = P * Q
    ha, hf, meta = [], [], []
    for pr in range(P):
        for pc in range(Q):
            la, ri, ci = orc.local_block(a, nb, nb, grid, (pr, pc), src)
            lf, _, _ = orc.local_block(f, nb, nb, grid, (pr, pc), src)
            ha.append(colmajor(la, max(1, la.shape[0]) + pad, dt(-7.7)))
            hf.append(colmajor(lf, max(1, lf.shape[0]) + pad + 1, dt(-5.5)))
            meta.append((ri, ci, la.shape))
    hf0 = [h.copy() for h in hf]
    pa = (ctypes.c_void_p * nr)(*[h.ctypes.data for h in ha])
    pf = (ctypes.c_void_p * nr)(*[h.ctypes.data for h in hf])
    lda = (ctypes.c_longlong * nr)(*[h.shape[1] for h in ha])
    ldf = (ctypes.c_longlong * nr)(*[h.shape[1] for h in hf])
    r = getattr(emu, f"emu_gen_to_std_dist_{x}")(P, Q, src[0], src[1], uplo.encode(), n, nb, pa, lda, pf, ldf)
    assert r == 0, r
    out = np.zeros((n, n), dtype=dt)
    for h, hfk, h0, (ri, ci, shp) in zip(ha, hf, hf0, meta):
        assert np.array_equal(hfk, h0, equal_nan=True)   # the factor is never written
        assert np.all(h[:, shp[0]:] == dt(-7.7))          # padding rows untouched
        if ri.size and ci.size:
            out[np.ix_(ri, ci)] = h[:shp[1], :shp[0]].T
    return out


def src_of(grid):
    return (max(0, grid[0] - 1), min(1, grid[1] - 1))


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("grid", GRIDS)
@pytest.mark.parametrize("uplo", "LU")
def test_reference_known_answer(emu, x, grid, uplo):
    dt = TYPES[x]
    for m, mb in orc.GEN_TO_STD_TEST_SIZES:
        t, a, res = orc.gen_to_std_setters(m, uplo, dt)
        got = run(emu, x, grid, src_of(grid), uplo, a, t, mb, pad=m % 3)
        tol = 10 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(got, res, tol, 0.0) == 0, (m, mb)


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("grid,src", [((2, 2), (0, 0)), ((2, 4), (1, 3)), ((4, 2), (3, 0)), ((1, 4), (0, 2)), ((3, 1), (1, 0)),
                                      ((2, 3), (1, 1))])
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n,nb", [(45, 8), (64, 16), (41, 5)])
def test_random_against_oracle(emu, x, grid, src, uplo, n, nb):
    dt = TYPES[x]
    rng = np.random.default_rng(n + nb)
    cplx = np.dtype(dt).kind == "c"
    g = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a_full = (g + g.conj().T).astype(dt)
    a_full[np.diag_indices(n)] = a_full.diagonal().real
    h = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    lfac = np.linalg.cholesky((h @ h.conj().T + n * np.eye(n)).astype(dt))
    keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
    fac = lfac if uplo == "L" else lfac.conj().T
    want = orc.gen_to_std_local(uplo, np.asfortranarray(np.where(keep, a_full, 0).astype(dt)),
                                np.asfortranarray(np.where(keep, fac, 0).astype(dt)), nb)
    got = run(emu, x, grid, src, uplo, np.where(keep, a_full, np.nan).astype(dt), np.where(keep, fac, np.nan).astype(dt),
              nb, pad=1)
    assert np.all(np.isnan(got[~keep]))
    scale = np.abs(want[keep]).max()
    assert np.abs(got[keep] - want[keep]).max() <= 200 * n * np.finfo(dt).eps * scale
