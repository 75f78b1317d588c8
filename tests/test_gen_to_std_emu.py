"""CPU tests of the generalized-to-standard reduction's product code without a GPU: tests/emu/solver_emu.cu runs the
algorithm template (dla-future_b200/src/gen_to_std.h) over the host executor - same operation sequence, pointer
arithmetic and workspace layout as on the device, workspaces NaN-filled, every contraction checked for aliasing - and
the tile-inverse kernel body (csrc/solver_kernels.cuh). Checked against the reference's analytical case
(test/unit/eigensolver/test_gen_to_std.cpp:53-83, tolerance 10 (m+1) error) and against the oracle's restatement of the
reference tile loops on random problems. The GPU run of the same code is tests/test_zz_gen_to_std_gpu.py.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import dlaf_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TYPES = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so")
    deps = [os.path.join(ROOT, "tests", "emu", "solver_emu.cu")] + [
        os.path.join(ROOT, "dla-future_b200", d, f) for d, f in
        (("src", "gen_to_std.h"), ("src", "triangular_solver.h"), ("csrc", "solver_kernels.cuh"), ("csrc", "common.cuh"))]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    lib = ctypes.CDLL(so)
    for x in TYPES:
        f = getattr(lib, f"emu_gen_to_std_{x}")
        f.argtypes = [ctypes.c_char, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p,
                      ctypes.c_longlong]
        f.restype = ctypes.c_int
    return lib


def colmajor(a, ld, fill):
    m, n = a.shape
    h = np.full((max(n, 1), ld), fill, dtype=a.dtype)
    h[:n, :m] = a.T
    return h


def run(emu, x, uplo, a, f, nb, pad_a=0, pad_f=0):
    dt = TYPES[x]
    n = a.shape[0]
    ha = colmajor(a, max(1, n) + pad_a, dt(-7.7))
    hf = colmajor(f, max(1, n) + pad_f, dt(-5.5))
    hf0 = hf.copy()
    r = getattr(emu, f"emu_gen_to_std_{x}")(uplo.encode(), n, nb, ha.ctypes.data, ha.shape[1], hf.ctypes.data,
                                            hf.shape[1])
    assert r == 0
    assert np.array_equal(hf, hf0, equal_nan=True)  # the factor is never written
    assert np.all(ha[:, n:] == dt(-7.7))     # padding rows untouched
    return np.ascontiguousarray(ha[:n, :n].T)


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
def test_reference_known_answer(emu, x, uplo):
    dt = TYPES[x]
    for m, mb in orc.GEN_TO_STD_TEST_SIZES:
        t, a, res = orc.gen_to_std_setters(m, uplo, dt)
        got = run(emu, x, uplo, a, t, mb, pad_a=m % 3, pad_f=(m + 1) % 3)
        tol = 10 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(got, res, tol, 0.0) == 0, (m, mb)   # includes the -9.9 of the unreferenced triangle


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n,nb", [(45, 8), (64, 16), (50, 64), (33, 32)])
def test_random_against_oracle(emu, x, uplo, n, nb):
    """Random Hermitian A and a random Cholesky factor, unreferenced triangles poisoned with NaN."""
    dt = TYPES[x]
    rng = np.random.default_rng(n * 7 + nb)
    cplx = np.dtype(dt).kind == "c"
    g = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a_full = (g + g.conj().T).astype(dt)
    a_full[np.diag_indices(n)] = a_full.diagonal().real
    h = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    lfac = np.linalg.cholesky((h @ h.conj().T + n * np.eye(n)).astype(dt))
    keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
    fac = lfac if uplo == "L" else lfac.conj().T
    a = np.where(keep, a_full, np.nan).astype(dt)
    f = np.where(keep, fac, np.nan).astype(dt)
    want = orc.gen_to_std_local(uplo, np.asfortranarray(np.where(keep, a_full, 0).astype(dt)),
                                np.asfortranarray(np.where(keep, fac, 0).astype(dt)), nb)
    got = run(emu, x, uplo, a, f, nb, pad_a=1)
    assert np.all(np.isnan(got[~keep]))
    scale = np.abs(want[keep]).max()
    assert np.abs(got[keep] - want[keep]).max() <= 200 * n * np.finfo(dt).eps * scale
