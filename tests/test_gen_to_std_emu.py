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
        g = getattr(lib, f"emu_gen_to_std_la_{x}")
        g.argtypes = f.argtypes + [ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        g.restype = ctypes.c_int
    return lib


def colmajor(a, ld, fill):
    m, n = a.shape
    h = np.full((max(n, 1), ld), fill, dtype=a.dtype)
    h[:n, :m] = a.T
    return h


def run(emu, x, uplo, a, f, nb, pad_a=0, pad_f=0, lookahead=None):
    """lookahead = None: single lane; 0 / 1: the two-lane variant replayed with that lane running whenever it can."""
    dt = TYPES[x]
    n = a.shape[0]
    ha = colmajor(a, max(1, n) + pad_a, dt(-7.7))
    hf = colmajor(f, max(1, n) + pad_f, dt(-5.5))
    hf0 = hf.copy()
    args = (uplo.encode(), n, nb, ha.ctypes.data, ha.shape[1], hf.ctypes.data, hf.shape[1])
    if lookahead is None:
        r = getattr(emu, f"emu_gen_to_std_{x}")(*args)
    else:
        races = ctypes.c_int(-1)
        r = getattr(emu, f"emu_gen_to_std_la_{x}")(*args, lookahead, ctypes.byref(races))
        assert races.value == 0, f"{races.value} unordered operation pairs with overlapping operands"
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


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("prefer", [0, 1])
def test_lookahead_reference_known_answer(emu, x, uplo, prefer):
    """Two-lane variant (DLAFB_G2S_LOOKAHEAD=1): replayed with either lane running ahead whenever the events allow it;
    no pair of operations that the events leave unordered may touch overlapping operands."""
    dt = TYPES[x]
    for m, mb in orc.GEN_TO_STD_TEST_SIZES:
        t, a, res = orc.gen_to_std_setters(m, uplo, dt)
        got = run(emu, x, uplo, a, t, mb, pad_a=m % 3, pad_f=(m + 1) % 3, lookahead=prefer)
        tol = 10 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(got, res, tol, 0.0) == 0, (m, mb)


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n,nb", [(45, 8), (64, 16), (61, 7)])
def test_lookahead_matches_single_lane(emu, x, uplo, n, nb):
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
    a = np.where(keep, a_full, np.nan).astype(dt)
    f = np.where(keep, fac, np.nan).astype(dt)
    base = run(emu, x, uplo, a, f, nb)
    for prefer in (0, 1):
        got = run(emu, x, uplo, a, f, nb, lookahead=prefer)
        assert np.all(np.isnan(got[~keep]))
        # the same operations on the same data: the split of the trailing update changes no element's arithmetic
        assert np.array_equal(got[keep], base[keep])


MUTATIONS = {
    # lane 0 updates the next block column without waiting for the bulk update of the previous step
    "no_wait_for_bulk": ("    if (k >= 1 && bulk_issued[(k - 1) & 1] && (e = ops.wait(kG2sBulk0 + ((k - 1) & 1))) != 0)\n      return e;\n",
                         ""),
    # lane 1 starts the bulk update without waiting for the panels of the step
    "no_wait_for_panel": ("      if ((e = ops.wait(kG2sPanel0 + (k & 1))) != 0)\n        return e;\n", ""),
    # one panel buffer for all the steps
    "single_panel": ("    T* pa = w.pa + (size_t)(k & 1) * w.panel_stride;\n    T* pb = w.pb + (size_t)(k & 1) * w.panel_stride;",
                     "    T* pa = w.pa;\n    T* pb = w.pb;"),
}


@pytest.mark.parametrize("name", sorted(MUTATIONS))
def test_race_detector_catches_seeded_faults(emu, name):
    old, new = MUTATIONS[name]
    src = open(os.path.join(ROOT, "dla-future_b200", "src", "gen_to_std.h")).read()
    assert src.count(old) == 1, name
    mut_dir = os.path.join(ROOT, "tests", "emu", "_build", "mutg_" + name)
    os.makedirs(mut_dir, exist_ok=True)
    with open(os.path.join(mut_dir, "gen_to_std.h"), "w") as f:
        f.write(src.replace(old, new))
    so = os.path.join(mut_dir, "libsolver_emu_mut.so")
    subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-O0", "-std=c++17", "-Wno-deprecated-gpu-targets", "-shared",
                           "-Xcompiler", "-fPIC,-pthread", "-I" + mut_dir,
                           "-I" + os.path.join(ROOT, "dla-future_b200", "csrc"),
                           "-I" + os.path.join(ROOT, "dla-future_b200", "src"), "-o", so,
                           os.path.join(ROOT, "tests", "emu", "solver_emu.cu")])
    lib = ctypes.CDLL(so)
    g = lib.emu_gen_to_std_la_d
    g.argtypes = [ctypes.c_char, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p,
                  ctypes.c_longlong, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    g.restype = ctypes.c_int
    m, mb = 61, 7
    rng = np.random.default_rng(2)
    gm = rng.standard_normal((m, m))
    a = np.tril(gm + gm.T)
    lf = np.tril(rng.standard_normal((m, m)) / np.sqrt(m), -1) + 2 * np.eye(m)
    want = orc.gen_to_std_local("L", np.asfortranarray(a.copy()), np.asfortranarray(lf.copy()), mb)
    total_races, wrong = 0, 0
    for prefer in (0, 1):
        ha, hf = colmajor(a, m, 0.0), colmajor(lf, m, 0.0)
        races = ctypes.c_int(-1)
        r = g(b"L", m, mb, ha.ctypes.data, m, hf.ctypes.data, m, prefer, ctypes.byref(races))
        assert r == 0
        total_races += races.value
        wrong += int(not (np.abs(np.tril(ha.T) - np.tril(want)).max() <= 1e-9))
    assert total_races > 0
    assert wrong > 0
