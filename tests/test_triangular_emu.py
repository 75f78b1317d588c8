"""CPU tests of the triangular solver's product code without a GPU: tests/emu/solver_emu.cu runs the algorithm
template (dla-future_b200/src/triangular_solver.h) and the kernel bodies (csrc/solver_kernels.cuh) on the host, one
call per thread of the launch geometry. Checked against the reference's analytical system
(test/unit/solver/test_triangular.cpp:57-106, tolerance 40 (m+1) error) and against the oracle on random systems.
The GPU run of the same code is tests/test_triangular_gpu.py.
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
    src = os.path.join(ROOT, "tests", "emu", "solver_emu.cu")
    deps = [src] + [os.path.join(ROOT, "dla-future_b200", d, f) for d, f in
                    (("src", "triangular_solver.h"), ("csrc", "solver_kernels.cuh"), ("csrc", "common.cuh"))]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    lib = ctypes.CDLL(so)
    for x in TYPES:
        f = getattr(lib, f"emu_triangular_solver_{x}")
        f.argtypes = [ctypes.c_char] * 4 + [ctypes.c_int] * 3 + [ctypes.c_void_p] * 2 + [ctypes.c_longlong,
                                                                                       ctypes.c_void_p, ctypes.c_longlong]
        f.restype = ctypes.c_int
        h = getattr(lib, f"emu_triangular_solver_la_{x}")
        h.argtypes = f.argtypes + [ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        h.restype = ctypes.c_int
        g = getattr(lib, f"emu_tri_inverse_{x}")
        g.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p]
        g.restype = ctypes.c_int
    return lib


def colmajor(a, ld, fill):
    """Column-major storage with leading dimension ld as a C array h[j, i] = a[i, j]."""
    m, n = a.shape
    h = np.full((max(n, 1), ld), fill, dtype=a.dtype)
    h[:n, :m] = a.T
    return h


def run(emu, x, side, uplo, op, diag, alpha, a, b, nb, pad_a=0, pad_b=0, lookahead=None):
    """lookahead = None: single lane; 0 / 1: the two-lane variant replayed with that lane running whenever it can."""
    dt = TYPES[x]
    m, n = b.shape
    na = a.shape[0]
    ha = colmajor(a, max(1, na) + pad_a, dt(-7.7))
    hb = colmajor(b, max(1, m) + pad_b, dt(-5.5))
    al = np.array([alpha], dtype=dt)
    args = (side.encode(), uplo.encode(), op.encode(), diag.encode(), m, n, nb, al.ctypes.data, ha.ctypes.data,
            ha.shape[1], hb.ctypes.data, hb.shape[1])
    if lookahead is None:
        r = getattr(emu, f"emu_triangular_solver_{x}")(*args)
    else:
        races = ctypes.c_int(-1)
        r = getattr(emu, f"emu_triangular_solver_la_{x}")(*args, lookahead, ctypes.byref(races))
        assert races.value == 0, f"{races.value} unordered operation pairs with overlapping operands"
    assert r == 0
    assert np.all(hb[:, m:] == dt(-5.5))  # padding rows untouched
    return np.ascontiguousarray(hb[:n, :m].T)


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
@pytest.mark.parametrize("diag", "NU")
def test_reference_known_answer(emu, x, side, uplo, op, diag):
    dt = TYPES[x]
    alpha = orc.triangular_test_alpha(dt)
    for m, n, mb, nb in orc.TRIANGULAR_TEST_SIZES:
        a, b, want = orc.triangular_system(side, uplo, op, diag, alpha, m, n, dt)
        got = run(emu, x, side, uplo, op, diag, alpha, a, b, mb if side == "L" else nb, pad_a=m % 3, pad_b=n % 2)
        tol = 40 * (m + 1) * orc.type_error(dt)
        assert orc.check_near(got, want, tol, tol) == 0, (m, n, mb, nb)


def random_system(rng, dt, side, m, n):
    na = m if side == "L" else n
    cplx = np.dtype(dt).kind == "c"
    a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if cplx else 0)
    a = a + 2 * np.sqrt(na) * np.eye(na) * np.sign(rng.standard_normal(na))
    b = rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cplx else 0)
    return np.asfortranarray(a.astype(dt)), np.asfortranarray(b.astype(dt))


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
def test_random_against_oracle(emu, x, side, uplo, op):
    """Poisoned unreferenced triangle (NaN), alpha = 1 and alpha != 1, ragged last tile, padded leading dimensions."""
    dt = TYPES[x]
    rng = np.random.default_rng(11)
    for (m, n, nb, alpha) in [(70, 45, 16, 1.0), (45, 70, 32, orc.triangular_test_alpha(dt)), (33, 33, 33, 0.5)]:
        a, b = random_system(rng, dt, side, m, n)
        for diag in "NU":
            want = orc.triangular_solver_local(side, uplo, op, diag, alpha, a.copy(order="F"), b.copy(order="F"), nb, nb)
            ap = a.copy()
            unref = np.triu(np.ones(a.shape, bool), 1) if uplo == "L" else np.tril(np.ones(a.shape, bool), -1)
            if diag == "U":
                unref |= np.eye(a.shape[0], dtype=bool)
            ap[unref] = np.nan
            got = run(emu, x, side, uplo, op, diag, alpha, ap, b, nb, pad_a=3, pad_b=1)
            scale = np.abs(want).max()
            assert np.abs(got - want).max() <= 1e-11 * scale * max(m, n), (m, n, nb, diag)


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("upper", [0, 1])
@pytest.mark.parametrize("unit", [0, 1])
def test_tile_inverse(emu, x, upper, unit):
    """tri_inverse on several diagonal tiles (more columns than threads of a CTA, ragged last tile): W T = I."""
    dt = TYPES[x]
    rng = np.random.default_rng(3)
    n, nb, ld = 700, 300, 703
    cplx = np.dtype(dt).kind == "c"
    t = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    t = (t / np.sqrt(n) + 2 * np.eye(n)).astype(dt)
    poisoned = t.copy()
    unref = np.tril(np.ones((n, n), bool), -1) if upper else np.triu(np.ones((n, n), bool), 1)
    if unit:
        unref |= np.eye(n, dtype=bool)
    poisoned[unref] = np.nan
    ht = colmajor(poisoned, ld, dt(np.nan))
    nt = -(-n // nb)
    w = np.full((nt, nb, nb), np.nan, dtype=dt)  # slot k: w[k, j, i] = W_k[i, j]
    assert getattr(emu, f"emu_tri_inverse_{x}")(upper, unit, n, nb, ht.ctypes.data, ld, w.ctypes.data) == 0
    eps = np.finfo(dt).eps
    for k in range(nt):
        k0 = k * nb
        o = min(nb, n - k0)
        tk = t[k0:k0 + o, k0:k0 + o]
        tk = np.triu(tk) if upper else np.tril(tk)
        if unit:
            np.fill_diagonal(tk, 1)
        wk = w[k].T[:o, :o]
        assert np.all(np.isfinite(wk))
        assert np.all((np.tril(wk, -1) if upper else np.triu(wk, 1)) == 0)
        assert np.abs(wk @ tk - np.eye(o)).max() <= 50 * o * eps


# ---------------------------------------------------------------------------------- look-ahead on two lanes
@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
@pytest.mark.parametrize("prefer", [0, 1])
def test_lookahead_reference_known_answer(emu, x, side, uplo, op, prefer):
    """The two-lane variant under both extreme interleavings: same answers, no unordered conflicting operations."""
    dt = TYPES[x]
    alpha = orc.triangular_test_alpha(dt)
    for diag in "NU":
        for m, n, mb, nb in orc.TRIANGULAR_TEST_SIZES:
            a, b, want = orc.triangular_system(side, uplo, op, diag, alpha, m, n, dt)
            got = run(emu, x, side, uplo, op, diag, alpha, a, b, mb if side == "L" else nb, pad_a=m % 3, pad_b=n % 2,
                      lookahead=prefer)
            tol = 40 * (m + 1) * orc.type_error(dt)
            assert orc.check_near(got, want, tol, tol) == 0, (m, n, mb, nb, diag)


@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NC")
@pytest.mark.parametrize("prefer", [0, 1])
def test_lookahead_random_many_steps(emu, side, uplo, op, prefer):
    """Enough steps for every event parity to be reused several times; ragged last tile."""
    dt = np.float64
    rng = np.random.default_rng(17)
    m, n, nb = (61, 23, 7) if side == "L" else (23, 61, 7)
    a, b = random_system(rng, dt, side, m, n)
    want = orc.triangular_solver_local(side, uplo, op, "N", 0.75, a.copy(order="F"), b.copy(order="F"), nb, nb)
    ap = a.copy()
    ap[np.triu(np.ones(a.shape, bool), 1) if uplo == "L" else np.tril(np.ones(a.shape, bool), -1)] = np.nan
    got = run(emu, "d", side, uplo, op, "N", 0.75, ap, b, nb, lookahead=prefer)
    assert np.abs(got - want).max() <= 1e-11 * np.abs(want).max() * max(m, n)


MUTATIONS = {
    # lane 0 updates the next block without waiting for the previous bulk update, which writes that block too
    "no_wait_for_bulk": ("      if (s >= 1 && bulk_issued[(s - 1) & 1] && (e = ops.wait(kEvBulk0 + ((s - 1) & 1))) != 0)\n"
                         "        return e;\n", ""),
    # lane 1 starts the bulk update without waiting for the solve that produces its operand
    "no_wait_for_solve": ("        if ((e = ops.wait(kEvSolve0 + (s & 1))) != 0)\n          return e;\n", ""),
    # single panel buffer: lane 0 overwrites X_s^H while lane 1 still reads it (left-side solves)
    "single_panel": ("      T* pan = w.panel + (size_t)(s & 1) * w.panel_stride;", "      T* pan = w.panel;"),
}


@pytest.mark.parametrize("name", sorted(MUTATIONS))
def test_race_detector_catches_seeded_faults(emu, name):
    """The CPU replay is only worth something if it fails when the event logic is wrong: seed three faults into a
    copy of the algorithm header, rebuild the emulation against it, and expect races and/or wrong answers."""
    old, new = MUTATIONS[name]
    src = open(os.path.join(ROOT, "dla-future_b200", "src", "triangular_solver.h")).read()
    assert old in src
    mut_dir = os.path.join(ROOT, "tests", "emu", "_build", "mut_" + name)
    os.makedirs(mut_dir, exist_ok=True)
    with open(os.path.join(mut_dir, "triangular_solver.h"), "w") as f:
        f.write(src.replace(old, new))
    so = os.path.join(mut_dir, "libsolver_emu_mut.so")
    subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-O0", "-std=c++17", "-Wno-deprecated-gpu-targets", "-shared",
                           "-Xcompiler", "-fPIC", "-I" + mut_dir, "-I" + os.path.join(ROOT, "dla-future_b200", "csrc"),
                           "-I" + os.path.join(ROOT, "dla-future_b200", "src"), "-o", so,
                           os.path.join(ROOT, "tests", "emu", "solver_emu.cu")])
    lib = ctypes.CDLL(so)
    f = lib.emu_triangular_solver_la_d
    f.argtypes = [ctypes.c_char] * 4 + [ctypes.c_int] * 3 + [ctypes.c_void_p] * 2 + [
        ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    f.restype = ctypes.c_int
    rng = np.random.default_rng(1)
    m, n, nb = 61, 23, 7
    a, b = random_system(rng, np.float64, "L", m, n)
    want = orc.triangular_solver_local("L", "L", "N", "N", 1.0, a.copy(order="F"), b.copy(order="F"), nb, nb)
    total_races, wrong = 0, 0
    for prefer in (0, 1):
        ha, hb = colmajor(a, m, 0.0), colmajor(b, m, 0.0)
        al = np.array([1.0])
        races = ctypes.c_int(-1)
        r = f(b"L", b"L", b"N", b"N", m, n, nb, al.ctypes.data, ha.ctypes.data, m, hb.ctypes.data, m, prefer,
              ctypes.byref(races))
        assert r == 0
        total_races += races.value
        err = np.abs(hb[:n, :m].T - want).max()
        wrong += int(not (err <= 1e-9))
    assert total_races > 0
    assert wrong > 0
