"""CPU tests of the DISTRIBUTED triangular solver's product code without a GPU: tests/emu/solver_emu.cu runs the
algorithm template (dla-future_b200/src/triangular_solver_dist.h) for every rank of a simulated P x Q grid - one host
thread per rank, broadcasts and reductions through shared memory with NCCL's semantics (equal counts on every member or
the run fails, non-root reduce buffers poisoned), workspaces NaN-filled, contractions checked for aliasing - over the
product's index algebra (src/schedule.h) and kernel bodies (csrc/solver_kernels.cuh).

Mirrors test/unit/solver/test_triangular.cpp:106-175 (CorrectnessDistributed): the analytical system for side x uplo x
op x diag, the reference's sizes, four element types, the grids of its 6-rank environment with source rank
(max(0, P-1), min(1, Q-1)), CHECK_MATRIX_NEAR with 20 (m+1) error. The GPU run of the same code is
tests/test_zz_triangular_dist_gpu.py -> tests/dist_trsolve_check.py.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import scipy.linalg as sla

import dlaf_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TYPES = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
GRIDS = [(3, 2), (2, 3), (3, 1), (1, 2), (1, 1)]  # test/include/dlaf_test/comm_grids/grids_6_ranks.h


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "emu", "_build", "libsolver_emu.so")
    deps = [os.path.join(ROOT, "tests", "emu", "solver_emu.cu")] + [
        os.path.join(ROOT, "dla-future_b200", d, f) for d, f in
        (("src", "triangular_solver_dist.h"), ("src", "schedule.h"), ("src", "gen_to_std.h"),
         ("src", "triangular_solver.h"), ("csrc", "solver_kernels.cuh"), ("csrc", "common.cuh"))]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    lib = ctypes.CDLL(so)
    for x in TYPES:
        f = getattr(lib, f"emu_triangular_solver_dist_{x}")
        f.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char] * 4 + [ctypes.c_int] * 4 + [
            ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong),
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong)]
        f.restype = ctypes.c_int
    return lib


def colmajor(a, ld, fill):
    m, n = a.shape
    h = np.full((max(n, 1), ld), fill, dtype=a.dtype)
    h[:n, :m] = a.T
    return h


def run(emu, x, grid, src, side, uplo, op, diag, alpha, a, b, mb, nb, pad=0):
    """Scatters a and b, runs all the ranks, gathers X."""
    dt = TYPES[x]
    P, Q = grid
    m, n = b.shape
    ta = mb if side == "L" else nb
    nr = P * Q
    ha, hb, meta = [], [], []
    for pr in range(P):
        for pc in range(Q):
            la, _, _ = orc.local_block(a, ta, ta, grid, (pr, pc), src)
            lb, ri, ci = orc.local_block(b, mb, nb, grid, (pr, pc), src)
            ha.append(colmajor(la, max(1, la.shape[0]) + pad, dt(-7.7)))
            hb.append(colmajor(lb, max(1, lb.shape[0]) + pad, dt(-5.5)))
            meta.append((ri, ci, lb.shape))
    ha0 = [h.copy() for h in ha]
    pa = (ctypes.c_void_p * nr)(*[h.ctypes.data for h in ha])
    pb = (ctypes.c_void_p * nr)(*[h.ctypes.data for h in hb])
    lda = (ctypes.c_longlong * nr)(*[h.shape[1] for h in ha])
    ldb = (ctypes.c_longlong * nr)(*[h.shape[1] for h in hb])
    al = np.array([alpha], dtype=dt)
    r = getattr(emu, f"emu_triangular_solver_dist_{x}")(P, Q, src[0], src[1], side.encode(), uplo.encode(), op.encode(),
                                                        diag.encode(), m, n, mb, nb, al.ctypes.data, pa, lda, pb, ldb)
    assert r == 0, r
    out = np.zeros((m, n), dtype=dt)
    for h, h0, hbk, (ri, ci, shp) in zip(ha, ha0, hb, meta):
        assert np.array_equal(h, h0, equal_nan=True)  # A is read-only
        assert np.all(hbk[:, shp[0]:] == dt(-5.5))      # padding rows untouched
        if ri.size and ci.size:
            out[np.ix_(ri, ci)] = hbk[:shp[1], :shp[0]].T
    return out


def src_of(grid):
    return (max(0, grid[0] - 1), min(1, grid[1] - 1))


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("grid", GRIDS)
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
def test_reference_known_answer(emu, x, grid, side, uplo, op):
    dt = TYPES[x]
    alpha = orc.triangular_test_alpha(dt)
    for diag in "NU":
        for m, n, mb, nb in orc.TRIANGULAR_TEST_SIZES:
            a, b, want = orc.triangular_system(side, uplo, op, diag, alpha, m, n, dt)
            got = run(emu, x, grid, src_of(grid), side, uplo, op, diag, alpha, a, b, mb, nb, pad=(m + n) % 3)
            tol = 20 * (m + 1) * orc.type_error(dt)
            assert orc.check_near(got, want, tol, tol) == 0, (diag, m, n, mb, nb)


@pytest.mark.parametrize("x", "dz")
@pytest.mark.parametrize("grid,src", [((2, 2), (0, 0)), ((2, 4), (1, 3)), ((4, 2), (3, 0)), ((1, 4), (0, 2)), ((3, 1), (1, 0))])
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
def test_random_against_lapack(emu, x, grid, src, side, uplo, op):
    """Random systems with the unreferenced triangle of A poisoned with NaN, ragged last tiles, more tiles than ranks."""
    dt = TYPES[x]
    rng = np.random.default_rng(17)
    m, n, mb, nb = 52, 43, 8, 6
    na = m if side == "L" else n
    cplx = np.dtype(dt).kind == "c"
    a = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if cplx else 0)
    a = (a / np.sqrt(na) + 2 * np.eye(na) * np.sign(rng.standard_normal(na))).astype(dt)
    b = (rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if cplx else 0)).astype(dt)
    tri = np.tril(a) if uplo == "L" else np.triu(a)
    unref = np.triu(np.ones(a.shape, bool), 1) if uplo == "L" else np.tril(np.ones(a.shape, bool), -1)
    a[unref] = np.nan
    alpha = orc.triangular_test_alpha(dt)
    opa = tri if op == "N" else tri.T if op == "T" else tri.conj().T
    want = sla.solve(opa, alpha * b) if side == "L" else sla.solve(opa.T, (alpha * b).T).T
    got = run(emu, x, grid, src, side, uplo, op, "N", alpha, a, b, mb, nb, pad=1)
    assert np.abs(got - want).max() <= 1e-11 * np.abs(want).max() * max(m, n)


MUTATIONS = {
    # a non-owner of block k adds its partial sum to whatever its buffer held instead of overwriting it
    "reduce_accumulates_into_garbage": ("triangular_solver_dist.h", "ldar, own_col ? R(1) : R(0), w.xb, ldx)", "ldar, R(1), w.xb, ldx)"),
    # the solved block travels down the grid column from the wrong rank row
    "bcast_from_wrong_root": ("triangular_solver_dist.h", "ops.bcast_col(pr_k, w.xb, (size_t)ldx * wk)",
                              "ops.bcast_col((pr_k + 1) % d.P, w.xb, (size_t)ldx * wk)"),
    # the update forgets that the panel starts at the first row beyond... the diagonal tile, not at the local origin
    "panel_offset_dropped": ("triangular_solver_dist.h", "R(-1), w.pan + (rem_begin - r_begin), ldap, w.xb, ldx,", "R(-1), w.pan, ldap, w.xb, ldx,"),
    # the column panel of the distributed reduction is sent without swapping its halves
    "g2s_column_panel_not_swapped": ("gen_to_std_dist.h", "ops.lacpy('G', wj, nb, xl + rloc, ldxr, slot, ldk)",
                                     "ops.lacpy('G', wj, nb, xp + rloc, ldxr, slot, ldk)"),
}


@pytest.mark.parametrize("name", sorted(MUTATIONS))
def test_simulated_grid_catches_seeded_faults(emu, name):
    """The simulated grid is only worth something if it fails when the distributed logic is wrong: seed a fault into a
    copy of the algorithm header, rebuild the emulation against it, and expect an error code or a wrong answer."""
    fname, old, new = MUTATIONS[name]
    src = open(os.path.join(ROOT, "dla-future_b200", "src", fname)).read()
    assert src.count(old) == 1, (name, src.count(old))
    mut_dir = os.path.join(ROOT, "tests", "emu", "_build", "mutd_" + name)
    os.makedirs(mut_dir, exist_ok=True)
    # the two headers include each other with quotes: keep the pair together in the mutated directory
    for h in ("triangular_solver_dist.h", "gen_to_std_dist.h"):
        text = open(os.path.join(ROOT, "dla-future_b200", "src", h)).read()
        with open(os.path.join(mut_dir, h), "w") as f:
            f.write(text.replace(old, new) if h == fname else text)
    so = os.path.join(mut_dir, "libsolver_emu_mut.so")
    subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-O0", "-std=c++17", "-Wno-deprecated-gpu-targets", "-shared",
                           "-Xcompiler", "-fPIC,-pthread", "-I" + mut_dir,
                           "-I" + os.path.join(ROOT, "dla-future_b200", "csrc"),
                           "-I" + os.path.join(ROOT, "dla-future_b200", "src"), "-o", so,
                           os.path.join(ROOT, "tests", "emu", "solver_emu.cu")])
    lib = ctypes.CDLL(so)
    rng = np.random.default_rng(3)
    caught = 0
    if fname == "triangular_solver_dist.h":
        f = lib.emu_triangular_solver_dist_d
        f.argtypes = getattr(emu, "emu_triangular_solver_dist_d").argtypes
        f.restype = ctypes.c_int
        m, n, mb, nb = 37, 29, 5, 4
        for side, uplo, op in (("L", "L", "N"), ("R", "L", "C"), ("L", "U", "C"), ("R", "U", "N")):
            na = m if side == "L" else n
            a = (rng.standard_normal((na, na)) / np.sqrt(na) + 2 * np.eye(na)).astype(np.float64)
            b = rng.standard_normal((m, n))
            tri = np.tril(a) if uplo == "L" else np.triu(a)
            opa = tri if op == "N" else tri.T
            want = sla.solve(opa, b) if side == "L" else sla.solve(opa.T, b.T).T

            class Fake:
                pass
            fake = Fake()
            setattr(fake, "emu_triangular_solver_dist_d", f)
            try:
                got = run(fake, "d", (2, 3), (1, 1), side, uplo, op, "N", 1.0, a, b, mb, nb)
                caught += int(not (np.abs(got - want).max() <= 1e-9))
            except AssertionError:
                caught += 1
    else:
        import test_gen_to_std_dist_emu as g2s
        f = lib.emu_gen_to_std_dist_d
        f.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char] + [ctypes.c_int] * 2 + [
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong),
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong)]
        f.restype = ctypes.c_int

        class Fake:
            pass
        fake = Fake()
        setattr(fake, "emu_gen_to_std_dist_d", f)
        t, a, res = orc.gen_to_std_setters(32, "L", np.float64)
        try:
            got = g2s.run(fake, "d", (2, 3), (1, 1), "L", a, t, 5)
            caught += int(orc.check_near(got, res, 10 * 33 * orc.type_error(np.float64), 0.0) != 0)
        except AssertionError:
            caught += 1
    assert caught > 0
