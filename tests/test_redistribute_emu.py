"""CPU tests of the matrix redistribution's product code without a GPU (dla-future_b200/src/redistribute.h: the
rectangle lists and segment bookkeeping; csrc/copy_rects.cuh: the gather / scatter kernel body, run once per thread of
the launch geometry), for every rank of a simulated grid - one host thread per rank, the all-to-all through shared
memory with a size check between every pair, exchange buffers guarded.

Reference semantics: matrix::copy(src, dst, grid), include/dlaf/matrix/copy.h:118-231 - same global matrix, same grid,
the smaller block size divides the larger one in each dimension; checked against the oracle's owner lists
(oracle/dlaf_oracle.py: local_block) - bit exact, this is byte movement. The reference's own test of this function
(test/unit/matrix/test_matrix.cpp:1671-1730, MatrixTest.CopyReshuffling / GPUCopyReshuffling: three size / tile
configurations on the grids of its 6-rank environment, different source ranks for src and dst, element (i, j) = 1000 i + j,
CHECK_MATRIX_EQ) is replayed verbatim by test_reference_reshuffling_cases. GPU: tests/test_zz_redistribute_gpu.py.
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
        (("src", "redistribute.h"), ("csrc", "copy_rects.cuh"), ("src", "schedule.h"), ("src", "gen_to_std_dist.h"),
         ("src", "triangular_solver_dist.h"), ("csrc", "common.cuh"))]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        subprocess.check_call(["make", "emu"], cwd=ROOT)
    lib = ctypes.CDLL(so)
    for x in TYPES:
        f = getattr(lib, f"emu_redistribute_{x}")
        f.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_char, ctypes.c_longlong, ctypes.c_longlong] + [ctypes.c_int] * 8 + [
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong),
            ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_longlong)]
        f.restype = ctypes.c_int
    return lib


def linear(grid, ordering, pr, pc):
    return pc * grid[0] + pr if ordering == "C" else pr * grid[1] + pc


def redistribute(emu, x, grid, ordering, a, src_tile, src_rank, dst_tile, dst_rank, pad=0):
    """Scatters `a` in the source layout, runs all ranks, returns {rank: (got, want)} in the destination layout."""
    dt = TYPES[x]
    P, Q = grid
    m, n = a.shape
    nr = P * Q
    hs, hd, want = [None] * nr, [None] * nr, [None] * nr
    for pr in range(P):
        for pc in range(Q):
            r = linear(grid, ordering, pr, pc)
            ls, _, _ = orc.local_block(a, src_tile[0], src_tile[1], grid, (pr, pc), src_rank)
            ld_, _, _ = orc.local_block(a, dst_tile[0], dst_tile[1], grid, (pr, pc), dst_rank)
            h = np.full((max(1, ls.shape[1]), max(1, ls.shape[0]) + pad), dt(-7.7))
            h[:ls.shape[1], :ls.shape[0]] = ls.T
            hs[r] = h
            hd[r] = np.full((max(1, ld_.shape[1]), max(1, ld_.shape[0]) + pad + 1), dt(-5.5))
            want[r] = ld_
    hs0 = [h.copy() for h in hs]
    ps = (ctypes.c_void_p * nr)(*[h.ctypes.data for h in hs])
    pd = (ctypes.c_void_p * nr)(*[h.ctypes.data for h in hd])
    lds = (ctypes.c_longlong * nr)(*[h.shape[1] for h in hs])
    ldd = (ctypes.c_longlong * nr)(*[h.shape[1] for h in hd])
    r = getattr(emu, f"emu_redistribute_{x}")(P, Q, ordering.encode(), m, n, src_tile[0], src_tile[1], src_rank[0],
                                              src_rank[1], dst_tile[0], dst_tile[1], dst_rank[0], dst_rank[1], ps, lds,
                                              pd, ldd)
    assert r == 0, r
    out = {}
    for k in range(nr):
        assert np.array_equal(hs[k], hs0[k])                                  # the source is read-only
        rows, cols = want[k].shape
        assert np.all(hd[k][:, rows:] == dt(-5.5))                            # padding rows untouched
        assert np.all(hd[k][cols:, :] == dt(-5.5))
        out[k] = (np.ascontiguousarray(hd[k][:cols, :rows].T), want[k])
    return out


CASES = [  # (m, n, src tile, dst tile): in each dimension the smaller block size divides the larger one
    (37, 29, (4, 3), (8, 6)), (37, 29, (8, 6), (4, 3)), (64, 64, (4, 4), (16, 16)), (50, 41, (5, 41), (10, 1)),
    (33, 70, (3, 10), (3, 10)), (16, 16, (16, 16), (2, 2)), (1, 1, (2, 2), (4, 4)), (45, 7, (6, 2), (2, 4)),
]


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("grid,ordering", [((1, 1), "R"), ((2, 1), "R"), ((1, 3), "C"), ((2, 3), "R"), ((3, 2), "C"),
                                           ((2, 4), "R"), ((4, 2), "C")])
def test_destination_layout_is_bit_exact(emu, x, grid, ordering):
    dt = TYPES[x]
    rng = np.random.default_rng(grid[0] * 7 + grid[1])
    for m, n, st, dtile in CASES:
        a = rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if np.dtype(dt).kind == "c" else 0)
        a = a.astype(dt)
        for src_rank, dst_rank in [((0, 0), (0, 0)), ((grid[0] - 1, min(1, grid[1] - 1)), (0, 0)),
                                   ((0, grid[1] - 1), (grid[0] // 2, grid[1] // 2))]:
            out = redistribute(emu, x, grid, ordering, a, st, src_rank, dtile, dst_rank, pad=(m + n) % 3)
            for k, (got, want) in out.items():
                assert np.array_equal(got, want), (m, n, st, dtile, src_rank, dst_rank, k)


def test_round_trip_and_empty(emu):
    rng = np.random.default_rng(5)
    a = rng.standard_normal((53, 47))
    grid = (2, 3)
    there = redistribute(emu, "d", grid, "R", a, (4, 4), (1, 2), (8, 16), (0, 1))
    # gather the destination layout and send it back
    b = np.zeros_like(a)
    for pr in range(2):
        for pc in range(3):
            _, ri, ci = orc.local_block(a, 8, 16, grid, (pr, pc), (0, 1))
            got, _ = there[linear(grid, "R", pr, pc)]
            if ri.size and ci.size:
                b[np.ix_(ri, ci)] = got
    assert np.array_equal(a, b)
    back = redistribute(emu, "d", grid, "R", b, (8, 16), (0, 1), (4, 4), (1, 2))
    for got, want in back.values():
        assert np.array_equal(got, want)
    # empty matrices are a no-op
    for m, n in [(0, 0), (0, 5), (7, 0)]:
        out = redistribute(emu, "d", grid, "C", np.zeros((m, n)), (2, 2), (0, 0), (4, 4), (1, 1))
        assert all(got.size == 0 for got, _ in out.values())


def test_block_sizes_that_do_not_divide_are_rejected(emu):
    a = np.zeros((10, 10))
    with pytest.raises(AssertionError):
        redistribute(emu, "d", (2, 2), "R", a, (4, 4), (0, 0), (6, 4), (0, 0))


REF_RESHUFFLING = [((10, 10), (3, 3), (3, 3)), ((10, 5), (5, 10), (10, 2)), ((26, 13), (10, 3), (5, 6))]  # test_matrix.cpp:1676-1680


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("grid,ordering", [((3, 2), "R"), ((2, 3), "C"), ((3, 1), "R"), ((1, 2), "R"), ((1, 1), "R")])
def test_reference_reshuffling_cases(emu, x, grid, ordering):
    """MatrixTest.CopyReshuffling: src origin (max(0, P-1), max(0, Q-1)), dst origin (min(P-1, ceil(P/2)),
    min(Q-1, ceil(Q/2))), values 1000 i + j, exact comparison (test_matrix.cpp:1683-1712)."""
    dt = TYPES[x]
    P, Q = grid
    src_origin = (max(0, P - 1), max(0, Q - 1))
    dst_origin = (min(P - 1, -(-P // 2)), min(Q - 1, -(-Q // 2)))
    for (m, n), st, dtile in REF_RESHUFFLING:
        i = np.arange(m)[:, None]
        j = np.arange(n)[None, :]
        a = (i * 1000 + j).astype(dt)
        out = redistribute(emu, x, grid, ordering, a, st, src_origin, dtile, dst_origin)
        for k, (got, want) in out.items():
            assert np.array_equal(got, want), ((m, n), st, dtile, k)
