"""CPU tests of the wide float contraction kernel (dla-future_b200/csrc/gemm_simt_wide.cuh, DLAFB_SGEMM_WIDE=1) without a
GPU: tests/emu/simt_wide_emu.cu runs the kernel body with one host thread per CUDA thread of a block, a real barrier and
a real (NaN-initialised) shared-memory slice per block. The result must equal the scalar chain
acc = fmaf(a_ik, b_jk, acc), k = 0, 1, ... bit for bit - the arithmetic of the 64 x 64 kernel that the GPU tests have
verified (csrc/gemm_simt.cu) - in RECT mode (gemm_nt as tile_ops.cu issues it: clipping, lower mask, K not a multiple
of the slice, odd leading dimensions) and in LIST mode (the factorization's update lists: diagonal tiles that take both
operands from the row panel, operands stored as contiguous tiles)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
c_f, c_i, c_ll, c_vp = ctypes.c_float, ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "emu", "_build", "libsimt_wide_emu.so")
    deps = [os.path.join(ROOT, "tests", "emu", "simt_wide_emu.cu")] + [
        os.path.join(ROOT, "dla-future_b200", "csrc", f) for f in ("gemm_simt_wide.cuh", "gemm_simt.cuh", "common.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        subprocess.check_call(["make", "tests/emu/_build/libsimt_wide_emu.so"], cwd=ROOT)
    lib = ctypes.CDLL(so)
    sig = [c_i, c_i, c_i, c_f, c_vp, c_ll, c_vp, c_ll, c_f, c_vp, c_ll, c_i]
    lib.emu_sgemm_wide_rect.argtypes = sig
    lib.emu_sgemm_wide_rect.restype = c_i
    lib.emu_sgemm_scalar.argtypes = sig
    lib.emu_sgemm_scalar.restype = None
    lib.emu_sgemm_wide_list.argtypes = [c_i, c_i, c_i, c_f, c_f, c_vp, c_ll, c_i, c_vp, c_ll, c_i, c_vp, c_ll, c_vp, c_i]
    lib.emu_sgemm_wide_list.restype = c_i
    return lib


def cm(a, ld, fill=np.float32(-7.7)):
    """Column-major storage with leading dimension ld: h[j, i] = a[i, j]."""
    m, n = a.shape
    h = np.full((max(n, 1), ld), fill, dtype=np.float32)
    h[:n, :m] = a.T
    return h


@pytest.mark.parametrize("M,N,K,lower", [(128, 128, 8, 0), (128, 128, 64, 0), (200, 131, 37, 0), (1, 1, 1, 0), (64, 300, 16, 0),
                                         (257, 257, 24, 1), (130, 130, 5, 1), (96, 40, 128, 0)])
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (-1.0, 1.0), (0.75, -0.5)])
def test_rect_mode_is_bit_exact(emu, M, N, K, lower, alpha, beta):
    rng = np.random.default_rng(M * 3 + N + K)
    a = rng.standard_normal((M, K)).astype(np.float32)
    b = rng.standard_normal((N, K)).astype(np.float32)
    c0 = rng.standard_normal((M, N)).astype(np.float32)
    if beta == 0.0:
        c0[:] = np.nan   # beta = 0 must not read C
    ha, hb = cm(a, M + 3), cm(b, N + 1)
    got, want = cm(c0, M + 2), cm(c0, M + 2)
    emu.emu_sgemm_scalar(M, N, K, alpha, ha.ctypes.data, ha.shape[1], hb.ctypes.data, hb.shape[1], beta, want.ctypes.data,
                         want.shape[1], lower)
    r = emu.emu_sgemm_wide_rect(M, N, K, alpha, ha.ctypes.data, ha.shape[1], hb.ctypes.data, hb.shape[1], beta,
                                got.ctypes.data, got.shape[1], lower)
    assert r == 0
    assert np.array_equal(got, want, equal_nan=True)          # includes the padding rows and (lower) the upper triangle
    ref = alpha * (a.astype(np.float64) @ b.astype(np.float64).T) + (beta * c0 if beta != 0 else 0)
    mask = np.tril(np.ones((M, N), bool)) if lower else np.ones((M, N), bool)
    assert np.abs(got[:N, :M].T[mask] - ref[mask]).max() <= 1e-4 * max(1.0, np.abs(ref[mask]).max())


@pytest.mark.parametrize("nb,tiled", [(128, 0), (128, 1), (256, 1), (192, 0), (64, 1)])
def test_list_mode_matches_the_engine_update(emu, nb, tiled):
    """C -= A B^T on a list of nb x nb tiles of a local block, diagonal tiles lower-only with both operands from the row
    panel - the shape of the factorization's trailing update for float."""
    rng = np.random.default_rng(nb + tiled)
    rt, ct = 3, 2                                   # local tile rows / columns
    mloc, nloc = rt * nb - 5, ct * nb - 9           # ragged last tiles
    c0 = rng.standard_normal((mloc, nloc)).astype(np.float32)
    if tiled:
        # panels as contiguous nb x nb tiles (slot i = local tile row / column i), lda = ldb = nb
        pr = rng.standard_normal((rt * nb, nb)).astype(np.float32)
        pc = rng.standard_normal((ct * nb, nb)).astype(np.float32)
        a_mem = np.concatenate([cm(pr[i * nb:(i + 1) * nb], nb).reshape(-1) for i in range(rt)])
        b_mem = np.concatenate([cm(pc[j * nb:(j + 1) * nb], nb).reshape(-1) for j in range(ct)])
        lda = ldb = nb
    else:
        pr = rng.standard_normal((rt * nb, nb)).astype(np.float32)
        pc = pr                                      # local case: both operands are column k of the local block
        a_mem = cm(pr, rt * nb + 2).reshape(-1)
        b_mem = a_mem
        lda = ldb = rt * nb + 2
    # tiles: (0,0) diagonal, (1,0), (2,0), (1,1) diagonal-like flagged tile, (2,1)
    entries = [(0, 0, 1), (nb, 0, 0), (2 * nb, 0, 0), (nb, nb, 1), (2 * nb, nb, 0)]
    lst = np.array([[r, c | (d << 31)] for r, c, d in entries], dtype=np.uint32).astype(np.uint32).view(np.int32)
    got = cm(c0, mloc + 1)
    r = emu.emu_sgemm_wide_list(mloc, nloc, nb, -1.0, 1.0, a_mem.ctypes.data, lda, tiled, b_mem.ctypes.data, ldb, tiled,
                                got.ctypes.data, got.shape[1], lst.ctypes.data, len(entries))
    assert r == 0
    want = c0.astype(np.float64).copy()
    for r0, c0_, d in entries:
        rows = slice(r0, min(mloc, r0 + nb))
        cols = slice(c0_, min(nloc, c0_ + nb))
        a_rows = pr[rows].astype(np.float64)
        # a diagonal tile takes its second operand from the ROW panel too (rows of the tile), the others from the
        # column panel (rows = the tile's columns)
        b_rows = (pr[r0:r0 + nb] if d else pc[c0_:c0_ + nb])[:cols.stop - cols.start].astype(np.float64)
        upd = a_rows @ b_rows.T
        if d:
            upd = np.tril(upd)
        want[rows, cols] -= upd
    out = got[:nloc, :mloc].T
    assert np.abs(out - want).max() <= 2e-4 * np.abs(want).max()
    assert np.all(got[:, mloc:] == np.float32(-7.7))
    # tiles that are not in the list are untouched: (0, 1) here
    assert np.array_equal(out[0:nb, nb:nloc], c0[0:nb, nb:nloc])
