"""-m gpu: the matrix redistribution (include/dlafb_matrix.h) on one GPU (1 x 1 grid: the rectangle-list kernel alone,
every sub-tile stays on the rank) and on several GPUs (tests/dist_redistribute_check.py under torchrun).
(The file name sorts last on purpose: written after the round's GPU budget was spent.)"""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("x", "sdcz")
def test_single_rank_is_a_bit_exact_copy(pkg, L, x):
    from gpu_util import NP, DevMat, stream_ptr, sync
    dt = NP[x]
    rng = np.random.default_rng(6)
    grid = pkg.Grid(L)
    for m, n, st, dtile in [(37, 29, (4, 3), (8, 6)), (37, 29, (8, 6), (4, 3)), (1000, 700, (64, 64), (256, 128)),
                            (2048, 2048, (512, 512), (64, 64)), (1, 1, (2, 2), (4, 4)), (0, 5, (2, 2), (4, 4))]:
        a = rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if np.dtype(dt).kind == "c" else 0)
        a = a.astype(dt)
        S = DevMat(a, ld=max(1, m) + 1)
        D = DevMat(np.full((m, n), dt(-5.5)), ld=max(1, m) + 3)
        plan = pkg.RedistributePlan(L, grid, x, m, n, st, (0, 0), dtile, (0, 0))
        assert (plan.src_rows, plan.src_cols, plan.dst_rows, plan.dst_cols) == (m, n, m, n)
        plan.run(S.ptr, S.ld, D.ptr, D.ld, stream_ptr())
        plan.run(S.ptr, S.ld, D.ptr, D.ld, stream_ptr())   # a second run reuses the uploaded lists
        sync()
        plan.close()
        assert np.array_equal(D.get(), a), (m, n, st, dtile)
        assert np.all(D.pad() == dt(-7.7))
        assert np.array_equal(S.get(), a)
    grid.close()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_distributed_redistribution(world):
    import torch
    env = os.environ.get("DLAFB_DIST_WORLDS")
    worlds = [int(w) for w in env.split(",")] if env else [2, 4, 8]
    if world not in worlds or torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29900 + world),
           os.path.join(ROOT, "tests", "dist_redistribute_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=1500)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "total_failures=0" in r.stdout, tail


@pytest.mark.parametrize("x,n,nb", [("d", 2304, 64), ("d", 2100, 48), ("z", 2048, 32), ("s", 2048, 128)])
@pytest.mark.parametrize("uplo", "LU")
def test_pxpotrf_with_redistribution_to_large_blocks(pkg, L, monkeypatch, x, n, nb, uplo):
    """DLAFB_POTRF_REDISTRIBUTE=1: a caller with ScaLAPACK-sized blocks gets redistribute -> Cholesky at <= 512 ->
    redistribute back inside dlaf_p?potrf; the result must be what the plain path gives (same tolerance to the oracle),
    and everything outside the referenced triangle must come back untouched."""
    import ctypes

    import dlaf_oracle as orc
    from gpu_util import NP
    dt = NP[x]
    ctx = 47
    L.dlaf_initialize(0, None, 0, None)
    pkg.check(L.dlafb_mpi_init(0, 1, None), "mpi_init")
    L.dlaf_create_grid_from_blacs_info(ctx, L.dlafb_mpi_comm_world(), 1, 1, b"R")
    try:
        a = orc.gen_hpd(n, nb, dt)
        ref = a.copy(order="F")
        assert orc.cholesky_local(uplo, ref, nb) == 0
        keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
        a_in = np.where(keep, a, dt(-9.9))
        tol = 50 * n * float(np.finfo(dt).eps) * np.sqrt(2.0)
        for flag in ("1", "0"):
            monkeypatch.setenv("DLAFB_POTRF_REDISTRIBUTE", flag)
            ld = n + 5
            h = np.full((n, ld), dt(-7.7))
            h[:, :n] = a_in.T
            desca = pkg.DESC9(1, ctx, n, n, nb, nb, 0, 0, ld)
            info = ctypes.c_int(-1)
            getattr(L, f"dlaf_p{x}potrf")(uplo.encode(), n, h.ctypes.data, 1, 1, desca, ctypes.byref(info))
            assert info.value == 0, flag
            got = np.ascontiguousarray(h[:, :n].T)
            assert np.abs(got[keep] - ref[keep]).max() <= tol * max(1.0, np.abs(ref[keep]).max()), flag
            assert np.all(got[~keep] == dt(-9.9)), flag
            assert np.all(h[:, n:] == dt(-7.7)), flag
    finally:
        L.dlaf_free_grid(ctx)
        L.dlaf_finalize()


@pytest.mark.parametrize("uplo", "LU")
def test_pdpotrf_device_pointer_small_blocks_redistributes_by_default(pkg, L, monkeypatch, uplo):
    """Without DLAFB_POTRF_REDISTRIBUTE in the environment a DEVICE-pointer caller with nb < 128 (n >= 2048) goes
    through redistribute -> Cholesky at 512 -> redistribute back (2.5x faster at nb = 64, profiles/r02/small_blocks.txt):
    same result as the oracle, nothing outside the referenced triangle touched, and the C API keeps the entry cached."""
    import ctypes

    import dlaf_oracle as orc
    from gpu_util import DevMat, sync
    monkeypatch.delenv("DLAFB_POTRF_REDISTRIBUTE", raising=False)
    ctx = 48
    L.dlaf_initialize(0, None, 0, None)
    pkg.check(L.dlafb_mpi_init(0, 1, None), "mpi_init")
    L.dlaf_create_grid_from_blacs_info(ctx, L.dlafb_mpi_comm_world(), 1, 1, b"R")
    try:
        n, nb = 2304, 64
        a = orc.gen_hpd(n, nb, np.float64)
        ref = a.copy(order="F")
        assert orc.cholesky_local(uplo, ref, nb) == 0
        keep = (np.tril if uplo == "L" else np.triu)(np.ones((n, n))) > 0
        a_in = np.where(keep, a, -9.9)
        tol = 50 * n * float(np.finfo(np.float64).eps) * np.sqrt(2.0)
        before = L.dlafb_capi_cached_plans()
        for rep in range(2):   # the second call reuses the cached plans and buffers
            A = DevMat(a_in, ld=n + 8)
            desca = pkg.DESC9(1, ctx, n, n, nb, nb, 0, 0, A.ld)
            info = ctypes.c_int(-1)
            L.dlaf_pdpotrf(uplo.encode(), n, A.ptr, 1, 1, desca, ctypes.byref(info))
            sync()
            assert info.value == 0
            got = A.get()
            assert np.abs(got[keep] - ref[keep]).max() <= tol * max(1.0, np.abs(ref[keep]).max())
            assert np.all(got[~keep] == -9.9) and np.all(A.pad() == -7.7)
            assert L.dlafb_capi_cached_plans() == before + 1
    finally:
        L.dlaf_free_grid(ctx)
        L.dlaf_finalize()
