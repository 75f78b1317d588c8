"""-m gpu: the distributed triangular solver on real GPUs (tests/dist_trsolve_check.py under torchrun), plus the
single-rank grid through the same distributed entry point on one GPU. Multi-rank cases are skipped on boxes with a
single GPU. (The file name sorts last on purpose: written after the round's GPU budget was spent.)"""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("x", "sdcz")
@pytest.mark.parametrize("side", "LR")
def test_single_rank_grid_matches_reference_known_answer(pkg, L, x, side):
    """A 1 x 1 grid runs the distributed algorithm (panels, transposed blocks, strided tile inverse) without NCCL."""
    import dlaf_oracle as orc
    from gpu_util import NP, DevMat, stream_ptr, sync
    dt = NP[x]
    alpha = orc.triangular_test_alpha(dt)
    grid = pkg.Grid(L)
    for uplo in "LU":
        for op in "NTC":
            for diag in "NU":
                for m, n, mb, nb in orc.TRIANGULAR_TEST_SIZES:
                    a, b, want = orc.triangular_system(side, uplo, op, diag, alpha, m, n, dt)
                    A = DevMat(a, ld=max(1, a.shape[0]) + 1)
                    B = DevMat(b, ld=max(1, m) + 2)
                    plan = pkg.TriangularSolverDistPlan(L, grid, x, side, uplo, op, diag, m, n, mb, nb)
                    plan.run(alpha, A.ptr, A.ld, B.ptr, B.ld, stream_ptr())
                    sync()
                    plan.close()
                    tol = 20 * (m + 1) * orc.type_error(dt)
                    assert orc.check_near(B.get(), want, tol, tol) == 0, (uplo, op, diag, m, n, mb, nb)
                    assert np.all(B.pad() == dt(-7.7))
    grid.close()


def _worlds():
    import torch
    n = torch.cuda.device_count()
    env = os.environ.get("DLAFB_DIST_WORLDS")
    ws = [int(w) for w in env.split(",")] if env else [2, 4, 6, 8]
    return [w for w in ws if w <= n]


@pytest.mark.parametrize("world", [2, 4, 6, 8])
def test_distributed_parity(world):
    if world not in _worlds():
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ)
    if world > 2:
        env.setdefault("DLAFB_DIST_QUICK", "1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + world), os.path.join(ROOT, "tests",
                                                                                          "dist_trsolve_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=2400, env=env)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "total_failures=0" in r.stdout, tail
