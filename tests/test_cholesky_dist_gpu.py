"""-m gpu: multi-GPU parity of the distributed factorization (tests/dist_gpu_check.py under torchrun).
Skipped on boxes with a single GPU. DLAFB_DIST_WORLDS (e.g. "2,4,6,8") selects the world sizes; default: 2, and
every further size in {4, 6, 8} the box can host."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


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
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + world), os.path.join(ROOT, "tests",
                                                                                          "dist_gpu_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=1500)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "total_failures=0" in r.stdout, tail
