"""Helper of tests/test_zz_sgemm_wide_gpu.py: runs float contractions and one float factorization through the C ABI in
THIS process (the switch DLAFB_SGEMM_WIDE is read once per process) and stores the results in an .npz file.
    python tests/wide_sgemm_probe.py OUT.npz
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from conftest import load_package  # noqa: E402
from gpu_util import DevMat, stream_ptr, sync  # noqa: E402
import dlaf_oracle as orc  # noqa: E402

SHAPES = [(128, 128, 8, 0), (200, 131, 37, 0), (1, 1, 1, 0), (64, 300, 16, 0), (257, 257, 24, 1), (1000, 777, 512, 0),
          (2048, 2048, 256, 1)]


def main():
    out = {}
    pkg = load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    rng = np.random.default_rng(12)
    for i, (m, n, k, lower) in enumerate(SHAPES):
        a = rng.standard_normal((m, k)).astype(np.float32)
        b = rng.standard_normal((n, k)).astype(np.float32)
        c = rng.standard_normal((m, n)).astype(np.float32)
        A, B, C = DevMat(a, ld=m + 3), DevMat(b, ld=n + 1), DevMat(c, ld=m + 2)
        err = L.dlafb_gemm_nt_s(stream_ptr(), m, n, k, -1.0, A.ptr, A.ld, B.ptr, B.ld, 1.0, C.ptr, C.ld, lower)
        sync()
        assert err == 0
        out[f"gemm{i}"] = C.get()
        out[f"gemm{i}_pad_ok"] = np.array(bool(np.all(C.pad() == np.float32(-7.7))))
    # the factorization's list-mode update for float: n = 1536, nb = 256 (tile = 2 wide blocks per side)
    n, nb = 1536, 256
    a = orc.gen_hpd(n, nb, np.float32)
    A = DevMat(a)
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, "s", "L", n, nb)
    plan.run(A.ptr, A.ld, stream_ptr())
    out["potrf_info"] = np.array(plan.info())
    sync()
    plan.close()
    grid.close()
    out["potrf"] = A.get()
    np.savez(sys.argv[1], **out)


if __name__ == "__main__":
    main()
