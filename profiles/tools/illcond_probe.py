"""Panel solve on ill-conditioned diagonal tiles: the product path multiplies by the explicitly
inverted 128 x 128 diagonal blocks, LAPACK substitutes. For cond(A_kk) = 1e4 .. 1e12 prints the normwise backward errors
||X L^T - B|| / (||X|| ||L||) of both and the forward difference, in units of eps.   python profiles/tools/illcond_probe.py"""
import json
import os
import sys

import numpy as np
import scipy.linalg as sla
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from gpu_util import DevMat, stream_ptr, sync  # noqa: E402


def main():
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    nk, m = 512, 300
    eps = np.finfo(np.float64).eps
    rng = np.random.default_rng(1)
    q, _ = np.linalg.qr(rng.standard_normal((nk, nk)))
    out = []
    for cond in (1e2, 1e4, 1e8, 1e10, 1e12):
        ev = np.logspace(0, -np.log10(cond), nk)
        a = (q * ev) @ q.T
        a = (a + a.T) / 2
        b = rng.uniform(-1, 1, (m, nk))
        Akk, B = DevMat(a), DevMat(b)
        ib = L.dlafb_inner_block(b"d", Akk.ptr, Akk.ld, nk)
        work = DevMat(np.zeros((L.dlafb_potrf_work_elems(b"d", nk, ib), 1)))
        info = DevMat(np.zeros((1, 1), dtype=np.int32))
        s = stream_ptr()
        assert L.dlafb_potrf_tile_d(s, b"L", nk, Akk.ptr, Akk.ld, work.ptr, ib, info.ptr, 0) == 0
        assert L.dlafb_trsm_panel_d(s, nk, Akk.ptr, Akk.ld, work.ptr, ib, B.ptr, B.ld, m) == 0
        sync()
        lkk = np.tril(Akk.get())
        x = B.get()
        xref = sla.solve_triangular(lkk, b.T, lower=True).T          # X L^T = B  <=>  L X^T = B^T
        lref = np.linalg.cholesky(a)

        def bwd(xx):
            return np.linalg.norm(xx @ lkk.T - b) / (np.linalg.norm(xx) * np.linalg.norm(lkk)) / eps
        out.append({"cond_A": cond, "cond_L": float(np.linalg.cond(lkk)), "info": int(info.get()[0, 0]),
                    "potrf_rel_err_eps": float(np.linalg.norm(lkk - lref) / np.linalg.norm(lref) / eps),
                    "potrf_bwd_err_eps": float(np.linalg.norm(lkk @ lkk.T - a) / np.linalg.norm(a) / eps),
                    "bwd_err_product_eps": float(bwd(x)), "bwd_err_substitution_eps": float(bwd(xref)),
                    "fwd_diff_rel_eps": float(np.linalg.norm(x - xref) / np.linalg.norm(xref) / eps)})
        print(json.dumps(out[-1]), flush=True)


if __name__ == "__main__":
    main()
