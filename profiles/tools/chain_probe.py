"""Times the kernels of the factorization's dependent chain alone, through the C ABI (include/dlafb_kernels.h):
the diagonal-tile factorization `dlafb_potrf_tile_d` (nb x nb) and the panel solve `dlafb_trsm_panel_d` (m x nb), and
checks both against torch (potrf: L L^T = A; solve: X L^T = B). Also the command ncu is pointed at for the two
latency kernels (`--once`: two calls of each, no timing loop).

  python profiles/tools/chain_probe.py [--nb 512] [--once] [--only potrf|trsm]

Algorithmic bytes: potrf tile 2 * 8 * nb(nb+1)/2; panel solve 8*nb^2/2 + 2*8*m*nb.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nb", type=int, default=512)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--only", default="")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--rows", default="1024,4096,16384")
    args = ap.parse_args()
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    nb = args.nb
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device="cuda").manual_seed(3)
    m0 = torch.rand(nb, nb, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
    a0 = (m0 + m0.T) / 2 + 2 * nb * torch.eye(nb, dtype=torch.float64, device="cuda")
    a0 = a0.T.contiguous()  # column-major storage of a symmetric matrix
    a = torch.empty_like(a0)
    ib = L.dlafb_inner_block(b"d", a.data_ptr(), nb, nb)
    work = torch.zeros(L.dlafb_potrf_work_elems(b"d", nb, ib) + nb * nb, dtype=torch.float64, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    out = {"nb": nb, "ib": ib}

    def potrf():
        a.copy_(a0)
        pkg.check(L.dlafb_potrf_tile_d(stream, b"L", nb, a.data_ptr(), nb, work.data_ptr(), ib, info.data_ptr(), 0),
                  "potrf_tile")

    def timed(f, reps):
        for _ in range(3):
            f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            f()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1000.0 / reps

    potrf()
    torch.cuda.synchronize()
    lf = a.T.tril()  # a holds column-major data: a.T is the matrix
    ref = torch.linalg.cholesky(a0.T)
    out["potrf_max_err"] = (lf - ref).abs().max().item()
    out["info"] = int(info.item())
    if args.only in ("", "potrf"):
        if args.once:
            potrf()
        else:
            t_reset = timed(lambda: a.copy_(a0), args.reps)
            t = timed(potrf, args.reps) - t_reset
            out["potrf_tile_us"] = round(t, 1)
            out["potrf_tile_alg_gbs"] = round(2 * 8 * nb * (nb + 1) / 2 / t / 1e3, 2)
    if args.only in ("", "trsm"):
        for m in [int(x) for x in args.rows.split(",")]:
            b0 = torch.rand(nb, m, dtype=torch.float64, device="cuda", generator=g)  # column-major m x nb
            b = b0.clone()

            def solve():
                pkg.check(L.dlafb_trsm_panel_d(stream, nb, a.data_ptr(), nb, work.data_ptr(), ib, b.data_ptr(), m, m),
                          "trsm_panel")
            solve()
            torch.cuda.synchronize()
            x = b.T  # m x nb
            res = (x @ lf.T - b0.T).abs().max().item()
            out[f"trsm_m{m}_max_res"] = res
            if args.once:
                continue
            # repeated in-place solves keep the values bounded (||L^-1|| < 1 here)
            t = timed(solve, args.reps)
            out[f"trsm_m{m}_us"] = round(t, 1)
            out[f"trsm_m{m}_tflops"] = round(m * nb * nb / t / 1e6, 2)
            out[f"trsm_m{m}_alg_gbs"] = round((8 * nb * nb / 2 + 16 * m * nb) / t / 1e3, 1)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
