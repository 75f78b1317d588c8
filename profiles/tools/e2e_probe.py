"""Where the end-to-end time of a host-mirrored run goes (1 GPU): wall clock of dlafb_cholesky_plan_run_host with a
pinned host block, the engine's own step timeline for that run, and the raw pinned->device copy rate for reference.

  python profiles/tools/e2e_probe.py [n nb]
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    nb = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
    ld = n
    dev = torch.empty(ld * n + 64, dtype=torch.float64, device="cuda")
    src = torch.empty(ld * n + 64, dtype=torch.float64, device="cuda")
    host = torch.empty(ld * n, dtype=torch.float64).pin_memory()
    stream = torch.cuda.current_stream().cuda_stream
    out = {"n": n, "nb": nb}
    # raw copy rates
    torch.cuda.synchronize()
    for name, f in (("h2d_full_GBps", lambda: dev[: ld * n].copy_(host, non_blocking=True)),
                    ("d2h_full_GBps", lambda: host.copy_(dev[: ld * n], non_blocking=True))):
        f()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        f()
        torch.cuda.synchronize()
        out[name] = round(ld * n * 8 / (time.perf_counter() - t0) / 1e9, 1)
    plan.set_profiling(True)
    walls = []
    for i in range(3):
        pkg.check(L.dlafb_gen_hpd_d(stream, n, nb, n, n, 0, 1, 0, 1, src.data_ptr(), ld, 11 + i), "gen")
        host.copy_(src[: ld * n])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pkg.check(L.dlafb_cholesky_plan_run_host(plan.h, host.data_ptr(), ld, dev.data_ptr(), ld, stream), "run_host")
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        walls.append((round((t1 - t0) * 1e3, 2), round((t2 - t0) * 1e3, 2)))
    assert plan.info() == 0
    tl = plan.timeline()
    out["wall_ms_enqueue_total"] = walls
    out["steps"] = {k: [round(x, 2) for x in tl[k]] for k in (0, 1, 2, 3, 5, 8, 10, 11, 12, 16, 24, 32, 48, 60, 62)
                    if k < len(tl)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
