"""Timing of the distributed "next" rows under torchrun, one process per GPU: the Cholesky factor of a synthetic HPD
matrix (the engine, as bench.py), then with that factor
  - the distributed triangular solver   L X = B and L^H X = B   (flops m n k per solve, miniapp_triangular_solver.cpp:143-144)
  - the distributed generalized-to-standard reduction A <- L^-1 A L^-H   (flops n^3, miniapp_gen_to_std.cpp)
Device-resident operands, CUDA events around each call, max over ranks, best of --steps.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29911 \
      profiles/tools/bench_dist_rows.py --n 65536 --nb 512 [--nrhs 65536] [--steps 3]
  (also runs with a single process: 1 x 1 grid, no NCCL)

Sanity check only (finite results and, for the solves, max|L (L^H X) - B| through the same solver pair is NOT formed
here): parity is the job of tests/dist_trsolve_check.py and tests/dist_g2s_check.py.
"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", "--size", dest="n", type=int, default=32768)
    ap.add_argument("--nb", type=int, default=512)
    ap.add_argument("--nrhs", type=int, default=None)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    pkg = bench.load_package()
    L = pkg.lib()
    P, Q = bench.grid_shape(world)
    n, nb = args.n, args.nb
    nrhs = args.nrhs or n
    uid = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        buf = ctypes.create_string_buffer(128)
        if rank == 0:
            pkg.check(L.dlafb_nccl_unique_id(buf), "unique_id")
        t = torch.tensor(list(buf.raw), dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().tolist())
    grid = pkg.Grid(L, P, Q, rank, world, uid, "R")
    stream = torch.cuda.current_stream().cuda_stream
    chol = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
    rows, cols = chol.local_rows, chol.local_cols
    ld = max(2, (rows + 63) // 64 * 64)
    row_first = L.dlafb_dist_global_tile_from_local(0, P, grid.pr, 0)
    col_first = L.dlafb_dist_global_tile_from_local(0, Q, grid.pc, 0)
    fac = torch.empty(ld * max(1, cols) + 64, dtype=torch.float64, device="cuda")
    pkg.check(L.dlafb_gen_hpd_d(stream, n, nb, rows, cols, row_first, P, col_first, Q, fac.data_ptr(), ld, 1), "gen")
    chol.run(fac.data_ptr(), ld, stream)
    assert chol.info() == 0

    def timed(fn, prep):
        best = None
        for _ in range(args.steps + 1):
            prep()
            grid.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ms = grid.allreduce_max(e0.elapsed_time(e1))
            best = ms if best is None or _ == 1 else min(best, ms)   # the first pass is the warm-up
        return best

    out = {"n": n, "nb": nb, "nrhs": nrhs, "grid": [P, Q], "gpus": world}
    # ---- triangular solves with the factor
    for op in "NC":
        plan = pkg.TriangularSolverDistPlan(L, grid, "d", "L", "L", op, "N", n, nrhs, nb, nb)
        brows, bcols = plan.b_rows, plan.b_cols
        ldb = max(2, (brows + 63) // 64 * 64)
        b0 = torch.randn(ldb * max(1, bcols) + 64, dtype=torch.float64, device="cuda")
        b = torch.empty_like(b0)
        ms = timed(lambda: plan.run(1.0, fac.data_ptr(), ld, b.data_ptr(), ldb, stream), lambda: b.copy_(b0))
        out[f"trsolve_LL{op}_ms"] = round(ms, 3)
        out[f"trsolve_LL{op}_gflops"] = round(float(n) * n * nrhs / ms / 1e6, 1)
        out[f"trsolve_LL{op}_finite"] = bool(torch.isfinite(b[:ldb * max(1, bcols)].view(-1, ldb)[:, :brows]).all().item())
        out[f"trsolve_LL{op}_workspace_MiB"] = round(plan.workspace_bytes() / 2 ** 20, 1)
        plan.close()
        del b, b0
        torch.cuda.empty_cache()
    # ---- generalized-to-standard reduction with the factor (A: another synthetic Hermitian matrix, lower part used)
    g2s = pkg.GenToStdDistPlan(L, grid, "d", "L", n, nb)
    a0 = torch.empty(ld * max(1, cols) + 64, dtype=torch.float64, device="cuda")
    pkg.check(L.dlafb_gen_hpd_d(stream, n, nb, rows, cols, row_first, P, col_first, Q, a0.data_ptr(), ld, 2), "gen")
    a = torch.empty_like(a0)
    ms = timed(lambda: g2s.run(a.data_ptr(), ld, fac.data_ptr(), ld, stream), lambda: a.copy_(a0))
    out["gen_to_std_ms"] = round(ms, 3)
    out["gen_to_std_gflops"] = round(float(n) ** 3 / ms / 1e6, 1)
    out["gen_to_std_workspace_MiB"] = round(g2s.workspace_bytes() / 2 ** 20, 1)
    g2s.close()
    chol.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    grid.close()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
