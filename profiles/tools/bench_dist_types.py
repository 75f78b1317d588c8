"""Factorization timing for the BASELINE.json configurations that bench.py's headline line does not cover, on any number
of GPUs (torchrun, one process per GPU; also runs as a single process):

  configs[3]  complex<double> n=40960 nb=512, grid 2x4:   --type z --n 40960 --nb 512
  configs[4]  weak scaling double n=32768*sqrt(P) nb=1024: --type d --n 32768|46080|65536|92160 --nb 1024
              (also the un-rounded 46341 / 92682 once, SURVEY.md section 8: ragged edge tiles, no residual check)
  configs[4]  float n=65536 at 8 GPUs:                     --type s --n 65536 --nb 512   (DLAFB_SGEMM_WIDE=0/1)

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29951 \
      profiles/tools/bench_dist_types.py --type z --n 40960 --nb 512

Metric as miniapp_cholesky.cpp:157-162: flops = n^3/3 (x4 for complex), time = CUDA events between grid barriers, max over
ranks, best of --steps after 2 warm-ups, fresh matrix every step. The distributed device-side residual check
(dlafb_cholesky_plan_residual, uplo L) runs once after the timing when n is a multiple of nb.
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

TORCH = {"s": torch.float32, "d": torch.float64, "c": torch.complex64, "z": torch.complex128}
EPS = {"s": 2.0 ** -23, "d": 2.0 ** -52, "c": 2.0 ** -23, "z": 2.0 ** -52}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="z")
    ap.add_argument("--n", "--size", dest="n", type=int, default=40960)
    ap.add_argument("--nb", type=int, default=512)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    pkg = bench.load_package()
    L = pkg.lib()
    P, Q = bench.grid_shape(world)
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
    x, n, nb = args.type, args.n, args.nb
    plan = pkg.CholeskyPlan(L, grid, x, "L", n, nb)
    rows, cols = plan.local_rows, plan.local_cols
    ld = max(2, (rows + 63) // 64 * 64)
    a = torch.empty(ld * max(1, cols) + 64, dtype=TORCH[x], device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rf = L.dlafb_dist_global_tile_from_local(0, P, grid.pr, 0)
    cf = L.dlafb_dist_global_tile_from_local(0, Q, grid.pc, 0)
    gen = getattr(L, f"dlafb_gen_hpd_{x}")
    times = []
    for i in range(2 + args.steps):
        pkg.check(gen(stream, n, nb, rows, cols, rf, P, cf, Q, a.data_ptr(), ld, 11 + i), "gen")
        grid.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.run(a.data_ptr(), ld, stream)
        e1.record()
        info = plan.info()
        torch.cuda.synchronize()
        ms = grid.allreduce_max(e0.elapsed_time(e1))
        assert info == 0, info
        if i >= 2:
            times.append(ms)
    check = None
    if not args.no_check and n % nb == 0:
        a0 = torch.empty_like(a)
        pkg.check(gen(stream, n, nb, rows, cols, rf, P, cf, Q, a0.data_ptr(), ld, 11 + 1 + args.steps), "gen")
        amax, dmax = plan.residual(a0.data_ptr(), ld, a.data_ptr(), ld, stream)
        check = {"max_norm_ratio": dmax / amax if amax > 0 else 0.0, "threshold_n_eps": n * EPS[x]}
        check["pass"] = bool(check["max_norm_ratio"] <= check["threshold_n_eps"])
    mode = plan.exchange_mode()
    plan.close()
    best = min(times)
    flops = (4.0 if x in "cz" else 1.0) * float(n) ** 3 / 3
    if rank == 0:
        print(json.dumps({"what": "cholesky", "type": x, "n": n, "nb": nb, "grid": [P, Q], "gpus": world, "ms_best": round(best, 3),
                          "ms_all": [round(t, 3) for t in times], "gflops_real_equiv": round(flops / best / 1e6, 1),
                          "exchange": mode, "sgemm_wide": os.environ.get("DLAFB_SGEMM_WIDE", "0"), "check": check}),
              flush=True)
    grid.close()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
