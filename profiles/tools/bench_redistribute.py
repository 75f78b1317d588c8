"""Timing of the matrix redistribution (include/dlafb_matrix.h): achieved bytes/s of the rectangle-list kernel against
the measured HBM copy bandwidth (MEASURED_PEAKS.json), and the whole collective under torchrun.

  python profiles/tools/bench_redistribute.py [--n 32768] [--nb-src 64] [--nb-dst 512] [--type d]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29921 \
      profiles/tools/bench_redistribute.py --n 65536

On one GPU every sub-tile stays on the rank: one launch of copy_rects_kernel moves the whole matrix (algorithmic bytes
= 2 * s * n^2: every element read once and written once). On N GPUs: gather + grouped NCCL send/recv + scatter
(bytes per rank = 2 * s * local elements for each of the two kernels, s * moved elements over NVLink each way).
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", "--size", dest="n", type=int, default=32768)
    ap.add_argument("--nb-src", type=int, default=64)
    ap.add_argument("--nb-dst", type=int, default=512)
    ap.add_argument("--type", default="d")
    ap.add_argument("--steps", type=int, default=5)
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
    n, x = args.n, args.type
    plan = pkg.RedistributePlan(L, grid, x, n, n, (args.nb_src, args.nb_src), (0, 0), (args.nb_dst, args.nb_dst), (0, 0))
    lds, ldd = max(2, plan.src_rows), max(2, plan.dst_rows)
    src = torch.randn(lds * max(1, plan.src_cols), dtype=torch.float64, device="cuda").to(TORCH[x])
    dst = torch.empty(ldd * max(1, plan.dst_cols), dtype=TORCH[x], device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    best = None
    for i in range(args.steps + 2):
        grid.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.run(src.data_ptr(), lds, dst.data_ptr(), ldd, stream)
        e1.record()
        torch.cuda.synchronize()
        ms = grid.allreduce_max(e0.elapsed_time(e1))
        if i >= 2:
            best = ms if best is None else min(best, ms)
    plan.close()
    es = src.element_size()
    local_elems = plan.src_rows * plan.src_cols
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm = peaks.get("hbm_gbs")
    # one GPU: one kernel, 2*s bytes per element; N GPUs: gather + scatter = 4*s bytes per local element through HBM
    algo_bytes = (2 if world == 1 else 4) * es * local_elems
    gbs = algo_bytes / best / 1e6
    if rank == 0:
        print(json.dumps({"what": "redistribute", "type": x, "n": n, "nb_src": args.nb_src, "nb_dst": args.nb_dst,
                          "grid": [P, Q], "ms_best": round(best, 3), "algorithmic_bytes_per_rank": algo_bytes,
                          "achieved_GBs_per_rank": round(gbs, 1), "hbm_peak_GBs": hbm,
                          "frac_of_hbm": round(gbs / hbm, 3) if hbm else None,
                          "note": "multi-rank time includes the NVLink exchange between the two kernels"}), flush=True)
    grid.close()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
