#!/usr/bin/env python
"""Benchmark of the distributed tiled Cholesky path (BASELINE.json metric: pdpotrf GFLOP/s, FP64, n^3/3).

  python bench.py --gpus 1 --steps K --warmup W          # double n=32768 nb=512 lower on 1 B200 (configs[1])
  torchrun ... bench.py --gpus N ...                     # double n=65536 nb=512 on a 1x2 / 2x2 / 2x4 grid (configs[2])
  python bench.py --impl reference ...                   # CPU restatement of the reference MC path (oracle port)

A step = one factorization of a fresh synthetic Hermitian positive definite matrix (regenerated on the device,
outside the timed region, before every step because the factorization is in place). Timing = CUDA events around
the factorization between two grid barriers, max over ranks per step (the miniapp's rule,
miniapp/miniapp_cholesky.cpp:136-154). Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

# read at CUDA context creation: the engine uses more streams than the default 8 hardware queues (see grid.cu)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.abspath(__file__))
FP64_PEAK_FILE = os.path.join(ROOT, "profiles", "fp64_peak_r01.json")


def load_package():
    if "dlaf_b200" in sys.modules:
        return sys.modules["dlaf_b200"]
    path = os.path.join(ROOT, "dla-future_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location("dlaf_b200", path, submodule_search_locations=[os.path.dirname(path)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dlaf_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def flops_cholesky(n, complex_=False):
    """total_ops(n^3/6 adds, n^3/6 muls): miniapp_cholesky.cpp:157-162, types.h:115-133."""
    return (4.0 if complex_ else 1.0) * n ** 3 / 3.0


def fp64_peak_tflops():
    """Measured register-resident DMMA peak on this pool's B200 (profiles/tools/fp64_peak.cu); MEASURED_PEAKS.json
    has no FP64 entry. Fallback: nominal 40 TFLOP/s."""
    try:
        d = json.load(open(FP64_PEAK_FILE))
        return max(max(v) for k, v in d.items() if k.startswith("dmma_w8")), "measured (profiles/fp64_peak_r01.json)"
    except Exception:
        return 40.0, "nominal (no measurement file)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, dev):
        self.dev, self.rows, self.p = dev, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={self.dev}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=2)
        except Exception:
            self.p.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for nme, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        sm.sort()
        busy = [x for x in sm if mx and x > 0.3 * mx] or sm
        return {"sm_mhz": busy[len(busy) // 2] if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def ncu_traffic(wl):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant launch from the committed ncu capture (it is of
    the N=1 default workload; null for any other workload)."""
    try:
        if wl["n"] != 32768 or wl["nb"] != 512 or wl["grid"] != [1, 1]:
            return None
        return json.load(open(os.path.join(ROOT, "profiles", "r01", "ncu_update_kernel.json")))["traffic_bytes_per_launch"]
    except Exception:
        return None


def cpu_baseline(n, nb, threads=None):
    """Oracle port of the reference MC path (tile tasks, one single-threaded BLAS call per tile) on the host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import dlaf_oracle as orc
    threads = threads or (os.cpu_count() or 1)
    a = orc.gen_hpd(n, nb, np.float64)
    t0 = time.perf_counter()
    info = orc.cholesky_local_threaded("L", a, nb, task_threads=threads, blas_threads=1)
    dt = time.perf_counter() - t0
    assert info == 0
    return flops_cholesky(n) / dt / 1e9, dt, threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n, nb = (16384, 512) if args.cpu_n is None else (args.cpu_n, args.nb or 512)
    vals = []
    for i in range(args.warmup + args.steps):
        g, dt, cores = cpu_baseline(n, nb)
        if i >= args.warmup:
            vals.append((g, dt))
    gf = sum(flops_cholesky(n) for _ in vals) / sum(dt for _, dt in vals) / 1e9
    ms = 1e3 * sum(dt for _, dt in vals) / len(vals)
    wl = workload(args)
    sample = f"double n={n} nb={nb} lower, one factorization per step (bounded sample of the {wl['workload']} workload)"
    print(json.dumps({
        "impl": "reference", "metric": "pdpotrf GFLOP/s (FP64, n^3/3)", "value": round(gf, 2), "unit": "GFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
        "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": wl,
        "cpu_baseline": {"value": round(gf, 2), "unit": "GFLOP/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(gf, 2), "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def grid_shape(n):
    return {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}[n]


def workload(args):
    P, Q = grid_shape(args.gpus)
    n = args.n or (32768 if args.gpus == 1 else 65536)
    nb = args.nb or 512
    return {"workload": f"double n={n} nb={nb} lower, grid {P}x{Q} on {args.gpus} B200", "n": n, "nb": nb,
            "grid": [P, Q], "uplo": "L", "l2": "inputs larger than L2 (fresh matrix every step)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--n", type=int, default=None)
    ap.add_argument("--nb", type=int, default=None)
    ap.add_argument("--cpu-n", type=int, default=None)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--timeline", default=None, help="write every rank's per-step timeline of the last step to PREFIX.rankR.json")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    pkg = load_package()
    L = pkg.lib()
    wl = workload(args)
    n, nb = wl["n"], wl["nb"]
    P, Q = wl["grid"]

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
    plan = pkg.CholeskyPlan(L, grid, "d", "L", n, nb)
    rows, cols = plan.local_rows, plan.local_cols
    ld = max(1, (rows + 63) // 64 * 64)
    a = torch.empty(ld * max(1, cols) + 64, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    row_first = L.dlafb_dist_global_tile_from_local(0, P, grid.pr, 0)
    col_first = L.dlafb_dist_global_tile_from_local(0, Q, grid.pc, 0)

    def gen(seed):
        pkg.check(L.dlafb_gen_hpd_d(stream, n, nb, rows, cols, row_first, P, col_first, Q, a.data_ptr(), ld, seed), "gen")

    def one_step(seed):
        gen(seed)
        torch.cuda.synchronize()
        grid.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.run(a.data_ptr(), ld, stream)
        e1.record()
        torch.cuda.synchronize()
        grid.barrier()
        return grid.allreduce_max(e0.elapsed_time(e1))

    for i in range(args.warmup):
        one_step(1 + i)
    assert plan.info() == 0
    plan.set_profiling(True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = L.dlafb_launch_count()
    times, upd_ms, upd_fl, upd_n = [], 0.0, 0.0, 0
    for i in range(args.steps):
        times.append(one_step(100 + i))
        m, f, c = plan.profile()
        upd_ms, upd_fl, upd_n = upd_ms + m, upd_fl + f, upd_n + c
    launches = L.dlafb_launch_count() - l0 - args.steps  # minus the generator launch of every step
    clocks = sampler.stop() if rank == 0 else None
    if args.timeline:
        with open(f"{args.timeline}.rank{rank}.json", "w") as fh:
            json.dump({"rank": rank, "grid": [P, Q], "n": n, "nb": nb,
                       "ms": times[-1], "columns": ["start", "potrf", "panel", "row_bcast", "col_bcast", "lookahead",
                                                     "bulk_begin", "bulk_end"], "steps": plan.timeline()}, fh)
    plan.set_profiling(False)
    exchange = plan.exchange_mode()
    info = plan.info()
    ms = sum(times) / len(times)
    gflops = flops_cholesky(n) / (ms * 1e-3) / 1e9

    # ---- residual of the last factorization on the device, distributed like the matrix (any number of ranks):
    #      max|A - L L^T| / max|A| <= n*eps, miniapp_cholesky.cpp:408-442 ----
    check = None
    engine_check = None
    if not args.no_check:
        try:
            a0buf = torch.empty_like(a)
            pkg.check(L.dlafb_gen_hpd_d(stream, n, nb, rows, cols, row_first, P, col_first, Q, a0buf.data_ptr(), ld,
                                        100 + args.steps - 1), "gen")
            amax_g, dmax_g = plan.residual(a0buf.data_ptr(), ld, a.data_ptr(), ld, stream)
            eps = 2.220446049250313e-16
            engine_check = {"max_norm_ratio": dmax_g / amax_g, "threshold_n_eps": n * eps,
                            "pass": bool(dmax_g / amax_g <= n * eps), "info": info,
                            "by": "dlafb_cholesky_plan_residual (distributed, on the device)"}
            del a0buf
        except Exception as ex:
            engine_check = {"error": str(ex)[:200]}
        if world > 1:
            check = engine_check

    # ---- the same residual with torch.matmul as an independent checker (1 GPU: whole matrix in one place) ----
    if world == 1 and not args.no_check:
        try:
            l_ = torch.as_strided(a, (n, n), (1, ld)).tril()
            gen(100 + args.steps - 1)
            a0 = torch.as_strided(a, (n, n), (1, ld))
            amax = a0.abs().max().item()
            ainf = a0.abs().sum(dim=1).max().item()
            blk, dmax, rowsum = 8192, 0.0, torch.zeros(n, dtype=torch.float64, device="cuda")
            for j0 in range(0, n, blk):          # checker arithmetic: torch.matmul (not the product path)
                j1 = min(n, j0 + blk)
                d = (a0[:, j0:j1] - l_ @ l_[j0:j1, :].T).abs_()
                low = d.tril(-j0)                 # referenced triangle: global row >= global column
                dmax = max(dmax, low.max().item())
                rowsum += low.sum(dim=1)
                rowsum[j0:j1] += d.tril(-j0 - 1).sum(dim=0)   # mirror of the strictly lower part
                del d, low
            eps = 2.220446049250313e-16
            check = {"max_norm_ratio": dmax / amax, "threshold_n_eps": n * eps, "pass": bool(dmax / amax <= n * eps),
                     "inf_norm_ratio_over_n_eps": rowsum.max().item() / (ainf * n * eps), "info": info}
            del l_
        except Exception as ex:  # the check must never hide the timing
            check = {"error": str(ex)[:200]}

    # ---- end to end through the reference-facing C API with HOST buffers (dlaf_pdpotrf) ----
    e2e = None
    if not args.no_e2e:
        try:
            L.dlaf_initialize(0, None, 0, None)
            pkg.check(L.dlafb_mpi_init(rank, world, ctypes.create_string_buffer(_second_uid(pkg, L, rank, world), 128)
                                       if world > 1 else None), "mpi_init")
            ctx = 77
            L.dlaf_create_grid_from_blacs_info(ctx, L.dlafb_mpi_comm_world(), P, Q, b"R")
            host = torch.empty(ld * max(1, cols), dtype=torch.float64).pin_memory()
            desca = pkg.DESC9(1, ctx, n, n, nb, nb, 0, 0, ld)
            info_c = ctypes.c_int(0)
            et = []
            esteps = max(1, min(args.steps, 3))
            for i in range(1 + esteps):
                gen(500 + i)
                host.copy_(a[: ld * max(1, cols)])
                torch.cuda.synchronize()
                grid.barrier()
                t0 = time.perf_counter()
                L.dlaf_pdpotrf(b"L", n, host.data_ptr(), 1, 1, desca, ctypes.byref(info_c))
                dt = (time.perf_counter() - t0) * 1e3
                grid.barrier()
                if i > 0:
                    et.append(grid.allreduce_max(dt))
            ems = sum(et) / len(et)
            # only the referenced (lower) triangle travels, whole tile columns from the diagonal tile down, all ranks
            nt_ = -(-n // nb)
            nbytes = 8 * sum(min(nb, n - gj * nb) * (n - gj * nb) for gj in range(nt_))
            e2e = {"value": round(flops_cholesky(n) / (ems * 1e-3) / 1e9, 1), "unit": "GFLOP/s",
                   "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes, "ms_per_step": round(ems, 2),
                   "steps": esteps, "api": "dlaf_pdpotrf (host pointer, pinned; lower triangle only, D2H of finished column panels overlapped)", "info": info_c.value}
            L.dlaf_free_grid(ctx)
            L.dlaf_finalize()
        except Exception as ex:
            e2e = {"error": str(ex)[:200]}

    if rank == 0:
        peak, peak_src = fp64_peak_tflops()
        ach = upd_fl / (upd_ms * 1e-3) / 1e12 if upd_ms > 0 else None
        out = {
            "metric": "pdpotrf GFLOP/s (FP64, n^3/3)", "value": round(gflops, 1), "unit": "GFLOP/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
            "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": wl,
            "frac_of_fp64_tensor_peak": round(gflops / 1e3 / (peak * world), 4),
            "fp64_tensor_peak_tflops_per_gpu": peak, "fp64_peak_source": peak_src,
            "roofline": {"bound": "tensor", "kernel": "gemm_nt_dmma_kernel (bulk trailing update, rank 0)",
                         "achieved": round(ach, 2) if ach else None, "peak": peak, "unit": "TFLOP/s",
                         "frac": round(ach / peak, 4) if ach else None, "traffic": ncu_traffic(wl),
                         "launches_timed": upd_n,
                         "note": "FP64 DMMA peak: " + peak_src + "; achieved = algorithmic flops of the bulk launches / "
                                 "their CUDA-event time, while the high-priority stream shares the SMs; traffic = "
                                 "dram bytes read+written by the step-0 bulk launch of the n=32768 workload "
                                 "(profiles/r01/ncu_update_kernel.json: 94.5 % tensor pipe for the kernel alone)"},
            "clocks": clocks, "gpu_launches": int(launches), "exchange": exchange, "e2e": e2e, "check": check,
            "engine_check": engine_check,
            "times_ms": [round(t, 2) for t in times],
        }
        if not args.no_cpu and world == 1:
            try:
                g, dt, cores = cpu_baseline(args.cpu_n or 28672, nb)
                out["cpu_baseline"] = {"value": round(g, 2), "unit": "GFLOP/s", "cores": cores, "kind": "port",
                                       "sample": f"double n={args.cpu_n or 28672} nb={nb} lower, 1 factorization, "
                                                 f"{dt:.1f} s (oracle port of the reference MC tile-task path)"}
            except Exception as ex:
                out["cpu_baseline"] = {"error": str(ex)[:200]}
        print(json.dumps(out))
    plan.close()
    grid.close()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


def _second_uid(pkg, L, rank, world):
    """A fresh NCCL id for the C-API world communicator, distributed with torch.distributed."""
    import torch
    import torch.distributed as dist
    buf = ctypes.create_string_buffer(128)
    if rank == 0:
        pkg.check(L.dlafb_nccl_unique_id(buf), "unique_id")
    t = torch.tensor(list(buf.raw), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    return bytes(t.cpu().tolist())


if __name__ == "__main__":
    main()
