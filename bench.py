#!/usr/bin/env python
"""Benchmark of the distributed tiled Cholesky path (BASELINE.json metric: pdpotrf GFLOP/s, FP64, n^3/3).

  python bench.py --gpus 1 --steps K --warmup W          # double n=32768 nb=512 lower on 1 B200 (configs[1])
  torchrun ... bench.py --gpus N ...                     # double n=65536 nb=512 on a 1x2 / 2x2 / 2x4 grid (configs[2])
  ... bench.py --config zpotrf|float|weak ...            # the other BASELINE.json configurations (configs[3], [4])
  python bench.py --impl reference ...                   # CPU restatement of the reference MC path (oracle port)

At N > 1 an untimed `dist_parity` leg follows the timed steps: the reference's analytical known-answer cases
(test/unit/factorization/test_cholesky.cpp:54-101) on the run's own grid, compared element-wise with the closed form,
and a residual of the timed factor by an INDEPENDENT checker (torch.distributed all_gather + torch.matmul, none of the
product's exchange code or kernels).

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


def kernel_source_hash():
    """sha256 of the update kernel's source files: a committed ncu capture is only quoted while it still describes
    the kernel that runs."""
    import hashlib
    h = hashlib.sha256()
    for f in ("gemm_dmma.cu", "gemm_dmma.cuh"):
        with open(os.path.join(ROOT, "dla-future_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_traffic(wl):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant launch from the committed ncu capture (of the N=1
    default workload; null for any other workload, and null when the kernel source has changed since the capture)."""
    try:
        if wl["n"] != 32768 or wl["nb"] != 512 or wl["grid"] != [1, 1] or wl.get("type", "d") != "d":
            return None
        d = json.load(open(os.path.join(ROOT, "profiles", "r01", "ncu_update_kernel.json")))
        if d.get("kernel_source_sha256_16") != kernel_source_hash():
            return None
        return d["traffic_bytes_per_launch"]
    except Exception:
        return None


def cpu_baseline(n, nb, threads=None, x="d"):
    """Oracle port of the reference MC path (tile tasks, one single-threaded BLAS call per tile) on the host cores."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import dlaf_oracle as orc
    threads = threads or (os.cpu_count() or 1)
    a = orc.gen_hpd(n, nb, {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}[x])
    t0 = time.perf_counter()
    info = orc.cholesky_local_threaded("L", a, nb, task_threads=threads, blas_threads=1)
    dt = time.perf_counter() - t0
    assert info == 0
    return flops_cholesky(n, x in "cz") / dt / 1e9, dt, threads


def run_reference(args):
    """The reference's own CPU (MC) path = the oracle port, all host threads, on this arm's config. The CPU cannot
    finish K + W factorizations of n = 65536 in minutes, so each step factorizes the LARGEST n <= the config's n
    (multiple of 4096) for which the whole run stays within ~150 s, chosen from a short calibration run; the line says
    which n it ran (`config.n_run`, `config.same_size`), and GFLOP/s of the tile-task path is nearly flat in n beyond
    16384 (1.0 - 1.2 TFLOP/s on 16 cores between n = 16384 and 28672, round 1)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = workload(args)
    x, nb, n_cfg = wl["type"], wl["nb"], wl["n"]
    cplx = x in "cz"
    if args.cpu_n is not None:
        n = args.cpu_n
    else:
        g_cal, _, _ = cpu_baseline(4096, nb, x=x)          # ~0.1 s: calibrates the host
        g_cal = max(g_cal, 1.0) * 1.6                      # larger n runs faster per flop than n = 4096 (measured ~1.6x)
        budget = 150.0 / max(1, args.steps + args.warmup)  # seconds per step
        n = 4096
        for cand in range(4096, n_cfg + 1, 4096):
            if flops_cholesky(cand, cplx) / (g_cal * 1e9) <= budget:
                n = cand
        n = min(n, n_cfg)
    vals = []
    for i in range(args.warmup + args.steps):
        g, dt, cores = cpu_baseline(n, nb, x=x)
        if i >= args.warmup:
            vals.append((g, dt))
    gf = sum(flops_cholesky(n, cplx) for _ in vals) / sum(dt for _, dt in vals) / 1e9
    ms = 1e3 * sum(dt for _, dt in vals) / len(vals)
    wl["n_run"] = n
    wl["same_size"] = bool(n == n_cfg)
    wl["grid_run"] = [1, 1]
    if n != n_cfg:
        wl["workload"] = (f"{TYPE_NAME[x]} n={n} nb={nb} lower on the host cores (bounded sample: the config's "
                          f"n={n_cfg} does not finish {args.steps}+{args.warmup} CPU steps within minutes)")
    else:
        wl["workload"] = f"{TYPE_NAME[x]} n={n} nb={nb} lower on the host cores"
    sample = f"{TYPE_NAME[x]} n={n} nb={nb} lower, one factorization per step, {ms / 1e3:.1f} s per step"
    print(json.dumps({
        "impl": "reference", "metric": metric_name(x), "value": round(gf, 2), "unit": "GFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
        "higher_is_better": True, "scaling": "weak" if (args.gpus == 1 or wl["config"] == "weak") else "strong",
        "vs_baseline": None, "dtype": DTYPE[x], "data": "synthetic", "config": wl,
        "cpu_baseline": {"value": round(gf, 2), "unit": "GFLOP/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(gf, 2), "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def grid_shape(n):
    return {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}[n]


# BASELINE.json configs. headline = configs[1] (1 GPU) / configs[2] (2, 4, 8 GPUs); zpotrf = configs[3];
# weak / float = configs[4]. Weak scaling: n = 32768 * sqrt(P) rounded DOWN to a multiple of nb = 1024 so that the
# reference's checker applies (miniapp_cholesky.cpp:90-95); --n 46341 / 92682 runs the un-rounded sizes.
WEAK_N = {1: 32768, 2: 46080, 4: 65536, 8: 92160}
TYPE_NAME = {"s": "float", "d": "double", "c": "complex<float>", "z": "complex<double>"}
DTYPE = {"s": "f32", "d": "f64", "c": "c64", "z": "c128"}


def workload(args):
    P, Q = grid_shape(args.gpus)
    cfg = getattr(args, "config", "headline")
    x, nb = "d", 512
    if cfg == "zpotrf":
        x, n = "z", 40960
    elif cfg == "float":
        x, n = "s", 65536
    elif cfg == "weak":
        n, nb = WEAK_N[args.gpus], 1024
    else:
        n = 32768 if args.gpus == 1 else 65536
    n = args.n or n
    nb = args.nb or nb
    return {"workload": f"{TYPE_NAME[x]} n={n} nb={nb} lower, grid {P}x{Q} on {args.gpus} B200", "n": n, "nb": nb,
            "grid": [P, Q], "uplo": "L", "type": x, "config": cfg,
            "l2": "inputs larger than L2 (fresh matrix every step)"}


def metric_name(x):
    return {"d": "pdpotrf GFLOP/s (FP64, n^3/3)", "s": "pspotrf GFLOP/s (FP32, n^3/3)",
            "z": "pzpotrf GFLOP/s (complex FP64, 4n^3/3)", "c": "pcpotrf GFLOP/s (complex FP32, 4n^3/3)"}[x]


# ---- closed-form test matrices of the reference (test/include/dlaf_test/matrix/util_generic_lapack.h:27-68),
#      restated here for the dist_parity leg: t_ij = 2^-|i-j| e^{I(j-i)}, a_ij = (4^(min(i,j)+1) - 1) / (3 2^(i+j))
#      e^{I(j-i)} on the referenced triangle, -9.9 elsewhere ----
def analytic_pair(n, uplo, dtype, rows, cols):
    import numpy as np
    i = np.asarray(rows, dtype=np.float64)[:, None]
    j = np.asarray(cols, dtype=np.float64)[None, :]
    mn = np.minimum(i, j)
    mag_a = (np.exp2(2 * mn + 2 - i - j) - np.exp2(-(i + j))) / 3
    mag_t = np.exp2(-np.abs(i - j))
    if np.dtype(dtype).kind == "c":
        ph = np.exp(1j * (j - i))
        a, t = mag_a * ph, mag_t * ph
    else:
        a, t = mag_a + 0 * j, mag_t + 0 * j
    unref = (i < j) if uplo == "L" else (i > j)
    return np.where(unref, -9.9, a).astype(dtype), np.where(unref, -9.9, t).astype(dtype)


REF_SIZES = [(0, 2), (5, 8), (34, 34), (4, 3), (16, 10), (34, 13), (32, 5)]  # test_cholesky.cpp:54-62


def dist_parity_cases(pkg, L, grid, rank, world, P, Q, new_uid):
    """The reference's distributed known-answer cases on this run's grid (test_cholesky.cpp:54-101): (m, mb) table,
    both uplo, double and complex<double>, source rank (max(0, P-1), min(1, Q-1)); every rank compares its own block
    element-wise with the closed form at 4(m+1)*err (CHECK_MATRIX_NEAR: absolute OR relative)."""
    import numpy as np
    import torch
    stream = torch.cuda.current_stream().cuda_stream
    src = (max(0, P - 1), min(1, Q - 1))
    cases = failures = 0
    for x, dt in (("d", np.float64), ("z", np.complex128)):
        eps = float(np.finfo(dt).eps)
        err = (8 if x == "z" else 2) * eps           # TypeUtilities<T>::error, util_types.h:40,:62
        for uplo in "LU":
            for (m, mb) in REF_SIZES:
                def idx(Pn, r, s_):
                    tiles = [L.dlafb_dist_global_tile_from_local(l, Pn, r, s_)
                             for l in range(L.dlafb_dist_local_nr_tiles(m, mb, Pn, r, s_))]
                    return np.concatenate([np.arange(g * mb, min(m, (g + 1) * mb)) for g in tiles] +
                                          [np.arange(0)]).astype(int)
                rows, cols = idx(P, grid.pr, src[0]), idx(Q, grid.pc, src[1])
                a, t = analytic_pair(m, uplo, dt, rows, cols)
                ld = max(1, len(rows)) + 3
                host = np.full((max(1, len(cols)), ld), -7.7, dtype=dt)
                host[:len(cols), :len(rows)] = a.T
                plan = pkg.CholeskyPlan(L, grid, x, uplo, m, mb, src)
                dev = torch.from_numpy(host.reshape(-1).copy()).cuda()
                plan.run(dev.data_ptr(), ld, stream)
                info = plan.info()
                torch.cuda.synchronize()
                out = dev.cpu().numpy().reshape(host.shape)
                plan.close()
                got = out[:len(cols), :len(rows)].T
                tol = 4 * (m + 1) * err
                diff = np.abs(got - t)
                with np.errstate(divide="ignore", invalid="ignore"):
                    bad = int(np.count_nonzero(~((diff <= tol) | (diff / np.abs(t) <= tol)))) if got.size else 0
                pad_ok = bool(np.all(out[:len(cols), len(rows):] == dt(-7.7)))
                cases += 1
                failures += int(info != 0 or bad > 0 or not pad_ok)
    flag = torch.tensor([failures], device="cuda", dtype=torch.int64)
    import torch.distributed as dist
    dist.all_reduce(flag)
    return cases, int(flag.item())


def independent_residual(a0, lfac, ld, rows, cols, n, nb, P, Q, pr, pc, world, complex_):
    """max|A - L L^H| / max|A| over the referenced triangle of a factor distributed 2D block-cyclically (source rank
    (0, 0)), computed WITHOUT the product: block columns of L are assembled on every rank with torch.distributed
    all_gather (own index arithmetic below), the products are torch.matmul. a0 (this rank's block of the original
    matrix) is overwritten with the residual."""
    import torch
    import torch.distributed as dist
    dev = a0.device

    def gidx(nloc_tiles, Pn, r):  # global element indices of the local tiles of rank coordinate r
        if nloc_tiles == 0:
            return torch.zeros(0, dtype=torch.long, device=dev)
        g = torch.arange(nloc_tiles, device=dev) * Pn + r
        e = (g[:, None] * nb + torch.arange(nb, device=dev)[None, :]).reshape(-1)
        return e[e < n]
    nt = -(-n // nb)
    ntl = lambda Pn, r: len(range(r, nt, Pn))
    ridx = [gidx(ntl(P, r), P, r) for r in range(P)]
    cidx = [gidx(ntl(Q, c), Q, c) for c in range(Q)]
    assert len(ridx[pr]) == rows and len(cidx[pc]) == cols
    A = torch.as_strided(a0, (rows, cols), (1, ld))
    Lm = torch.as_strided(lfac, (rows, cols), (1, ld))
    # single precision: the checker works in double, so that its own rounding (sqrt(n) eps32 in a float matmul) does not
    # hide the factorization's
    wide = {torch.float32: torch.float64, torch.complex64: torch.complex128}.get(a0.dtype)
    if wide is not None:
        A = A.to(wide)

    def tri_max(M):  # max |M_ij| over global row >= global column, in column chunks (bounded temporaries)
        v = 0.0
        for j0 in range(0, cols, 2048):
            j1 = min(cols, j0 + 2048)
            blk = M[:, j0:j1].abs()
            blk.masked_fill_(ridx[pr][:, None] < cidx[pc][None, j0:j1], 0.0)
            if blk.numel():
                v = max(v, blk.max().item())
            del blk
        return v
    amax = torch.tensor([tri_max(A)], device=dev, dtype=torch.float64)
    gw = 8  # tile columns per block column
    rmax = max(len(x) for x in ridx)
    for g0 in range(0, nt, gw):
        g1 = min(nt, g0 + gw)
        c_lo, c_hi = g0 * nb, min(n, g1 * nb)
        w = c_hi - c_lo
        wmax = max(int(((cidx[c] >= c_lo) & (cidx[c] < c_hi)).sum().item()) for c in range(Q))
        sel = (cidx[pc] >= c_lo) & (cidx[pc] < c_hi)
        piece = torch.zeros(rmax, wmax, dtype=a0.dtype, device=dev)
        mine = Lm[:, sel]
        piece[:rows, :mine.shape[1]] = mine
        allp = torch.empty(world, rmax, wmax, dtype=a0.dtype, device=dev)
        dist.all_gather_into_tensor(allp.view(-1), piece.view(-1))
        Lk = torch.zeros(n, w, dtype=A.dtype, device=dev)
        for r in range(world):
            rp, rq = r // Q, r % Q       # row-major grid (bench.py creates the grid with ordering 'R')
            cs = cidx[rq][(cidx[rq] >= c_lo) & (cidx[rq] < c_hi)] - c_lo
            if len(cs) and len(ridx[rp]):
                Lk[ridx[rp][:, None], cs[None, :]] = allp[r, :len(ridx[rp]), :len(cs)].to(Lk.dtype)
        Lk.tril_(-c_lo)                  # the factor is lower triangular: global row >= global column
        # rows / columns of this rank at or below the block column's first row
        r0 = int(torch.searchsorted(ridx[pr], c_lo).item())
        q0 = int(torch.searchsorted(cidx[pc], c_lo).item())
        if r0 >= rows or q0 >= cols:
            continue
        X = Lk[ridx[pr][r0:]]
        step = 4096
        for j0 in range(q0, cols, step):
            j1 = min(cols, j0 + step)
            Y = Lk[cidx[pc][j0:j1]]
            A[r0:, j0:j1] -= X @ (Y.conj().T if complex_ else Y.T)
        del Lk, allp, piece
    dmax = torch.tensor([tri_max(A)], device=dev, dtype=torch.float64)
    dist.all_reduce(amax, op=dist.ReduceOp.MAX)
    dist.all_reduce(dmax, op=dist.ReduceOp.MAX)
    return amax.item(), dmax.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", default="headline", choices=["headline", "zpotrf", "float", "weak"],
                    help="BASELINE.json configuration: headline (configs[1] / [2]), zpotrf (configs[3]), float / weak (configs[4])")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed dist_parity leg (N > 1)")
    ap.add_argument("--n", "--size", dest="n", type=int, default=None)
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
    x = wl["type"]
    cplx = x in "cz"
    tdt = {"s": torch.float32, "d": torch.float64, "c": torch.complex64, "z": torch.complex128}[x]
    eps = {"s": 1.1920928955078125e-07, "d": 2.220446049250313e-16}["s" if x in "sc" else "d"]
    gen_fn = getattr(L, f"dlafb_gen_hpd_{x}")
    plan = pkg.CholeskyPlan(L, grid, x, "L", n, nb)
    rows, cols = plan.local_rows, plan.local_cols
    ld = max(1, (rows + 63) // 64 * 64)
    a = torch.empty(ld * max(1, cols) + 64, dtype=tdt, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    row_first = L.dlafb_dist_global_tile_from_local(0, P, grid.pr, 0)
    col_first = L.dlafb_dist_global_tile_from_local(0, Q, grid.pc, 0)

    def gen(seed, into=None):
        pkg.check(gen_fn(stream, n, nb, rows, cols, row_first, P, col_first, Q, (a if into is None else into).data_ptr(),
                         ld, seed), "gen")

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
    gflops = flops_cholesky(n, cplx) / (ms * 1e-3) / 1e9

    # ---- residual of the last factorization on the device, distributed like the matrix (any number of ranks):
    #      max|A - L L^T| / max|A| <= n*eps, miniapp_cholesky.cpp:408-442 ----
    check = None
    engine_check = None
    if not args.no_check:
        try:
            a0buf = torch.empty_like(a)
            gen(100 + args.steps - 1, a0buf)
            amax_g, dmax_g = plan.residual(a0buf.data_ptr(), ld, a.data_ptr(), ld, stream)
            engine_check = {"max_norm_ratio": dmax_g / amax_g, "threshold_n_eps": n * eps,
                            "pass": bool(dmax_g / amax_g <= n * eps), "info": info,
                            "by": "dlafb_cholesky_plan_residual (distributed, on the device)"}
            del a0buf
        except Exception as ex:
            engine_check = {"error": str(ex)[:200]}

    # ---- N > 1: untimed parity leg. (a) the reference's analytical distributed cases on this grid, element-wise;
    #      (b) the residual of the timed factor by an independent checker (torch.distributed + torch.matmul) ----
    dist_parity = None
    if world > 1 and not args.no_check and not args.no_parity:
        dist_parity = {}
        try:
            a0buf = torch.empty_like(a)
            gen(100 + args.steps - 1, a0buf)
            torch.cuda.synchronize()
            amax_i, dmax_i = independent_residual(a0buf, a, ld, rows, cols, n, nb, P, Q, grid.pr, grid.pc, world, cplx)
            del a0buf
            ratio = dmax_i / amax_i
            check = {"max_norm_ratio": ratio, "threshold_n_eps": n * eps, "pass": bool(ratio <= n * eps), "info": info,
                     "by": "independent: torch.distributed all_gather of block columns + torch.matmul"}
            dist_parity["independent_max_norm_ratio"] = ratio
            if engine_check and "max_norm_ratio" in engine_check:
                e_ = engine_check["max_norm_ratio"]
                dist_parity["engine_max_norm_ratio"] = e_
                dist_parity["residuals_agree_within_2x"] = bool(max(e_, ratio) <= 2 * min(e_, ratio) or
                                                                 abs(e_ - ratio) <= 4 * eps)
        except Exception as ex:
            dist_parity["independent_error"] = str(ex)[:200]
            check = engine_check
        try:
            cases, fails = dist_parity_cases(pkg, L, grid, rank, world, P, Q, None)
            dist_parity.update({"cases": cases, "failures": fails,
                                "what": "test_cholesky.cpp:54-101 sizes x uplo x {d, z}, source rank (max(0,P-1), "
                                        "min(1,Q-1)), element-wise vs the closed form at 4(m+1)*err, every rank"})
        except Exception as ex:
            dist_parity["cases_error"] = str(ex)[:200]

    # ---- the same residual with torch.matmul as an independent checker (1 GPU: whole matrix in one place) ----
    if world == 1 and not args.no_check:
        try:
            l_ = torch.as_strided(a, (n, n), (1, ld)).tril()
            gen(100 + args.steps - 1)
            a0 = torch.as_strided(a, (n, n), (1, ld))
            amax = a0.abs().max().item()
            ainf = a0.abs().sum(dim=1).max().item()
            blk, dmax, rowsum = (4096 if cplx else 8192), 0.0, torch.zeros(n, dtype=torch.float64, device="cuda")
            for j0 in range(0, n, blk):          # checker arithmetic: torch.matmul (not the product path)
                j1 = min(n, j0 + blk)
                lj = l_[j0:j1, :]
                d = (a0[:, j0:j1] - l_ @ (lj.conj().T if cplx else lj.T)).abs_()
                low = d.tril(-j0)                 # referenced triangle: global row >= global column
                dmax = max(dmax, low.max().item())
                rowsum += low.sum(dim=1)
                rowsum[j0:j1] += d.tril(-j0 - 1).sum(dim=0)   # mirror of the strictly lower part
                del d, low
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
            host = torch.empty(ld * max(1, cols), dtype=tdt).pin_memory()
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
                getattr(L, f"dlaf_p{x}potrf")(b"L", n, host.data_ptr(), 1, 1, desca, ctypes.byref(info_c))
                dt = (time.perf_counter() - t0) * 1e3
                grid.barrier()
                if i > 0:
                    et.append(grid.allreduce_max(dt))
            ems = sum(et) / len(et)
            # only the referenced (lower) triangle travels, whole tile columns from the diagonal tile down, all ranks
            nt_ = -(-n // nb)
            nbytes = a.element_size() * sum(min(nb, n - gj * nb) * (n - gj * nb) for gj in range(nt_))
            e2e = {"value": round(flops_cholesky(n, cplx) / (ems * 1e-3) / 1e9, 1), "unit": "GFLOP/s",
                   "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes, "ms_per_step": round(ems, 2),
                   "steps": esteps, "api": f"dlaf_p{x}potrf (host pointer, pinned; lower triangle only, D2H of finished column panels overlapped)", "info": info_c.value}
            L.dlaf_free_grid(ctx)
            L.dlaf_finalize()
        except Exception as ex:
            e2e = {"error": str(ex)[:200]}

    if rank == 0:
        peak, peak_src = fp64_peak_tflops()
        kern = "gemm_nt_dmma_kernel (bulk trailing update, rank 0)"
        if x in "sc":
            # single precision runs on the FP32 FMA pipe (no FP32 tensor kind; TF32 fails the reference tolerance):
            # nominal 148 SM x 128 FMA x 2 x 1.965 GHz
            peak, peak_src = 74.5, "nominal FP32 FMA: 148 SM x 128 lanes x 2 flop x 1.965 GHz"
            kern = "gemm_simt(_wide)_kernel (bulk trailing update, rank 0)"
        ach = upd_fl / (upd_ms * 1e-3) / 1e12 if upd_ms > 0 else None
        out = {
            "metric": metric_name(x), "value": round(gflops, 1), "unit": "GFLOP/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
            "higher_is_better": True,
            "scaling": "weak" if (world == 1 or wl["config"] == "weak") else "strong", "vs_baseline": None,
            "dtype": DTYPE[x], "data": "synthetic", "config": wl,
            ("frac_of_fp64_tensor_peak" if x in "dz" else "frac_of_fp32_fma_peak"): round(gflops / 1e3 / (peak * world), 4),
            ("fp64_tensor_peak_tflops_per_gpu" if x in "dz" else "fp32_fma_peak_tflops_per_gpu"): peak,
            "peak_source": peak_src,
            "roofline": {"bound": "tensor" if x in "dz" else "fma", "kernel": kern,
                         "achieved": round(ach, 2) if ach else None, "peak": peak, "unit": "TFLOP/s",
                         "frac": round(ach / peak, 4) if ach else None, "traffic": ncu_traffic(wl),
                         "launches_timed": upd_n,
                         "note": "peak: " + peak_src + "; achieved = algorithmic flops of the bulk launches / "
                                 "their CUDA-event time, while the high-priority stream shares the SMs; traffic = "
                                 "dram bytes read+written by the step-0 bulk launch of the n=32768 double workload "
                                 "(profiles/r01/ncu_update_kernel.json; null when the kernel source changed since)"},
            "clocks": clocks, "gpu_launches": int(launches), "exchange": exchange, "e2e": e2e, "check": check,
            "engine_check": engine_check, "dist_parity": dist_parity,
            "times_ms": [round(t, 2) for t in times],
        }
        if not args.no_cpu and world == 1 and x == "d":
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
