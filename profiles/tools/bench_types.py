"""Single-GPU timing of the factorization for the other element types of BASELINE.json's configs (z: configs[3],
s: configs[5]); real-equivalent flops = n^3/3 (x4 for complex), miniapp_cholesky.cpp:157-162.

  python profiles/tools/bench_types.py [type n nb] ...     default: z 16384 512, s 32768 512, c 16384 512
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

TORCH = {"s": torch.float32, "d": torch.float64, "c": torch.complex64, "z": torch.complex128}


def run(pkg, L, x, n, nb, steps=3, warmup=2):
    grid = pkg.Grid(L)
    plan = pkg.CholeskyPlan(L, grid, x, "L", n, nb)
    ld = n
    a = torch.empty(ld * n + 64, dtype=TORCH[x], device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    gen = getattr(L, f"dlafb_gen_hpd_{x}")
    times = []
    for i in range(warmup + steps):
        pkg.check(gen(stream, n, nb, n, n, 0, 1, 0, 1, a.data_ptr(), ld, 7 + i), "gen")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.run(a.data_ptr(), ld, stream)
        e1.record()
        torch.cuda.synchronize()
        if i >= warmup:
            times.append(e0.elapsed_time(e1))
    info = plan.info()
    ms = sum(times) / len(times)
    fl = bench.flops_cholesky(n, x in "cz")
    # residual of the last run (miniapp check), blocked to bound memory
    l_ = torch.as_strided(a, (n, n), (1, ld)).tril()
    pkg.check(gen(stream, n, nb, n, n, 0, 1, 0, 1, a.data_ptr(), ld, 7 + warmup + steps - 1), "gen")
    a0 = torch.as_strided(a, (n, n), (1, ld))
    amax = a0.abs().max().item()
    dmax = 0.0
    blk = 4096
    for j0 in range(0, n, blk):
        j1 = min(n, j0 + blk)
        d = a0[j0:, j0:j1] - l_[j0:, :j1] @ l_[j0:j1, :j1].conj().T
        dmax = max(dmax, d.tril(0).abs().max().item())
    eps = torch.finfo(TORCH[x]).eps
    plan.close()
    grid.close()
    return {"type": x, "n": n, "nb": nb, "ms": round(ms, 2), "gflops_real_equiv": round(fl / ms / 1e6, 1), "info": info,
            "residual_max_ratio": dmax / amax, "threshold_n_eps": n * eps, "pass": bool(dmax / amax <= n * eps)}


def main():
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    args = sys.argv[1:]
    cases = [(args[i], int(args[i + 1]), int(args[i + 2])) for i in range(0, len(args), 3)] or \
        [("z", 16384, 512), ("s", 32768, 512), ("c", 16384, 512)]
    for x, n, nb in cases:
        print(json.dumps(run(pkg, L, x, n, nb)), flush=True)


if __name__ == "__main__":
    main()
