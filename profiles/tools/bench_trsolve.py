"""Single-GPU timing + residual of the local triangular solver (include/dlafb_solver.h), the reference's
miniapp_triangular_solver metric: flops = total_ops<T>(m n k / 2, m n k / 2), k = order of A
(miniapp/miniapp_triangular_solver.cpp:143-144, :170); device-resident operands, CUDA events around the solve.

  python profiles/tools/bench_trsolve.py [--both] [type side uplo op m n nb] ...
  --both: every case with DLAFB_TRSOLVE_LOOKAHEAD=0 and =1 (the variable is read when the plan is created)
  default: d L L N 16384 16384 512, d R L C 16384 16384 512, d L L C 16384 16384 512, d L L N 32768 4096 512,
           z L L N 8192 8192 512, s L L N 16384 16384 512, d L L N 32768 32768 512

The residual max|op(A) X - alpha B| / (max|A| max|X| k eps) is formed with torch.matmul as the checker only.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

TORCH = {"s": torch.float32, "d": torch.float64, "c": torch.complex64, "z": torch.complex128}


def run(pkg, L, x, side, uplo, op, m, n, nb, steps=3, warmup=1):
    dt = TORCH[x]
    k = m if side == "L" else n
    g = torch.Generator(device="cuda").manual_seed(5)
    a = torch.randn(k, k, dtype=dt, device="cuda", generator=g) / k ** 0.5
    a += 2 * torch.eye(k, dtype=dt, device="cuda")
    a_cm = a.t().contiguous()            # memory = column-major a, ld = k
    b = torch.randn(m, n, dtype=dt, device="cuda", generator=g)
    b_cm = b.t().contiguous()            # column-major b, ld = m
    work = torch.empty_like(b_cm)
    alpha = complex(-1.2, 0.7) if x in "cz" else -1.2
    plan = pkg.TriangularSolverPlan(L, x, side, uplo, op, "N", m, n, nb)
    stream = torch.cuda.current_stream().cuda_stream
    times = []
    for i in range(warmup + steps):
        work.copy_(b_cm)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.run(alpha, a_cm.data_ptr(), k, work.data_ptr(), m, stream)
        e1.record()
        torch.cuda.synchronize()
        if i >= warmup:
            times.append(e0.elapsed_time(e1))
    ws = plan.workspace_bytes()
    plan.close()
    xs = work.t()                        # m x n view of the solution
    tri = torch.tril(a) if uplo == "L" else torch.triu(a)
    opa = tri if op == "N" else tri.t() if op == "T" else tri.conj().t()
    lhs = opa @ xs if side == "L" else xs @ opa
    res = (lhs - alpha * b).abs().max().item()
    eps = torch.finfo(dt).eps
    ratio = res / (tri.abs().max().item() * xs.abs().max().item() * k * eps)
    ms = min(times)
    flops = (4.0 if x in "cz" else 1.0) * m * n * k
    return {"type": x, "side": side, "uplo": uplo, "op": op, "m": m, "n": n, "nb": nb, "ms_best": round(ms, 3),
            "ms_all": [round(t, 3) for t in times], "gflops_real_equiv": round(flops / ms / 1e6, 1),
            "residual_over_k_eps": ratio, "pass": bool(ratio <= 1.0), "workspace_MiB": round(ws / 2 ** 20, 1)}


def main():
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    args = sys.argv[1:]
    both = "--both" in args
    args = [a for a in args if a != "--both"]
    cases = [(args[i], args[i + 1], args[i + 2], args[i + 3], int(args[i + 4]), int(args[i + 5]), int(args[i + 6]))
             for i in range(0, len(args), 7)] or \
        [("d", "L", "L", "N", 16384, 16384, 512), ("d", "R", "L", "C", 16384, 16384, 512),
         ("d", "L", "L", "C", 16384, 16384, 512), ("d", "L", "L", "N", 32768, 4096, 512),
         ("z", "L", "L", "N", 8192, 8192, 512), ("s", "L", "L", "N", 16384, 16384, 512),
         ("d", "L", "L", "N", 32768, 32768, 512)]
    for c in cases:
        for la in (("0", "1") if both else (os.environ.get("DLAFB_TRSOLVE_LOOKAHEAD"),)):
            if la is not None:
                os.environ["DLAFB_TRSOLVE_LOOKAHEAD"] = la
            r = run(pkg, L, *c)
            r["lookahead"] = la if la is not None else "default"
            print(json.dumps(r), flush=True)
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
