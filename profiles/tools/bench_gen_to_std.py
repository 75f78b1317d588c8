"""Single-GPU timing + check of the generalized-to-standard reduction (include/dlafb_gen_to_std.h), the reference's
miniapp_gen_to_std metric: flops = total_ops<T>(n^3 / 2, n^3 / 2) (miniapp/miniapp_gen_to_std.cpp), i.e. n^3 real /
4 n^3 complex; device-resident operands, CUDA events around the reduction.

  python profiles/tools/bench_gen_to_std.py [--both] [type uplo n nb] ...
  --both: every case with DLAFB_G2S_LOOKAHEAD=0 and =1 (the variable is read when the plan is created)
  default: d L 16384 512, d U 16384 512, z L 8192 512, s L 16384 512, d L 32768 512

Check (torch.matmul / triangular solves as the checker only): with C the result, max|L C L^H - A| / (max|A| n eps)
over the stored triangle.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

TORCH = {"s": torch.float32, "d": torch.float64, "c": torch.complex64, "z": torch.complex128}


def run(pkg, L, x, uplo, n, nb, steps=3, warmup=1):
    dt = TORCH[x]
    g = torch.Generator(device="cuda").manual_seed(7)
    a = torch.randn(n, n, dtype=dt, device="cuda", generator=g)
    a = a + a.conj().t()
    f = torch.randn(n, n, dtype=dt, device="cuda", generator=g) / n ** 0.5
    f = torch.tril(f, -1) + 2 * torch.eye(n, dtype=dt, device="cuda")     # lower factor, well conditioned
    fac = f if uplo == "L" else f.conj().t()
    a_cm = a.t().contiguous()            # memory = column-major a, ld = n
    f_cm = fac.t().contiguous()
    work = torch.empty_like(a_cm)
    plan = pkg.GenToStdPlan(L, x, uplo, n, nb)
    stream = torch.cuda.current_stream().cuda_stream
    times = []
    for i in range(warmup + steps):
        work.copy_(a_cm)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.run(work.data_ptr(), n, f_cm.data_ptr(), n, stream)
        e1.record()
        torch.cuda.synchronize()
        if i >= warmup:
            times.append(e0.elapsed_time(e1))
    ws = plan.workspace_bytes()
    plan.close()
    c = work.t()
    tri = torch.tril if uplo == "L" else torch.triu
    c = tri(c)
    c = c + tri(c, -1 if uplo == "L" else 1).conj().t()   # full Hermitian image of the stored triangle
    back = f @ c @ f.conj().t()
    res = tri(back - a).abs().max().item()
    eps = torch.finfo(dt).eps
    ratio = res / (a.abs().max().item() * n * eps)
    ms = min(times)
    flops = (4.0 if x in "cz" else 1.0) * float(n) ** 3
    return {"type": x, "uplo": uplo, "n": n, "nb": nb, "ms_best": round(ms, 3), "ms_all": [round(t, 3) for t in times],
            "gflops_real_equiv": round(flops / ms / 1e6, 1), "residual_over_n_eps": ratio, "pass": bool(ratio <= 10.0),
            "workspace_MiB": round(ws / 2 ** 20, 1)}


def main():
    pkg = bench.load_package()
    L = pkg.lib()
    torch.cuda.set_device(0)
    args = sys.argv[1:]
    both = "--both" in args
    args = [a for a in args if a != "--both"]
    cases = [(args[i], args[i + 1], int(args[i + 2]), int(args[i + 3])) for i in range(0, len(args), 4)] or \
        [("d", "L", 16384, 512), ("d", "U", 16384, 512), ("z", "L", 8192, 512), ("s", "L", 16384, 512),
         ("d", "L", 32768, 512)]
    for c in cases:
        for la in (("0", "1") if both else (os.environ.get("DLAFB_G2S_LOOKAHEAD"),)):
            if la is not None:
                os.environ["DLAFB_G2S_LOOKAHEAD"] = la
            r = run(pkg, L, *c)
            r["lookahead"] = la if la is not None else "default"
            print(json.dumps(r), flush=True)
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
