"""Times the trailing-update contraction kernel alone (dlafb_gemm_nt_d, lower) on one B200."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
from conftest import load_package
L = load_package().lib()
res = {}
for (m, k) in [(4096, 512), (16384, 512), (32256, 512), (16384, 1024)]:
    a = torch.rand(k * m, dtype=torch.float64, device="cuda") - 0.5        # m x k, ld = m
    c = torch.zeros(m * m, dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for lower in (1, 0):
        if lower == 0 and m > 16384: continue
        for _ in range(2):
            assert L.dlafb_gemm_nt_d(s, m, m, k, -1.0, a.data_ptr(), m, a.data_ptr(), m, 1.0, c.data_ptr(), m, lower) == 0
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        reps = 3
        for _ in range(reps):
            L.dlafb_gemm_nt_d(s, m, m, k, -1.0, a.data_ptr(), m, a.data_ptr(), m, 1.0, c.data_ptr(), m, lower)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        nblk = (m // 128) * (m // 128 + 1) // 2 if lower else (m // 128) ** 2
        fl = nblk * 2.0 * 128 * 128 * k
        res[f"m{m}_k{k}_lower{lower}"] = {"ms": round(ms, 3), "tflops": round(fl / ms / 1e9, 2)}
    # correctness spot check on the last C
print(json.dumps(res))
