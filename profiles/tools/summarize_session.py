"""Condenses a session summary.txt (profiles/tools/r02_multi.sh) : one line per step, bench JSON lines reduced to the fields
that matter.   python profiles/tools/summarize_session.py gpurun_out/r02_n2/summary.txt"""
import json
import sys

KEYS = ("metric", "value", "ms_per_step", "n_gpus", "frac_of_fp64_tensor_peak", "frac_of_fp32_fma_peak", "exchange",
        "dist_parity", "times_ms")
for line in open(sys.argv[1]):
    line = line.rstrip()
    if line.startswith("{"):
        try:
            d = json.loads(line)
        except Exception:
            print(line[:400])
            continue
        if "metric" not in d:
            print(line[:600])
            continue
        keep = {k: d[k] for k in KEYS if k in d}
        cfg = d.get("config") or {}
        keep["workload"] = cfg.get("workload")
        keep["roofline_frac"] = (d.get("roofline") or {}).get("frac")
        keep["e2e"] = {k: v for k, v in (d.get("e2e") or {}).items() if k in ("value", "ms_per_step", "error")}
        keep["check"] = d.get("check")
        keep["engine_check"] = {k: v for k, v in (d.get("engine_check") or {}).items() if k in ("max_norm_ratio", "pass", "error")}
        keep["clocks"] = d.get("clocks")
        print(json.dumps(keep))
    else:
        print(line[:400])
