"""Condenses an .ncu-rep (read here, without a GPU: `ncu -i ... --page raw --csv`) into a small JSON: per captured launch
the duration, DRAM bytes, pipe utilisation, issue-slot utilisation, registers, shared-memory bank conflicts and the
stall reasons per issued instruction.   python profiles/tools/ncu_summary.py report.ncu-rep > summary.json"""
import csv
import json
import subprocess
import sys

KEEP = {
    "gpu__time_duration.sum": "duration",
    "sm__cycles_elapsed.max": "sm_cycles",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__occupancy_limit_shared_mem": "occupancy_limit_smem_blocks",
    "smsp__inst_executed.sum": "warp_instructions",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slot_pct",
    "sm__inst_executed_pipe_fp64.sum.pct_of_peak_sustained_active": "fp64_pipe_pct",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64_pipe_cycles_pct",
    "sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active": "fma_pipe_pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "smem_wavefronts",
    "sm__warps_active.avg.per_cycle_active": "warps_active_per_cycle",
    "lts__t_bytes.sum": "l2_bytes",
}


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")][:120]}
        stalls = {}
        for i, h in enumerate(hdr):
            if h in KEEP:
                try:
                    d[KEEP[h]] = float(r[i].replace(",", ""))
                except ValueError:
                    d[KEEP[h]] = r[i]
                if units[i]:
                    d[KEEP[h] + "_unit"] = units[i]
            elif h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    stalls[h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = round(float(r[i]), 3)
                except ValueError:
                    pass
        d["stalls_per_issue"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:6])
        out.append(d)
    json.dump({"report": rep, "launches": out}, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
