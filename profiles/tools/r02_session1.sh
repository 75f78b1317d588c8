#!/bin/bash
# Round-2 GPU session 1 (one GPU): first timings of what round 1 left untimed + ncu of the latency kernels.
set -u
OUT=gpurun_out/r02_s1
mkdir -p "$OUT"
PY=python
step() {
  local name=$1 secs=$2
  shift 2
  echo "=== $name: $*" | tee -a "$OUT/summary.txt"
  timeout "$secs" "$@" > "$OUT/$name.log" 2>&1
  local rc=$?
  echo "=== $name rc=$rc" | tee -a "$OUT/summary.txt"
  tail -n 4 "$OUT/$name.log" | cut -c1-600 | tee -a "$OUT/summary.txt"
  return $rc
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/gpu.txt" 2>&1
step chain_probe 300 $PY profiles/tools/chain_probe.py
step bench_types_default 600 $PY profiles/tools/bench_types.py s 32768 512 c 16384 512
step bench_types_wide 600 env DLAFB_SGEMM_WIDE=1 $PY profiles/tools/bench_types.py s 32768 512 c 16384 512
step tests_sgemm_wide 600 $PY -m pytest tests/test_zz_sgemm_wide_gpu.py -q -m gpu
step bench_gen_to_std 600 $PY profiles/tools/bench_gen_to_std.py --both
step bench_dist_rows_1x1 600 $PY profiles/tools/bench_dist_rows.py --n 32768 --nb 512
step bench_redistribute 300 $PY profiles/tools/bench_redistribute.py --n 32768 --nb-src 64 --nb-dst 512
if step probe_once_potrf 200 $PY profiles/tools/chain_probe.py --once --only potrf; then
  step ncu_potf2 600 ncu --set full --clock-control none --import-source on -k regex:potf2_kernel -s 4 -c 2 \
      -f -o "$OUT/ncu_potf2" $PY profiles/tools/chain_probe.py --once --only potrf
fi
if step probe_once_trsm 200 $PY profiles/tools/chain_probe.py --once --only trsm --rows 4096; then
  step ncu_trsm 600 ncu --set full --clock-control none --import-source on -k regex:gemm_nt_dmma -s 6 -c 7 \
      -f -o "$OUT/ncu_trsm" $PY profiles/tools/chain_probe.py --once --only trsm --rows 4096
fi
echo "=== done" | tee -a "$OUT/summary.txt"
