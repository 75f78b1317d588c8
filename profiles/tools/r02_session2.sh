#!/bin/bash
# Round-2 GPU session 2 (one GPU): whole -m gpu suite with the round's new tests, bench line, float kernel timing + ncu.
set -u
OUT=gpurun_out/r02_s2
mkdir -p "$OUT"
PY=python
step() {
  local name=$1 secs=$2
  shift 2
  echo "=== $name: $*" | tee -a "$OUT/summary.txt"
  timeout "$secs" "$@" > "$OUT/$name.log" 2>&1
  local rc=$?
  echo "=== $name rc=$rc" | tee -a "$OUT/summary.txt"
  tail -n 6 "$OUT/$name.log" | cut -c1-1200 | tee -a "$OUT/summary.txt"
  return $rc
}
step tests_gpu 1500 $PY -m pytest tests -x -q -m gpu
step bench 600 $PY bench.py --steps 5 --warmup 3
step bench_float 600 $PY bench.py --config float --steps 2 --warmup 2 --no-cpu --n 32768
step bench_types 400 $PY profiles/tools/bench_types.py s 32768 512 s 16384 512
if step sgemm_once 300 $PY profiles/tools/bench_types.py s 8192 512; then
  step ncu_sgemm 600 ncu --set full --clock-control none --import-source on -k regex:gemm_simt_wide -s 40 -c 2 \
      -f -o "$OUT/ncu_sgemm" $PY profiles/tools/bench_types.py s 8192 512
fi
echo "=== done" | tee -a "$OUT/summary.txt"
