#!/bin/bash
# Round-2 GPU session 3 (one GPU): panel timing stamps, small-block / upper timings, ncu of the round's new kernels and
# the launch list of bench.py.
set -u
OUT=gpurun_out/r02_s3
mkdir -p "$OUT"
PY=python
step() {
  local name=$1 secs=$2
  shift 2
  echo "=== $name: $*" | tee -a "$OUT/summary.txt"
  timeout "$secs" "$@" > "$OUT/$name.log" 2>&1
  local rc=$?
  echo "=== $name rc=$rc" | tee -a "$OUT/summary.txt"
  tail -n 8 "$OUT/$name.log" | cut -c1-1500 | tee -a "$OUT/summary.txt"
  return $rc
}
step potf2_timing 120 ./profiles/tools/potf2_bench_timing
step chain_probe 200 $PY profiles/tools/chain_probe.py
step small_blocks 600 $PY profiles/tools/bench_small_blocks.py 16384
step upper_32768 600 ./dla-future_b200/bin/miniapp_cholesky --matrix-size 32768 --block-size 512 --type d --uplo upper --nruns 3 --nwarmups 1 --local --check-result last
step lower_32768 600 ./dla-future_b200/bin/miniapp_cholesky --matrix-size 32768 --block-size 512 --type d --uplo lower --nruns 3 --nwarmups 1 --local --check-result last
step bench 600 $PY bench.py --steps 5 --warmup 3
step bench_fused16_all 300 env DLAFB_TRSM_FUSED=16 DLAFB_TRSM_FUSED_MAX_M=100000 $PY bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-check
step bench_fused16_20k 300 env DLAFB_TRSM_FUSED=16 DLAFB_TRSM_FUSED_MAX_M=20480 $PY bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-check
if step probe_once_potrf 200 $PY profiles/tools/chain_probe.py --once --only potrf; then
  step ncu_potf2 600 ncu --set full --clock-control none --import-source on -k regex:potf2_d128 -s 4 -c 2 \
      -f -o "$OUT/ncu_potf2_d128" $PY profiles/tools/chain_probe.py --once --only potrf
fi
if step probe_once_trsm 200 $PY profiles/tools/chain_probe.py --once --only trsm --rows 4096; then
  step ncu_trsm 600 ncu --set full --clock-control none --import-source on -k regex:trsm_fused -s 1 -c 2 \
      -f -o "$OUT/ncu_trsm_fused" $PY profiles/tools/chain_probe.py --once --only trsm --rows 4096
fi
if step bench_short 300 $PY bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-check --size 16384; then
  step ncu_launches 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv \
      --log-file "$OUT/launches_r02.csv" $PY bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-check --size 16384
fi
echo "=== done" | tee -a "$OUT/summary.txt"
