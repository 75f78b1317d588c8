#!/bin/bash
# quick one-GPU check of the diagonal-tile / panel-solve kernels: parity tests of the tile kernels, then timings;
# extra commands (quoted arguments) run afterwards
set -u
OUT=gpurun_out/r02_quick
mkdir -p "$OUT"
python -m pytest tests/test_tile_kernels.py tests/test_cholesky_gpu.py -x -q -m gpu > "$OUT/tests.log" 2>&1; echo "tests rc=$?" | tee "$OUT/summary.txt"
tail -n 3 "$OUT/tests.log" | tee -a "$OUT/summary.txt"
python profiles/tools/chain_probe.py > "$OUT/chain_probe.log" 2>&1; echo "probe rc=$?" | tee -a "$OUT/summary.txt"
cat "$OUT/chain_probe.log" | tee -a "$OUT/summary.txt"
for a in "$@"; do
  eval "$a" >> "$OUT/extra.log" 2>&1; echo "extra rc=$? : $a" | tee -a "$OUT/summary.txt"; tail -n 4 "$OUT/extra.log" | cut -c1-1500 | tee -a "$OUT/summary.txt"
done
