#!/bin/bash
# (torchrun's own parser takes a bare --n after the script name for an abbreviation of --nnodes: the tools accept --size)
# Round-2 multi-GPU session: usage  r02_multi.sh N [quick]   (run under gpurun --gpus N)
# distributed parity scripts at world N, bench.py at N (headline with dist_parity, then the other BASELINE configs),
# the distributed "next" rows (solver, reduction to standard form, redistribution, sub-matrix) and their timings.
set -u
N=${1:-2}
MODE=${2:-full}
OUT=gpurun_out/r02_n$N
mkdir -p "$OUT"
PY=python
HEAD=$(cat .git_head 2>/dev/null || echo unknown)
echo "head=$HEAD world=$N" | tee "$OUT/summary.txt"
PORT=29700
step() {
  local name=$1 secs=$2
  shift 2
  echo "=== $name: $*" | tee -a "$OUT/summary.txt"
  if [ "$1" = "trun" ]; then  # `timeout` cannot run a shell function: expand the torchrun prefix here
    shift
    PORT=$((PORT + 1))
    set -- $PY -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$PORT" "$@"
  fi
  timeout "$secs" "$@" > "$OUT/$name.log" 2>&1
  local rc=$?
  echo "=== $name rc=$rc" | tee -a "$OUT/summary.txt"
  tail -n 3 "$OUT/$name.log" | cut -c1-2500 | tee -a "$OUT/summary.txt"
  return $rc
}
export DLAFB_DIST_WORLDS=$N
if [ "$N" -ge 6 ]; then export DLAFB_DIST_QUICK=1; fi
step dist_gpu_check 900 trun tests/dist_gpu_check.py
step bench_headline 900 trun bench.py --gpus "$N" --steps 3 --warmup 3 --timeline "$OUT/timeline"
case "$MODE" in
full)   # 2 GPUs: everything
  step bench_zpotrf 900 trun bench.py --gpus "$N" --config zpotrf --steps 3 --warmup 2
  step bench_float 900 trun bench.py --gpus "$N" --config float --steps 3 --warmup 2
  step bench_weak 900 trun bench.py --gpus "$N" --config weak --steps 3 --warmup 2
  step bench_weak_ragged 600 trun bench.py --gpus "$N" --config weak --size 46341 --steps 2 --warmup 1 --no-check --no-e2e
  step dist_trsolve 900 trun tests/dist_trsolve_check.py
  step dist_g2s 900 trun tests/dist_g2s_check.py
  step dist_redistribute 900 trun tests/dist_redistribute_check.py
  step dist_submatrix 900 trun tests/dist_submatrix_check.py
  step bench_dist_rows 1200 trun profiles/tools/bench_dist_rows.py --size 32768 --nb 512
  step bench_redistribute 300 trun profiles/tools/bench_redistribute.py --size 32768
  ;;
n4)     # 2 x 2: parity of the distributed rows + their timings + the weak-scaling point
  step dist_trsolve 900 trun tests/dist_trsolve_check.py
  step dist_g2s 900 trun tests/dist_g2s_check.py
  step dist_redistribute 900 trun tests/dist_redistribute_check.py
  step dist_submatrix 900 trun tests/dist_submatrix_check.py
  step bench_weak 900 trun bench.py --gpus "$N" --config weak --steps 2 --warmup 2
  step bench_dist_rows 1200 trun profiles/tools/bench_dist_rows.py --size 65536 --nb 512
  ;;
n8)     # 2 x 4: the BASELINE configurations, the panel-solve variants, the distributed rows' timings
  step bench_zpotrf 900 trun bench.py --gpus "$N" --config zpotrf --steps 3 --warmup 2
  step bench_float 900 trun bench.py --gpus "$N" --config float --steps 3 --warmup 2
  step bench_weak 900 trun bench.py --gpus "$N" --config weak --steps 2 --warmup 2
  step bench_nofused 600 env DLAFB_TRSM_FUSED=0 $PY -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" \
      --master-addr 127.0.0.1 --master-port 29790 bench.py --gpus "$N" --steps 3 --warmup 2 --no-e2e --no-check --timeline "$OUT/timeline_nofused"
  step bench_fused_all 600 env DLAFB_TRSM_FUSED=16 DLAFB_TRSM_FUSED_MAX_M=100000 $PY -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" \
      --master-addr 127.0.0.1 --master-port 29791 bench.py --gpus "$N" --steps 3 --warmup 2 --no-e2e --no-check --timeline "$OUT/timeline_fusedall"
  step bench_weak_ragged 600 trun bench.py --gpus "$N" --config weak --size 92682 --steps 2 --warmup 1 --no-check --no-e2e
  step bench_dist_rows 1200 trun profiles/tools/bench_dist_rows.py --size 65536 --nb 512
  ;;
esac
echo "=== done" | tee -a "$OUT/summary.txt"
