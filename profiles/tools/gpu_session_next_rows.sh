#!/bin/bash
# First GPU sessions for the rows that have CPU parity only. Everything lands in
# gpurun_out/next_rows/. Usage, from the repository root:
#
#   gpurun --timeout 2400 -- 'bash profiles/tools/gpu_session_next_rows.sh one'        # 1 GPU: tests, timings, ncu
#   gpurun --gpus 2 --timeout 2400 -- 'bash profiles/tools/gpu_session_next_rows.sh two'
#   gpurun --gpus 8 --timeout 2400 -- 'bash profiles/tools/gpu_session_next_rows.sh eight'
#   gpurun --timeout 2400 -- 'bash profiles/tools/gpu_session_next_rows.sh sanitize'   # compute-sanitizer memcheck, small cases
#
# Every step is wrapped in its own `timeout` so that a hang cannot take the box with it, and continues after a
# failure (the logs say what failed). ncu comes last and only on commands that have just exited 0 without it.
set -u
OUT=gpurun_out/next_rows
mkdir -p "$OUT"
PY=python
step() {  # step NAME SECONDS COMMAND...
  local name=$1 secs=$2
  shift 2
  echo "=== $name: $*" | tee -a "$OUT/summary.txt"
  if [ "$1" = "torchrun_n" ]; then  # `timeout` cannot run a shell function: expand it here
    local n=$2 port=$3
    shift 3
    set -- $PY -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"
  fi
  timeout "$secs" "$@" > "$OUT/$name.log" 2>&1
  local rc=$?
  echo "=== $name rc=$rc" | tee -a "$OUT/summary.txt"
  tail -n 3 "$OUT/$name.log" | tee -a "$OUT/summary.txt"
  return $rc
}
torchrun_n() {  # torchrun_n N PORT SCRIPT ARGS...
  local n=$1 port=$2
  shift 2
  $PY -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"
}

case "${1:-one}" in
one)
  step verified_tests 900 $PY -m pytest tests -x -q -m gpu --deselect tests/test_zz_gen_to_std_gpu.py \
      --deselect tests/test_zz_max_norm_gpu.py --deselect tests/test_zz_redistribute_gpu.py \
      --deselect tests/test_zz_sgemm_wide_gpu.py --deselect tests/test_zz_submatrix_gpu.py \
      --deselect tests/test_zz_triangular_dist_gpu.py
  for t in gen_to_std triangular_dist submatrix redistribute max_norm sgemm_wide; do
    step "tests_$t" 900 $PY -m pytest "tests/test_zz_${t}_gpu.py" -q -m gpu
  done
  step bench_gen_to_std 900 $PY profiles/tools/bench_gen_to_std.py --both
  step bench_dist_rows_1x1 600 $PY profiles/tools/bench_dist_rows.py --n 32768 --nb 512
  step bench_redistribute 300 $PY profiles/tools/bench_redistribute.py --n 32768 --nb-src 64 --nb-dst 512
  step bench_types_default 600 $PY profiles/tools/bench_types.py
  step bench_types_wide 600 env DLAFB_SGEMM_WIDE=1 $PY profiles/tools/bench_types.py
  step bench 900 $PY bench.py --steps 3 --warmup 3
  # ncu: launch lists of the two new kernels' drivers, then one full capture each (only if the plain runs passed)
  if step plain_redistribute 300 $PY profiles/tools/bench_redistribute.py --n 16384 --steps 1; then
    step ncu_copy_rects 600 ncu --set full --clock-control none --import-source on -k regex:copy_rects -c 2 \
        -o "$OUT/copy_rects" $PY profiles/tools/bench_redistribute.py --n 16384 --steps 1
  fi
  ;;
two)
  export DLAFB_DIST_WORLDS=2
  for t in triangular_dist gen_to_std redistribute submatrix; do
    step "tests2_$t" 1500 $PY -m pytest "tests/test_zz_${t}_gpu.py" -q -m gpu -k "distributed or parity"
  done
  step bench_dist_rows_1x2 900 torchrun_n 2 29931 profiles/tools/bench_dist_rows.py --n 32768 --nb 512
  step bench_redistribute_1x2 300 torchrun_n 2 29932 profiles/tools/bench_redistribute.py --n 32768
  ;;
eight)
  export DLAFB_DIST_WORLDS=8 DLAFB_DIST_QUICK=1
  for t in triangular_dist gen_to_std redistribute; do
    step "tests8_$t" 1800 $PY -m pytest "tests/test_zz_${t}_gpu.py" -q -m gpu -k "distributed or parity"
  done
  step bench_dist_rows_2x4 1200 torchrun_n 8 29941 profiles/tools/bench_dist_rows.py --n 65536 --nb 512
  step bench_redistribute_2x4 300 torchrun_n 8 29942 profiles/tools/bench_redistribute.py --n 65536
  step bench8 900 torchrun_n 8 29943 bench.py --gpus 8 --steps 3 --warmup 3
  # the other BASELINE.json configurations at 8 GPUs: complex double, float (default and wide kernel), weak scaling
  step config3_z 900 torchrun_n 8 29944 profiles/tools/bench_dist_types.py --type z --n 40960 --nb 512
  step config4_s 900 torchrun_n 8 29945 profiles/tools/bench_dist_types.py --type s --n 65536 --nb 512
  step config4_s_wide 900 env DLAFB_SGEMM_WIDE=1 $PY -m torch.distributed.run --nnodes=1 --nproc-per-node 8 \
      --master-addr 127.0.0.1 --master-port 29946 profiles/tools/bench_dist_types.py --type s --n 65536 --nb 512
  step config4_weak8 1200 torchrun_n 8 29947 profiles/tools/bench_dist_types.py --type d --n 92160 --nb 1024
  step config4_weak8_ragged 1200 torchrun_n 8 29948 profiles/tools/bench_dist_types.py --type d --n 92682 --nb 1024 --no-check
  ;;
sanitize)
  # one tool per gpurun call (B200_PROFILING.md): memcheck on the smallest cases of the new code paths, and only after
  # the same commands have exited 0 without it
  T="tests/test_zz_redistribute_gpu.py::test_single_rank_is_a_bit_exact_copy tests/test_zz_gen_to_std_gpu.py::test_reference_known_answer tests/test_zz_triangular_dist_gpu.py::test_single_rank_grid_matches_reference_known_answer tests/test_zz_sgemm_wide_gpu.py"
  if step sanitize_plain 900 $PY -m pytest $T -q -m gpu -x; then
    step sanitize_memcheck 1800 compute-sanitizer --tool memcheck --target-processes all --error-exitcode 99 $PY -m pytest $T -q -m gpu -x
  fi
  ;;
*)
  echo "usage: $0 one|two|eight|sanitize" >&2
  exit 2
  ;;
esac
echo "=== done" | tee -a "$OUT/summary.txt"
