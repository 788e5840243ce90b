#!/bin/bash
# 1 GPU, end of round: whole GPU suite with -x, smoke, then the ncu launch list of a short bench command
# (run once without ncu first; a number printed under ncu is never a bench value)
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
grep -E "^FAILED|passed|failed|^rc=|Error" gpurun_out/gpu_tests.log | cut -c1-250 | tail -30
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-reference --no-phases --k1-reps 1 --e2e-steps 1"
$CMD > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; rc=$?; echo "bench short rc=$rc"
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_cfg2.csv $CMD > gpurun_out/ncu_bench_stdout.log 2>&1
  echo "ncu rc=$?"
  python tools/ncu_launch_summary.py gpurun_out/launches_cfg2.csv "ncu --metrics gpu__time_duration.sum --clock-control none  $CMD   (cfg2, 1 B200, HEAD at the end of round 2)" > gpurun_out/launches_cfg2_summary.txt
  head -14 gpurun_out/launches_cfg2_summary.txt
fi
