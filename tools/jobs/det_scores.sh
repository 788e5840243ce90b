#!/bin/bash
# deterministic scores (cxl_maxsim_fwd_ws, default) vs the atomicAdd entry: whole GPU suite on the new default, then K1 A/B
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
grep -E "^FAILED|passed|failed|^rc=|Error" gpurun_out/gpu_tests.log | cut -c1-250 | tail -30
{
for a in 0 1 0 1; do
  echo -n "CXL_S_ATOMIC=$a "; CXL_S_ATOMIC=$a python tools/k1_block_probe.py --b 16384 --bi 32768 --reps 2 2>&1 | tail -1
done
for a in 0 1; do
  echo -n "CXL_S_ATOMIC=$a "; CXL_S_ATOMIC=$a python tools/k1_block_probe.py --b 4096 --bi 4096 --reps 5 2>&1 | tail -1
done
} 2>&1 | tee gpurun_out/k1_det_scores_ab.txt
