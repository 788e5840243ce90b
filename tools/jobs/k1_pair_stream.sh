#!/bin/bash
# streamed-A CTA-pair K1 (P > 256 and / or D too large for a resident text tile): parity first, then A/B against the
# single-CTA kernel (CXL_K1_PAIR_STREAM=0) at the cfg-5 geometry and at P 196 / D 768
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests -x -q -m gpu --timeout 180 > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
grep -E "^FAILED|passed|failed|^rc=|Error|assert|Timeout" gpurun_out/gpu_tests.log | cut -c1-300 | tail -20
{
for v in 1 0 1 0; do
  echo -n "CXL_K1_PAIR_STREAM=$v  P576 D768 B2048: "; CXL_K1_PAIR_STREAM=$v timeout -s KILL 200 python tools/kbench.py --B 2048 --P 576 --D 768 --only k1 --reps 3 2>&1 | grep "K1 maxsim"
done
for v in 1 0; do
  echo -n "CXL_K1_PAIR_STREAM=$v  P196 D768 B4096: "; CXL_K1_PAIR_STREAM=$v timeout -s KILL 200 python tools/kbench.py --B 4096 --P 196 --D 768 --only k1 --reps 3 2>&1 | grep "K1 maxsim"
  echo -n "CXL_K1_PAIR_STREAM=$v  P576 D512 B2048: "; CXL_K1_PAIR_STREAM=$v timeout -s KILL 200 python tools/kbench.py --B 2048 --P 576 --D 512 --only k1 --reps 3 2>&1 | grep "K1 maxsim"
done
} 2>&1 | tee gpurun_out/k1_pair_stream_ab.txt
