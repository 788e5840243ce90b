#!/bin/bash
# dTxt row-tile pairs with multicast slabs: bit-identity test first (under a hard timeout), whole suite, then A/B
mkdir -p gpurun_out
timeout -s KILL 150 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k multicast --timeout 100 2>&1 | tail -5 | cut -c1-300 | tee gpurun_out/dtxt_mc_test.log
if ! grep -q "4 passed" gpurun_out/dtxt_mc_test.log; then echo "multicast test did not pass; stopping"; exit 0; fi
timeout -s KILL 400 python -m pytest tests -x -q -m gpu --timeout 180 > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
grep -E "^FAILED|passed|failed|^rc=|Error|assert|Timeout" gpurun_out/gpu_tests.log | cut -c1-300 | tail -20
{
for mc in 0 2 0 2; do
  echo -n "CXL_DTXT_MC=$mc P576 D768 B2048: "; CXL_DTXT_MC=$mc timeout -s KILL 200 python tools/kbench.py --B 2048 --P 576 --D 768 --only dtxt --reps 3 2>&1 | grep "K2 dTxt"
done
for mc in 0 2 0 2; do
  echo -n "CXL_DTXT_MC=$mc P196 D512 B4096: "; CXL_DTXT_MC=$mc timeout -s KILL 200 python tools/kbench.py --B 4096 --P 196 --D 512 --only dtxt --reps 3 2>&1 | grep "K2 dTxt"
done
} 2>&1 | tee gpurun_out/dtxt_mc_ab.txt
