#!/bin/bash
# dTxt with evenly split slab boxes at P > 256: parity (whole suite), then kbench at the cfg-5 geometry
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests -x -q -m gpu --timeout 180 > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
grep -E "^FAILED|passed|failed|^rc=|Error|assert|Timeout" gpurun_out/gpu_tests.log | cut -c1-300 | tail -20
{
echo "# kbench --B 2048 --P 576 --D 768 (before this change: K2 dTxt 80.999 ms, profiles/r4b_kbench_cfg5_geometry_B2048.txt)"
timeout -s KILL 300 python tools/kbench.py --B 2048 --P 576 --D 768 --reps 3 2>&1 | grep "K1\|K2"
echo "# kbench --B 2048 --P 300 --D 512"
timeout -s KILL 300 python tools/kbench.py --B 2048 --P 300 --D 512 --only dtxt --reps 3 2>&1 | grep "K2"
} 2>&1 | tee gpurun_out/dtxt_boxes.txt
