#!/bin/bash
# whole GPU suite (default kernels), then the backward tests again with the opt-in dImg v4
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu -s > gpurun_out/gpu_tests.log 2>&1
echo "rc=$?" >> gpurun_out/gpu_tests.log
tail -25 gpurun_out/gpu_tests.log
CXL_DIMG_V=4 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_scale.py -x -q -m gpu -k "backward or dimg or zero_text or d768" > gpurun_out/gpu_tests_v4.log 2>&1
tail -3 gpurun_out/gpu_tests_v4.log
python tools/kbench.py --B 2048 --P 576 --D 768 > gpurun_out/kbench_cfg5_geom.log 2>&1; cat gpurun_out/kbench_cfg5_geom.log
CXL_DIMG_V=4 python tools/kbench.py --B 2048 --P 576 --D 768 --only dimg 2>&1 | grep dImg
