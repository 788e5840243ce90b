#!/bin/bash
# whole GPU suite (default kernels), then the backward tests again with the opt-in dImg variants
mkdir -p gpurun_out
python -c "import torch; f,t=torch.cuda.mem_get_info(); print('mem_get_info free/total GiB', f/2**30, t/2**30, 'bytes', t)" > gpurun_out/meminfo.txt 2>&1
python -m pytest tests -q -m gpu -s > gpurun_out/gpu_tests.log 2>&1
echo "rc=$?" >> gpurun_out/gpu_tests.log
tail -25 gpurun_out/gpu_tests.log | cut -c1-300
CXL_DIMG_V=4 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_scale.py -q -m gpu -k "backward or dimg or zero_text or d768" > gpurun_out/gpu_tests_v4.log 2>&1
tail -3 gpurun_out/gpu_tests_v4.log
cat gpurun_out/meminfo.txt
