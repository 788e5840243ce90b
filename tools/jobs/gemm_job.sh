#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_widen.py tests/test_gpu_fp16.py -q -m gpu > gpurun_out/gpu_widen_fp16.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_widen_fp16.log
tail -25 gpurun_out/gpu_widen_fp16.log | cut -c1-250
python -m pytest tests -q -m gpu --deselect tests/test_gpu_widen.py --deselect tests/test_gpu_fp16.py > gpurun_out/gpu_tests_rest.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests_rest.log
tail -5 gpurun_out/gpu_tests_rest.log | cut -c1-250
python tools/widen_bench.py > gpurun_out/widen_bench.json 2> gpurun_out/widen_bench.err; cat gpurun_out/widen_bench.json; tail -3 gpurun_out/widen_bench.err
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
