#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_guards.py -x -q -m gpu 2>&1 | tail -15 | cut -c1-300 | tee gpurun_out/gpu_tests_guards.log
python tools/sanitize_probe.py 2>&1 | tail -3 | tee gpurun_out/kernel_family_probe.log
