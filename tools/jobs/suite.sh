#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
grep -E "^FAILED|passed|failed|^rc=" gpurun_out/gpu_tests.log | cut -c1-250 | tail -30
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
