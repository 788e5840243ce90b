#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 200 python -m pytest tests/test_gpu_parity_scale.py -x -q -m gpu -k "whole_matrix" --timeout 150 2>&1 | tail -6 | cut -c1-300 | tee gpurun_out/one_test.log
