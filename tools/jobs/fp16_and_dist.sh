#!/bin/bash
# 2-GPU job: fp16 tests + GPU suite on one GPU, then dist_check and a short 2-rank bench over NCCL
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fp16.py -q -m gpu > gpurun_out/gpu_fp16.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_fp16.log
tail -30 gpurun_out/gpu_fp16.log | cut -c1-260
python -m pytest tests -q -m gpu --deselect tests/test_gpu_fp16.py > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
tail -4 gpurun_out/gpu_tests.log | cut -c1-260
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py > gpurun_out/dist_check_2gpu.log 2>&1
echo "dist_check rc=$?"; tail -12 gpurun_out/dist_check_2gpu.log | cut -c1-600
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload mid --steps 2 --warmup 1 > gpurun_out/bench_mid_2gpu.json 2> gpurun_out/bench_mid_2gpu.err
echo "bench rc=$?"; cat gpurun_out/bench_mid_2gpu.json | cut -c1-1500; tail -5 gpurun_out/bench_mid_2gpu.err
