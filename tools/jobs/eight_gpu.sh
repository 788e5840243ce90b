#!/bin/bash
# 8-GPU box: dist_check at 4 and 8 ranks (NCCL vs reference-executed vectors), then the driver's N=8 bench command
mkdir -p gpurun_out
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py > gpurun_out/dist_check_4gpu.log 2>&1
echo "dist_check 4 rc=$?"; grep DIST_CHECK gpurun_out/dist_check_4gpu.log | cut -c1-500
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 tools/dist_check.py > gpurun_out/dist_check_8gpu.log 2>&1
echo "dist_check 8 rc=$?"; grep DIST_CHECK gpurun_out/dist_check_8gpu.log | cut -c1-500
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 3 --warmup 2 > gpurun_out/bench_cfg2_8gpu.json 2> gpurun_out/bench_cfg2_8gpu.err
echo "bench rc=$?"; cat gpurun_out/bench_cfg2_8gpu.json | cut -c1-3000; tail -5 gpurun_out/bench_cfg2_8gpu.err | cut -c1-300
