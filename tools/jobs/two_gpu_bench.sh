#!/bin/bash
# 2-GPU box: the driver's N=2 bench command at HEAD
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_cfg2_2gpu.json 2> gpurun_out/bench_cfg2_2gpu.err
echo "bench rc=$?"; cat gpurun_out/bench_cfg2_2gpu.json | cut -c1-4000; tail -5 gpurun_out/bench_cfg2_2gpu.err | cut -c1-300
