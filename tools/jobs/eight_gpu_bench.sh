#!/bin/bash
# 8-GPU box: the driver's N=8 bench command at HEAD (cfg-2, phases + dist_parity + the config-5 field inside the line)
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_cfg2_8gpu.json 2> gpurun_out/bench_cfg2_8gpu.err
echo "bench rc=$?"; cat gpurun_out/bench_cfg2_8gpu.json | cut -c1-5000; tail -5 gpurun_out/bench_cfg2_8gpu.err | cut -c1-300
