#!/bin/bash
# one rank of the 8-GPU cfg-5 step on one GPU (collectives stubbed): memory and kernel time at HEAD
mkdir -p gpurun_out
timeout -s KILL 600 python tools/rank_emulation.py --workload cfg5 --world 8 --steps 1 --warmup 1 > gpurun_out/cfg5_rank_emulation.json 2> gpurun_out/cfg5_rank_emulation.err
echo "rc=$?"; cat gpurun_out/cfg5_rank_emulation.json | cut -c1-1200; tail -3 gpurun_out/cfg5_rank_emulation.err | cut -c1-300
