#!/bin/bash
# streamed-A pair K1 vs single-CTA K1 at the cfg-5 per-shard launch shape (8192 texts x 8192 images, ~4 s per launch: sustained clocks)
mkdir -p gpurun_out
{
for v in 1 0 1; do
  echo -n "CXL_K1_PAIR_STREAM=$v  P576 D768 B8192: "; CXL_K1_PAIR_STREAM=$v timeout -s KILL 300 python tools/kbench.py --B 8192 --P 576 --D 768 --only k1 --reps 2 2>&1 | grep "K1 maxsim"
done
nvidia-smi --query-gpu=clocks.sm,power.draw --format=csv,noheader
} 2>&1 | tee gpurun_out/k1_pair_stream_long.txt
