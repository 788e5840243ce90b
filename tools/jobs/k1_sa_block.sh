#!/bin/bash
# streamed-A pair K1 at the cfg-5 per-shard shape: L2 image block size (default 40 MB = 47 images)
mkdir -p gpurun_out
{
for blk in 16 8 24 16 4; do
  if [ $blk -lt 0 ]; then unset CXL_K1_IMG_BLOCK; echo -n "CXL_K1_IMG_BLOCK=(default 47) "; else export CXL_K1_IMG_BLOCK=$blk; echo -n "CXL_K1_IMG_BLOCK=$blk "; fi
  timeout -s KILL 300 python tools/kbench.py --B 8192 --P 576 --D 768 --only k1 --reps 2 2>&1 | grep "K1 maxsim"
done
} 2>&1 | tee gpurun_out/k1_sa_block2.txt
