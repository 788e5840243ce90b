#!/bin/bash
# 1 GPU: the allocation-free entry point's test alone in a fresh process, then the K1 image-block A/B at the
# single-GPU shape (interleaved so that clock drift shows), one shard of the 8-GPU pipeline, and the per-rank shape.
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/zz_alone.log
{
for blk in 0 192 384 0 192 96; do
  CXL_K1_IMG_BLOCK=$blk python tools/k1_block_probe.py --b 16384 --bi 32768 --reps 2 2>&1 | tail -1
done
for blk in 0 192; do
  CXL_K1_IMG_BLOCK=$blk python tools/k1_block_probe.py --b 4096 --bi 4096 --reps 5 2>&1 | tail -1
  CXL_K1_IMG_BLOCK=$blk python tools/k1_block_probe.py --b 4096 --bi 32768 --reps 3 2>&1 | tail -1
done
} 2>&1 | tee gpurun_out/k1_img_block_ab_1gpu.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv | tee -a gpurun_out/k1_img_block_ab_1gpu.txt
