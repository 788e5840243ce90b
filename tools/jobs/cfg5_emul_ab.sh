#!/bin/bash
# same box: one rank of the cfg-5 step with the K1 image block at 47 images (40 MB) and at the default (14 MB), cold steps
mkdir -p gpurun_out
for blk in 47 default; do
  if [ $blk = default ]; then unset CXL_K1_IMG_BLOCK; else export CXL_K1_IMG_BLOCK=$blk; fi
  echo -n "CXL_K1_IMG_BLOCK=$blk: "
  timeout -s KILL 300 python tools/rank_emulation.py --workload cfg5 --world 8 --steps 1 --warmup 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k: round(d[k],1) for k in ('fwd_ms','bwd_ms','step_ms','fwd_tflops_rank')})"
done 2>&1 | tee gpurun_out/cfg5_emul_ab.txt
nvidia-smi --query-gpu=clocks.sm,power.draw,temperature.gpu --format=csv,noheader | tee -a gpurun_out/cfg5_emul_ab.txt
