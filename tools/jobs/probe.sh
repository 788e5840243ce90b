#!/bin/bash
# K1 image blocks: time + DRAM / L2 bytes per launch (ncu, one pass) at the per-rank shape of cfg-2 on 8 GPUs
mkdir -p gpurun_out
for blk in 0 96 192 384; do
  CXL_K1_IMG_BLOCK=$blk python tools/k1_block_probe.py 2>&1 | tail -1
  CXL_K1_IMG_BLOCK=$blk ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:maxsim_fwd_tc2 -c 1 --csv --log-file gpurun_out/k1_blk_$blk.csv python tools/k1_block_probe.py --reps 1 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/k1_blk_$blk.csv")) if len(r)>10 and r[0].isdigit()]
print("   ncu:", {r[-3]: float(r[-1].replace(',','')) for r in rows})
PY
done 2>&1 | tee gpurun_out/k1_img_block_ncu.txt
