#!/bin/bash
# K1 pair kernel: L2 eviction hints on the TMA loads (bit0) and streaming argmax stores (bit1), A/B interleaved
mkdir -p gpurun_out
{
for cfg in "192 0" "192 1" "192 3" "384 3" "192 0" "192 3" "384 1" "768 3"; do
  set -- $cfg
  echo -n "L2HINT=$2 "; CXL_K1_IMG_BLOCK=$1 CXL_K1_L2HINT=$2 python tools/k1_block_probe.py --b 16384 --bi 32768 --reps 2 2>&1 | tail -1
done
for cfg in "192 0" "192 3" "384 3"; do
  set -- $cfg
  echo -n "L2HINT=$2 "; CXL_K1_IMG_BLOCK=$1 CXL_K1_L2HINT=$2 python tools/k1_block_probe.py 2>&1 | tail -1
  CXL_K1_IMG_BLOCK=$1 CXL_K1_L2HINT=$2 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:maxsim_fwd_tc2 -c 1 --csv --log-file gpurun_out/k1_hint_$1_$2.csv python tools/k1_block_probe.py --reps 1 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/k1_hint_$1_$2.csv")) if len(r)>10 and r[0].isdigit()]
print("   ncu:", {r[-3]: float(r[-1].replace(',','')) for r in rows})
PY
done
} 2>&1 | tee gpurun_out/k1_l2hint_ab.txt
CXL_K1_L2HINT=3 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_scale.py -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/gpu_tests_l2hint3.log
