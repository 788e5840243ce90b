#!/bin/bash
# per-kernel launch times of the dImg path (v3 vs v4 variants) + one full capture of the v4 gather and steps kernels
mkdir -p gpurun_out
for tag in "v3:CXL_DIMG_V=3" "v4_1733:CXL_DIMG_CFG=1733 CXL_DIMG_PRED=0" "v4_1732:CXL_DIMG_CFG=1732 CXL_DIMG_PRED=0" "v4_1733p:CXL_DIMG_CFG=1733 CXL_DIMG_PRED=1"; do
  name=${tag%%:*}; envs=${tag#*:}
  env $envs ncu --metrics gpu__time_duration.sum --clock-control none -k regex:dimg -c 160 --csv \
    --log-file gpurun_out/launch_$name.csv python tools/kbench.py --B 4096 --only dimg --reps 1 > gpurun_out/ncu_$name.log 2>&1
  python - <<PY
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/launch_$name.csv")) if len(r)>10 and r[0].isdigit()]
d=collections.defaultdict(list)
for r in rows:
    d[r[4].split('(')[0][:60]].append(float(r[-1].replace(',','')))
for k,v in d.items(): print("$name",k,len(v),"launches avg",sum(v)/len(v)/1e3,"us")
PY
done
CXL_DIMG_CFG=1733 CXL_DIMG_PRED=0 ncu --set full --clock-control none --import-source on -k regex:dimg4_gather -s 3 -c 1 -o gpurun_out/d4_gather_1733 python tools/kbench.py --B 4096 --only dimg --reps 1 > gpurun_out/ncu_full1.log 2>&1
CXL_DIMG_CFG=1733 CXL_DIMG_PRED=0 ncu --set full --clock-control none --import-source on -k regex:dimg4_steps -s 3 -c 1 -o gpurun_out/d4_steps python tools/kbench.py --B 4096 --only dimg --reps 1 > gpurun_out/ncu_full2.log 2>&1
CXL_DIMG_V=3 ncu --set full --clock-control none --import-source on -k regex:dimg_v3 -s 3 -c 1 -o gpurun_out/d3_gather python tools/kbench.py --B 4096 --only dimg --reps 1 > gpurun_out/ncu_full3.log 2>&1
ls -la gpurun_out/*.ncu-rep
