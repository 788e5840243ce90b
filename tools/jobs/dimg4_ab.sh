#!/bin/bash
# A/B of the dImg gather variants (run under gpurun): correctness of the default first, then kbench timings.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "backward or dimg or zero_text" > gpurun_out/dimg4_tests.log 2>&1
tail -3 gpurun_out/dimg4_tests.log
CXL_DIMG_CFG=2522 CXL_DIMG_DEPTH=1 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "backward or dimg or zero_text" > gpurun_out/dimg4_tests2.log 2>&1
tail -3 gpurun_out/dimg4_tests2.log
{
echo "== v3"; CXL_DIMG_V=3 python tools/kbench.py --B 4096 --only dimg
for cfg in 1733 1732 2522 1343; do for depth in 2 1; do
echo "== v4 cfg=$cfg depth=$depth pred=0"; CXL_DIMG_CFG=$cfg CXL_DIMG_DEPTH=$depth python tools/kbench.py --B 4096 --only dimg
done; done
echo "== v4 cfg=2522 depth=2 pred=1"; CXL_DIMG_CFG=2522 CXL_DIMG_PRED=1 python tools/kbench.py --B 4096 --only dimg
echo "== v4 cfg=1732 depth=2 pred=1"; CXL_DIMG_CFG=1732 CXL_DIMG_PRED=1 python tools/kbench.py --B 4096 --only dimg
} > gpurun_out/dimg4_ab.log 2>&1
grep -E "==|dImg" gpurun_out/dimg4_ab.log
CXL_DIMG_CFG=2522 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:dimg -c 40 --csv --log-file gpurun_out/launch_2522.csv python tools/kbench.py --B 4096 --only dimg --reps 1 > /dev/null 2>&1
python - <<PY
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/launch_2522.csv")) if len(r)>10 and r[0].isdigit()]
d=collections.defaultdict(list)
for r in rows: d[r[4].split('(')[0][:60]].append(float(r[-1].replace(',','')))
for k,v in d.items(): print(k,len(v),"launches avg",sum(v)/len(v)/1e3,"us")
PY
