#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_bytes.sum,smsp__inst_executed.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,sm__cycles_elapsed.max --clock-control none -k regex:dimg4 --csv --log-file gpurun_out/dimg_cfg5.csv python tools/kbench.py --B 2048 --P 576 --D 768 --only dimg --reps 1 > gpurun_out/dimg_cfg5_stdout.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/dimg_cfg5.csv", errors="replace")) if len(r)>10 and r[0].isdigit()]
agg=collections.defaultdict(lambda: collections.defaultdict(float)); n=collections.Counter()
for r in rows:
    agg[r[4][:60]][r[-3]]+=float(r[-1].replace(',',''))
    if r[-3]=="gpu__time_duration.sum": n[r[4][:60]]+=1
for k,v in agg.items(): print(k, n[k], {m: f"{x:.4g}" for m,x in v.items()})
PY
