#!/bin/bash
# 1 GPU: whole GPU suite with -x (the allocation-free entry point's test runs last), smoke, sanitizer logs,
# K1 image-block ncu pass, K1 DRAM traffic at the cfg-2 single-GPU shape, then the default bench.
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
grep -E "^FAILED|passed|failed|^rc=" gpurun_out/gpu_tests.log | cut -c1-250 | tail -30
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
bash tools/jobs/sanitize.sh 2>&1 | tee gpurun_out/sanitize_summary.txt
bash tools/jobs/probe.sh
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:maxsim_fwd_tc2 -c 1 --csv --log-file gpurun_out/k1_cfg2_dram.csv python tools/k1_block_probe.py --b 32768 --bi 32768 --reps 1 > gpurun_out/k1_cfg2_ncu_stdout.log 2>&1
tail -3 gpurun_out/k1_cfg2_dram.csv | cut -c1-400
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_cfg2_1gpu.json 2> gpurun_out/bench_cfg2_1gpu.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_cfg2_1gpu.json
