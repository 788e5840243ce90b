#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-gpu-reference --k1-reps 1 --e2e-steps 1 > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_short.json; tail -3 gpurun_out/bench_short.err
python -c "import torch; print('max mem GiB', torch.cuda.max_memory_allocated()/2**30)" 
