#!/bin/bash
mkdir -p gpurun_out
# small guard run (1/8 of the per-rank batch), then the real cfg-5 rank: b = 8192 texts vs 65536 images
timeout 300 python tools/rank_emulation.py --workload cfg5 --world 8 --b 1024 > gpurun_out/emul_cfg5_b1024.json 2> gpurun_out/emul_cfg5_b1024.err; echo "rc=$?"; cat gpurun_out/emul_cfg5_b1024.json; tail -3 gpurun_out/emul_cfg5_b1024.err
timeout 900 python tools/rank_emulation.py --workload cfg5 --world 8 > gpurun_out/emul_cfg5_rank.json 2> gpurun_out/emul_cfg5_rank.err; echo "rc=$?"; cat gpurun_out/emul_cfg5_rank.json; tail -5 gpurun_out/emul_cfg5_rank.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
