"""Multi-GPU parity check (launch with torchrun, one rank per GPU, NCCL):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py
Runs tools/dist_parity.check: the NCCL path against vectors produced by executing the reference under the same number
of (gloo) ranks, and the tensor-core path against the single-GPU path of this package on the global batch."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from tools.dist_parity import check

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
res = check(rank, world, dev)
if res["failed"]:
    print(f"rank {rank}:", *res["failed"], sep="\n  ", file=sys.stderr)
if rank == 0:
    print("DIST_CHECK", "OK" if res["ok"] else "FAILED", json.dumps(res))
dist.destroy_process_group()
sys.exit(0 if res["ok"] else 1)
