"""Multi-GPU parity check (launch with torchrun, one rank per GPU, NCCL):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py
Every rank builds the same GLOBAL synthetic batch, feeds its own slice to colxlip_b200.ColClipLoss(world_size=W)
and compares loss / gradients with the single-process oracle on the global batch (SURVEY 5.8 scaling rules:
local grads x1 for gather_with_grad=False, x W for True; logit_scale grad identical on every rank)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from colxlip_b200 import ColClipLoss
from oracle import colxlip_oracle as O

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
b, P, T, D = 20, 196, 77, 256
B = b * world
inputs = O.synth_inputs(B, P, T, D, seed=99, pad_text=True)
rounded = [t.to(torch.bfloat16).float() for t in inputs]
ok = True
for gwg in (False, True):
    xs = [t[rank * b:(rank + 1) * b].clone().cuda().requires_grad_() for t in inputs]
    ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    crit = ColClipLoss(alpha=0.5, rank=rank, world_size=world, gather_with_grad=gwg)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = crit(*xs, ls, output_dict=True)
    out["total_loss"].backward()
    torch.cuda.synchronize()
    r_losses, r_grads = O.loss_and_grads(*rounded, 1 / 0.07, 0.5)
    mult = float(world) if gwg else 1.0
    for k, v in r_losses.items():
        e = abs(out[k].item() - v.item())
        if e > 2e-5 * max(1.0, abs(v.item())):
            ok = False
            print(f"rank {rank} gwg={gwg} {k}: {out[k].item()} vs {v.item()}")
    names = ["image_features", "text_features", "token_image_features", "token_text_features"]
    for x, n in zip(xs, names):
        ref = r_grads[n][rank * b:(rank + 1) * b] * mult
        err = (x.grad.float().cpu() - ref).abs().max().item()
        tol = (1e-2 if n.startswith("token") else 1e-4) * ref.abs().max().item()
        if err > tol:
            ok = False
            print(f"rank {rank} gwg={gwg} grad {n}: err {err:.3e} tol {tol:.3e}")
    e = abs(ls.grad.item() - r_grads["logit_scale"].item())
    if e > 2e-4 * max(1e-3, abs(r_grads["logit_scale"].item())):
        ok = False
        print(f"rank {rank} gwg={gwg} logit_scale grad {ls.grad.item()} vs {r_grads['logit_scale'].item()}")
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("DIST_CHECK", "OK" if flag.item() == 1 else "FAILED", f"world={world} B={B}")
dist.destroy_process_group()
sys.exit(0 if flag.item() == 1 else 1)
