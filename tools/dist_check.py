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
# grad_rows (accumulation slice) over NCCL: the live slice's token gradients equal the full backward's, zeros elsewhere
lo, hi = 16, 20         # lo * T = 1232 is a multiple of 16: the TMA-staged slice kernel, same bits as the full call
for gwg in (False, True):
    res = []
    for rows in (None, (lo, hi)):
        xs = [t[rank * b:(rank + 1) * b].clone().cuda().requires_grad_() for t in inputs]
        ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
        crit = ColClipLoss(alpha=0.5, rank=rank, world_size=world, gather_with_grad=gwg)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = crit(*xs, ls, **({"grad_rows": rows} if rows else {}))
        loss.backward()
        res.append([x.grad.float() for x in xs])
    for full, part, n in zip(res[0], res[1], names):
        err = (part[lo:hi] - full[lo:hi]).abs().max().item()
        if err > 2e-5 * full.abs().max().item():
            ok = False
            print(f"rank {rank} gwg={gwg} grad_rows {n}: err {err:.3e}")
        outside = torch.cat([part[:lo].reshape(-1), part[hi:].reshape(-1)])
        if n.startswith("token") and outside.numel() and outside.abs().max().item() != 0:
            ok = False
            print(f"rank {rank} gwg={gwg} grad_rows {n}: non-zero outside the slice")

# ClipLoss over NCCL against the REFERENCE's own 2-rank run (tests/golden/dist_w2.npz), world == 2 only
if world == 2:
    import numpy as np
    from colxlip_b200 import ClipLoss
    z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "dist_w2.npz"))
    gb = int(z["b"])
    for local_loss in (0, 1):
        for gwg in (0, 1):
            xs = [torch.from_numpy(z[n])[rank * gb:(rank + 1) * gb].clone().cuda().requires_grad_()
                  for n in ("image_features", "text_features")]
            ls = torch.tensor(float(z["scale"]), device="cuda", requires_grad=True)
            loss = ClipLoss(local_loss=bool(local_loss), gather_with_grad=bool(gwg), rank=rank, world_size=world)(*xs, ls)
            loss.backward()
            key = f"clip_l{local_loss}_g{gwg}_r{rank}"
            bad = abs(loss.item() - float(z[key + "_loss"])) > 1e-5 * abs(float(z[key + "_loss"]))
            for x, n in zip(xs, ("image_features", "text_features")):
                ref = torch.from_numpy(z[f"{key}_g_{n}"])
                bad |= (x.grad.cpu() - ref).abs().max().item() > 1e-5 * ref.abs().max().item() + 1e-9
            bad |= abs(ls.grad.item() - float(z[key + "_g_logit_scale"])) > 2e-5 * max(1e-3, abs(float(z[key + "_g_logit_scale"])))
            if bad:
                ok = False
                print(f"rank {rank} ClipLoss local_loss={local_loss} gwg={gwg}: mismatch vs the reference's rank-{rank} vectors")

flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("DIST_CHECK", "OK" if flag.item() == 1 else "FAILED", f"world={world} B={B}")
dist.destroy_process_group()
sys.exit(0 if flag.item() == 1 else 1)
