"""Golden vectors for the multi-rank semantics, produced by EXECUTING THE REFERENCE under 2 gloo ranks.

Run once in the build container (where /root/reference exists):  python tests/golden/make_golden_dist.py
Each rank loads /root/reference/src/colxlip/loss.py by path and runs the reference's own ClipLoss /
ColClipLoss on its slice of a seeded global batch; losses and local gradients of every rank are committed
as tests/golden/dist_w2.npz.  The only intervention: gradients entering torch.distributed.nn.all_gather's
backward are made contiguous (torch 2.11's gloo emulation of that backward mis-handles strided grads,
SURVEY 5.8) -- the reference code itself is untouched.
"""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
# python tests/golden/make_golden_dist.py [WORLD]   (2 -> dist_w2.npz, the historical file; 4 / 8 -> dist_w4 / dist_w8.npz)
WORLD = int(sys.argv[1]) if len(sys.argv) > 1 and __name__ == "__main__" else int(os.environ.get("CXL_GOLDEN_WORLD", "2"))
B_LOC = {2: 4, 4: 3, 8: 2}.get(WORLD, 2)
P, T, D, SEED, SCALE, ALPHA = 6, 5, 16, 9 + (WORLD if WORLD != 2 else 0), 12.0, 0.4


def _worker(rank, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    torch.set_num_threads(1)
    import torch.distributed.nn as dnn
    from tests.golden.make_golden import load_ref
    from oracle.colxlip_oracle import synth_inputs
    orig = dnn.all_gather

    def all_gather_contig(t, *a, **k):
        outs = orig(t, *a, **k)
        for o in outs:
            o.register_hook(lambda g: g.contiguous())
        return outs

    dnn.all_gather = all_gather_contig
    torch.distributed.nn.all_gather = all_gather_contig
    ref = load_ref()
    feats = synth_inputs(B_LOC * WORLD, P, T, D, seed=SEED, pad_text=True)
    sl = slice(rank * B_LOC, (rank + 1) * B_LOC)
    out = {}
    try:
        for local_loss in (False, True):
            for gwg in (False, True):
                xs = [t[sl].clone().requires_grad_() for t in feats[:2]]
                ls = torch.tensor(SCALE, requires_grad=True)
                crit = ref.ClipLoss(local_loss=local_loss, gather_with_grad=gwg, rank=rank, world_size=WORLD)
                loss = crit(*xs, ls)
                loss.backward()
                key = f"clip_l{int(local_loss)}_g{int(gwg)}_r{rank}"
                out[key + "_loss"] = loss.item()
                out[key + "_g_image_features"] = xs[0].grad.numpy().copy()
                out[key + "_g_text_features"] = xs[1].grad.numpy().copy()
                out[key + "_g_logit_scale"] = ls.grad.item()
        for gwg in (False, True):
            xs = [t[sl].clone().requires_grad_() for t in feats]
            ls = torch.tensor(SCALE, requires_grad=True)
            crit = ref.ColClipLoss(gather_with_grad=gwg, rank=rank, world_size=WORLD, alpha=ALPHA)
            res = crit(*xs, ls, output_dict=True)
            res["total_loss"].backward()
            key = f"colclip_g{int(gwg)}_r{rank}"
            for k, v in res.items():
                out[key + "_" + k] = v.item()
            for x, n in zip(xs, ("image_features", "text_features", "token_image_features", "token_text_features")):
                out[key + "_g_" + n] = x.grad.numpy().copy()
            out[key + "_g_logit_scale"] = ls.grad.item()
        # public gather_features / ColClipLoss.get_logits at world_size 2 (loss.py:48-92, 223-269): the gathered
        # tensors, the four logit matrices, and the local gradients of a seeded linear functional of them
        gw = torch.Generator().manual_seed(SEED + 100)
        w_img = torch.randn(B_LOC * WORLD, D, generator=gw)
        w_txt = torch.randn(B_LOC * WORLD, D, generator=gw)
        w_lpi = torch.randn(B_LOC * WORLD, B_LOC * WORLD, generator=gw)
        w_lpt = torch.randn(B_LOC * WORLD, B_LOC * WORLD, generator=gw)
        for gwg in (False, True):
            xs = [t[sl].clone().requires_grad_() for t in feats[:2]]
            ai, at = ref.gather_features(xs[0], xs[1], local_loss=False, gather_with_grad=gwg, rank=rank,
                                         world_size=WORLD)
            ((ai * w_img).sum() + (at * w_txt).sum()).backward()
            key = f"gather_g{int(gwg)}_r{rank}"
            out[key + "_all_image_features"] = ai.detach().numpy().copy()
            out[key + "_all_text_features"] = at.detach().numpy().copy()
            out[key + "_g_image_features"] = xs[0].grad.numpy().copy()
            out[key + "_g_text_features"] = xs[1].grad.numpy().copy()
            xs = [t[sl].clone().requires_grad_() for t in feats]
            crit = ref.ColClipLoss(gather_with_grad=gwg, rank=rank, world_size=WORLD, alpha=ALPHA)
            lg = crit.get_logits(*xs, torch.tensor(SCALE))
            ((lg["logits_per_image"] * w_lpi).sum() + (lg["logits_per_text_token"] * w_lpt).sum()).backward()
            key = f"logits_g{int(gwg)}_r{rank}"
            for k, v in lg.items():
                out[key + "_" + k] = v.detach().numpy().copy()
            for x, n in zip(xs, ("image_features", "text_features", "token_image_features", "token_text_features")):
                out[key + "_g_" + n] = x.grad.numpy().copy()
        q.put(out)
    finally:
        dist.destroy_process_group()


def main():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    os.environ["CXL_GOLDEN_WORLD"] = str(WORLD)          # spawned workers re-import this module
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, port, q)) for r in range(WORLD)]
    for p in procs:
        p.start()
    merged = {}
    for _ in range(WORLD):
        merged.update(q.get(timeout=300))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    from oracle.colxlip_oracle import synth_inputs
    feats = synth_inputs(B_LOC * WORLD, P, T, D, seed=SEED, pad_text=True)
    np.savez_compressed(os.path.join(HERE, f"dist_w{WORLD}.npz"), world=WORLD, b=B_LOC, P=P, T=T, D=D, seed=SEED,
                        scale=SCALE, alpha=ALPHA, image_features=feats[0].numpy(), text_features=feats[1].numpy(),
                        token_image_features=feats[2].numpy(), token_text_features=feats[3].numpy(), **merged)
    for k in sorted(merged):
        if k.endswith("loss"):
            print(k, merged[k])


if __name__ == "__main__":
    main()
