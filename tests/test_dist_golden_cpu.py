"""World-size-2 semantics against vectors produced by the REFERENCE ITSELF under 2 gloo ranks
(tests/golden/make_golden_dist.py -> dist_w2.npz): the oracle's single-process rules, and the host-side
sharding / collectives / gradient scaling of colxlip_b200.loss.ClipLoss and ColClipLoss (stand-in kernel
backend on CPU, gloo)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from colxlip_b200 import loss as L
from oracle import colxlip_oracle as O
from tests.ref_kernels import TorchRefKernels

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "dist_w2.npz")
GOLDEN_W = {w: os.path.join(os.path.dirname(__file__), "golden", f"dist_w{w}.npz") for w in (2, 4, 8)}
NAMES = ("image_features", "text_features", "token_image_features", "token_text_features")


def _close(a, ref, rel=1e-5):
    a, ref = torch.as_tensor(a), torch.as_tensor(ref)
    return (a - ref).abs().max().item() <= rel * ref.abs().max().item() + 1e-8


# ---- the oracle's rules vs the reference's multi-rank run ---------------------------------------------
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("gwg", [0, 1])
def test_oracle_colclip_scaling_rule_matches_reference_ranks(gwg, world):
    z = np.load(GOLDEN_W[world])
    W, b = int(z["world"]), int(z["b"])
    feats = [torch.from_numpy(z[n]) for n in NAMES]
    losses, grads = O.loss_and_grads(*feats, float(z["scale"]), float(z["alpha"]))
    mult = O.gather_scaling(bool(gwg), W)
    for r in range(W):
        key = f"colclip_g{gwg}_r{r}"
        for k, v in losses.items():
            assert abs(v.item() - float(z[f"{key}_{k}"])) < 1e-5
        sl = slice(r * b, (r + 1) * b)
        for n in NAMES:
            assert _close(grads[n][sl] * mult, z[f"{key}_g_{n}"]), (r, n)
        assert abs(grads["logit_scale"].item() - float(z[f"{key}_g_logit_scale"])) < 1e-5 * max(1e-3, abs(grads["logit_scale"].item()))


@pytest.mark.parametrize("local_loss,gwg", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_oracle_clip_loss_matches_reference_ranks(local_loss, gwg):
    z = np.load(GOLDEN)
    W, b = int(z["world"]), int(z["b"])
    img0, txt0 = torch.from_numpy(z["image_features"]), torch.from_numpy(z["text_features"])
    scale = float(z["scale"])
    for r in range(W):
        key = f"clip_l{local_loss}_g{gwg}_r{r}"
        img, txt = img0.clone().requires_grad_(), txt0.clone().requires_grad_()
        ls = torch.tensor(scale, requires_grad=True)
        sl = slice(r * b, (r + 1) * b)
        if not local_loss:
            loss = O.clip_loss(img, txt, ls)
            loss.backward()
            mult = O.gather_scaling(bool(gwg), W)
            g_img, g_txt, g_ls = img.grad[sl] * mult, txt.grad[sl] * mult, ls.grad
        elif not gwg:
            loss = O.clip_loss_local(img, txt, ls, r, W, img0, txt0)
            loss.backward()
            g_img, g_txt, g_ls = img.grad[sl], txt.grad[sl], ls.grad
        else:
            # every rank's loss reaches the gathered operands: local grads = d(sum over ranks)/d(local slice)
            per_rank = [O.clip_loss_local(img, txt, ls if q == r else ls.detach(), q, W) for q in range(W)]
            loss = per_rank[r]
            sum(per_rank).backward()
            g_img, g_txt, g_ls = img.grad[sl], txt.grad[sl], ls.grad
        assert abs(loss.item() - float(z[key + "_loss"])) < 1e-5
        assert _close(g_img, z[key + "_g_image_features"]), (r, "img")
        assert _close(g_txt, z[key + "_g_text_features"]), (r, "txt")
        assert abs(g_ls.item() - float(z[key + "_g_logit_scale"])) < 1e-5 * max(1e-3, abs(float(z[key + "_g_logit_scale"])))


# ---- our host logic under 2 gloo ranks vs the same vectors -------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["CXL_DIMG_REDUCE_MB"] = "0.0005"        # dImg partials go home in chunks of one or two images
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    out = {}
    try:
        z = np.load(GOLDEN_W[world])
        b, scale, alpha = int(z["b"]), float(z["scale"]), float(z["alpha"])
        sl = slice(rank * b, (rank + 1) * b)
        feats = [torch.from_numpy(z[n])[sl] for n in NAMES]
        for local_loss in (0, 1):
            for gwg in (0, 1):
                xs = [t.clone().requires_grad_() for t in feats[:2]]
                ls = torch.tensor(scale, requires_grad=True)
                crit = L.ClipLoss(local_loss=bool(local_loss), gather_with_grad=bool(gwg), rank=rank,
                                  world_size=world, kernels=TorchRefKernels())
                res = crit(*xs, ls, output_dict=True)
                assert set(res) == {"total_loss"}
                res["total_loss"].backward()
                key = f"clip_l{local_loss}_g{gwg}_r{rank}"
                out[key + "_loss"] = res["total_loss"].item()
                out[key + "_g_image_features"] = xs[0].grad.numpy().copy()
                out[key + "_g_text_features"] = xs[1].grad.numpy().copy()
                out[key + "_g_logit_scale"] = ls.grad.item()
        for gwg in (0, 1):
            xs = [t.clone().requires_grad_() for t in feats]
            ls = torch.tensor(scale, requires_grad=True)
            crit = L.ColClipLoss(alpha=alpha, gather_with_grad=bool(gwg), rank=rank, world_size=world,
                                 kernels=TorchRefKernels())
            res = crit(*xs, ls, output_dict=True)
            res["total_loss"].backward()
            key = f"colclip_g{gwg}_r{rank}"
            for k, v in res.items():
                out[key + "_" + k] = v.item()
            for x, n in zip(xs, NAMES):
                out[key + "_g_" + n] = x.grad.numpy().copy()
            out[key + "_g_logit_scale"] = ls.grad.item()
        # the PUBLIC gather_features / ColClipLoss.get_logits (reference loss.py:48-92, 223-269) at world_size 2:
        # gathered tensors, the four logit matrices, and local gradients of a seeded linear functional of them
        # (the generator applied the same contiguous-grad hook to torch.distributed.nn.all_gather)
        import torch.distributed.nn as dnn
        orig = dnn.all_gather

        def all_gather_contig(t, *a, **k):
            outs = orig(t, *a, **k)
            for o in outs:
                o.register_hook(lambda g: g.contiguous())
            return outs

        dnn.all_gather = all_gather_contig
        torch.distributed.nn.all_gather = all_gather_contig
        B, D = world * b, int(z["D"])
        gw = torch.Generator().manual_seed(int(z["seed"]) + 100)
        w_img, w_txt = torch.randn(B, D, generator=gw), torch.randn(B, D, generator=gw)
        w_lpi, w_lpt = torch.randn(B, B, generator=gw), torch.randn(B, B, generator=gw)
        for gwg in (0, 1):
            xs = [t.clone().requires_grad_() for t in feats[:2]]
            ai, at = L.gather_features(xs[0], xs[1], local_loss=False, gather_with_grad=bool(gwg), rank=rank,
                                       world_size=world)
            ((ai * w_img).sum() + (at * w_txt).sum()).backward()
            key = f"gather_g{gwg}_r{rank}"
            out[key + "_all_image_features"] = ai.detach().numpy().copy()
            out[key + "_all_text_features"] = at.detach().numpy().copy()
            out[key + "_g_image_features"] = xs[0].grad.numpy().copy()
            out[key + "_g_text_features"] = xs[1].grad.numpy().copy()
            xs = [t.clone().requires_grad_() for t in feats]
            crit = L.ColClipLoss(alpha=alpha, gather_with_grad=bool(gwg), rank=rank, world_size=world,
                                 kernels=TorchRefKernels())
            lg = crit.get_logits(*xs, torch.tensor(scale))
            ((lg["logits_per_image"] * w_lpi).sum() + (lg["logits_per_text_token"] * w_lpt).sum()).backward()
            key = f"logits_g{gwg}_r{rank}"
            for k, v in lg.items():
                out[key + "_" + k] = v.detach().numpy().copy()
            for x, n in zip(xs, NAMES):
                out[key + "_g_" + n] = x.grad.numpy().copy()
        if rank == 0:       # the reference raises for local_loss at world > 1 (loss.py:246-248); so does the mirror
            crit = L.ColClipLoss(local_loss=True, rank=rank, world_size=world, kernels=TorchRefKernels())
            try:
                crit.get_logits(*feats, torch.tensor(scale))
                out["local_loss_raises"] = 0
            except NotImplementedError:
                out["local_loss_raises"] = 1
        else:               # keep the collectives of rank 0's call matched
            crit = L.ColClipLoss(local_loss=True, rank=rank, world_size=world, kernels=TorchRefKernels())
            try:
                crit.get_logits(*feats, torch.tensor(scale))
            except NotImplementedError:
                pass
        q.put(out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_ranks_match_reference_ranks(world):
    z = np.load(GOLDEN_W[world])
    assert int(z["world"]) == world
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        got.update(q.get(timeout=180))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got.pop("local_loss_raises") == 1
    checked = 0
    for k, v in got.items():
        ref = z[k]
        if np.ndim(ref) == 0:
            assert abs(float(v) - float(ref)) <= 1e-5 * max(1.0, abs(float(ref))), k
        else:
            assert _close(v, ref), k
        checked += 1
    assert checked == len([k for k in z.files if k.startswith(("clip_", "colclip_", "gather_", "logits_"))])
