"""Multi-rank numerics of colxlip_b200 over the live NCCL group, without the oracle:

 1. against vectors produced by EXECUTING THE REFERENCE under 2 / 4 / 8 gloo ranks (tests/golden/dist_w{N}.npz,
    tests/golden/make_golden_dist.py): ColClipLoss (both gather_with_grad modes, whole-shard and chunked dImg
    reductions), ClipLoss (local_loss x gather_with_grad), and the public gather_features / get_logits;
 2. the tensor-core path (bf16 autocast, ViT-B/16 token geometry) against the SAME package run as one process on
    the global batch on this rank's GPU (that single-GPU path is what tests/test_gpu_parity*.py pin to the oracle):
    losses, all five gradients, both gather modes, and a grad_rows slice.

`check(rank, world, device)` is collective (every rank calls it) and returns
{"ok", "world", "max_rel_err", "max_rel_err_vs_reference_vectors", "max_rel_err_vs_single_gpu_path", "checks", "failed"}.
bench.py --gpus N puts it in its JSON line as "dist_parity"; tools/dist_check.py prints it.
"""
import os

import numpy as np
import torch
import torch.distributed as dist

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
NAMES = ("image_features", "text_features", "token_image_features", "token_text_features")


class _Acc:
    def __init__(self):
        self.worst = {"ref": 0.0, "self": 0.0}
        self.n = 0
        self.failed = []

    def add(self, kind, what, got, ref, tol):
        got = torch.as_tensor(got, dtype=torch.float32).detach().cpu()
        ref = torch.as_tensor(ref, dtype=torch.float32).detach().cpu()
        den = max(ref.abs().max().item(), 1e-30)
        rel = (got - ref).abs().max().item() / den
        self.worst[kind] = max(self.worst[kind], rel)
        self.n += 1
        if not rel <= tol:
            self.failed.append(f"{what}: rel {rel:.3e} > {tol:.1e}")


def _reference_vectors(acc, rank, world, dev):
    from colxlip_b200 import ClipLoss, ColClipLoss
    from colxlip_b200 import loss as L
    path = os.path.join(GOLDEN_DIR, f"dist_w{world}.npz")
    if not os.path.exists(path):
        return False
    z = np.load(path)
    b, scale, alpha = int(z["b"]), float(z["scale"]), float(z["alpha"])
    sl = slice(rank * b, (rank + 1) * b)
    feats = [torch.from_numpy(z[n])[sl].to(dev) for n in NAMES]
    for chunk_mb in (None, "0.0005"):                        # whole-shard and one-or-two-image dImg reductions
        old = os.environ.get("CXL_DIMG_REDUCE_MB")
        if chunk_mb is not None:
            os.environ["CXL_DIMG_REDUCE_MB"] = chunk_mb
        try:
            for gwg in (0, 1):
                xs = [t.clone().requires_grad_() for t in feats]
                ls = torch.tensor(scale, device=dev, requires_grad=True)
                res = ColClipLoss(alpha=alpha, gather_with_grad=bool(gwg), rank=rank, world_size=world)(*xs, ls, output_dict=True)
                res["total_loss"].backward()
                key = f"colclip_g{gwg}_r{rank}"
                for k, v in res.items():
                    acc.add("ref", f"{key}_{k}", v, float(z[f"{key}_{k}"]), 1e-5)
                for x, n in zip(xs, NAMES):
                    acc.add("ref", f"{key}_g_{n}", x.grad, z[f"{key}_g_{n}"], 2e-5)
                acc.add("ref", f"{key}_g_logit_scale", ls.grad, float(z[f"{key}_g_logit_scale"]), 5e-5)
        finally:
            if chunk_mb is not None:
                if old is None:
                    os.environ.pop("CXL_DIMG_REDUCE_MB", None)
                else:
                    os.environ["CXL_DIMG_REDUCE_MB"] = old
    for local_loss in (0, 1):
        for gwg in (0, 1):
            xs = [t.clone().requires_grad_() for t in feats[:2]]
            ls = torch.tensor(scale, device=dev, requires_grad=True)
            loss = ClipLoss(local_loss=bool(local_loss), gather_with_grad=bool(gwg), rank=rank, world_size=world)(*xs, ls)
            loss.backward()
            key = f"clip_l{local_loss}_g{gwg}_r{rank}"
            acc.add("ref", key + "_loss", loss, float(z[key + "_loss"]), 1e-5)
            acc.add("ref", key + "_g_image_features", xs[0].grad, z[key + "_g_image_features"], 2e-5)
            acc.add("ref", key + "_g_text_features", xs[1].grad, z[key + "_g_text_features"], 2e-5)
            acc.add("ref", key + "_g_logit_scale", ls.grad, float(z[key + "_g_logit_scale"]), 5e-5)
    # public gather_features / get_logits (reference loss.py:48-92, 223-269)
    B, D = world * b, int(z["D"])
    gw = torch.Generator().manual_seed(int(z["seed"]) + 100)
    w_img, w_txt = torch.randn(B, D, generator=gw).to(dev), torch.randn(B, D, generator=gw).to(dev)
    w_lpi, w_lpt = torch.randn(B, B, generator=gw).to(dev), torch.randn(B, B, generator=gw).to(dev)
    for gwg in (0, 1):
        xs = [t.clone().requires_grad_() for t in feats[:2]]
        ai, at = L.gather_features(xs[0], xs[1], local_loss=False, gather_with_grad=bool(gwg), rank=rank, world_size=world)
        ((ai * w_img).sum() + (at * w_txt).sum()).backward()
        key = f"gather_g{gwg}_r{rank}"
        acc.add("ref", key + "_all_image_features", ai, z[key + "_all_image_features"], 1e-6)
        acc.add("ref", key + "_all_text_features", at, z[key + "_all_text_features"], 1e-6)
        acc.add("ref", key + "_g_image_features", xs[0].grad, z[key + "_g_image_features"], 1e-5)
        acc.add("ref", key + "_g_text_features", xs[1].grad, z[key + "_g_text_features"], 1e-5)
        xs = [t.clone().requires_grad_() for t in feats]
        lg = ColClipLoss(alpha=alpha, gather_with_grad=bool(gwg), rank=rank, world_size=world).get_logits(
            *xs, torch.tensor(scale, device=dev))
        ((lg["logits_per_image"] * w_lpi).sum() + (lg["logits_per_text_token"] * w_lpt).sum()).backward()
        key = f"logits_g{gwg}_r{rank}"
        for k, v in lg.items():
            acc.add("ref", f"{key}_{k}", v, z[f"{key}_{k}"], 2e-5)
        for x, n in zip(xs, NAMES):
            acc.add("ref", f"{key}_g_{n}", x.grad, z[f"{key}_g_{n}"], 5e-5)
    return True


def _synth(B, P, T, D, dev, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    nrm = torch.nn.functional.normalize
    mk = lambda *s: nrm(torch.randn(*s, device=dev, generator=g), dim=-1).to(torch.bfloat16).float()   # noqa: E731
    img, txt, ti, tt = mk(B, D), mk(B, D), mk(B, P, D), mk(B, T, D)
    lens = torch.randint(4, T + 1, (B,), device=dev, generator=g)
    tt = tt * (torch.arange(T, device=dev)[None, :] < lens[:, None])[..., None]       # zero-padded text tails
    return img, txt, ti, tt


def _run(crit, xs, scale, dev, **kw):
    xs = [t.clone().requires_grad_() for t in xs]
    ls = torch.tensor(scale, device=dev, requires_grad=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = crit(*xs, ls, output_dict=True, **kw)
    out["total_loss"].backward()
    return out, [x.grad.float() for x in xs], ls.grad


def _single_gpu_path(acc, rank, world, dev):
    from colxlip_b200 import ColClipLoss
    b, P, T, D, scale = 24, 196, 77, 256, 1 / 0.07
    feats = _synth(b * world, P, T, D, dev, seed=4321)                                 # same global batch on every rank
    sl = slice(rank * b, (rank + 1) * b)
    one, g_one, ls_one = _run(ColClipLoss(alpha=0.5), feats, scale, dev)             # W = 1 on the global batch
    for gwg in (False, True):
        crit = ColClipLoss(alpha=0.5, rank=rank, world_size=world, gather_with_grad=gwg)
        out, grads, ls_g = _run(crit, [t[sl] for t in feats], scale, dev)
        mult = float(world) if gwg else 1.0
        for k in one:
            acc.add("self", f"tc gwg={gwg} {k}", out[k], one[k], 2e-5)
        for gr, ref, n in zip(grads, g_one, NAMES):
            # token grads: pair weights are rounded to bf16 from fp32 values that differ in their last bits between the
            # two row partitions of K3 -> a few weights land one bf16 step apart (2^-8 of one term)
            acc.add("self", f"tc gwg={gwg} g_{n}", gr, ref[sl] * mult, 4e-3)
        acc.add("self", f"tc gwg={gwg} g_logit_scale", ls_g, ls_one, 2e-4)
        lo, hi = 16, 20                                                                # lo * T % 16 == 0: streaming slice kernel
        _, part, _ = _run(crit, [t[sl] for t in feats], scale, dev, grad_rows=(lo, hi))
        for gr, full, n in zip(part, grads, NAMES):
            acc.add("self", f"tc gwg={gwg} grad_rows g_{n}", gr[lo:hi], full[lo:hi], 2e-5)
            if n.startswith("token"):
                outside = torch.cat([gr[:lo].reshape(-1), gr[hi:].reshape(-1)])
                if outside.abs().max().item() != 0:
                    acc.failed.append(f"tc gwg={gwg} grad_rows g_{n}: non-zero outside the slice")


def check(rank: int, world: int, dev: torch.device) -> dict:
    acc = _Acc()
    had_ref = _reference_vectors(acc, rank, world, dev)
    _single_gpu_path(acc, rank, world, dev)
    torch.cuda.synchronize(dev)
    t = torch.tensor([acc.worst["ref"], acc.worst["self"], float(len(acc.failed))], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res = {"ok": t[2].item() == 0, "world": world, "max_rel_err": max(t[0].item(), t[1].item()),
           "max_rel_err_vs_reference_vectors": t[0].item() if had_ref else None,
           "max_rel_err_vs_single_gpu_path": t[1].item(), "checks": acc.n,
           "reference_vectors": f"tests/golden/dist_w{world}.npz" if had_ref else None,
           "failed": acc.failed[:8]}
    return res
