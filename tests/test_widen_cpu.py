"""Host logic of the SURVEY 8(f) rows on CPU (stand-in kernel backend): ClipLoss (f-1), the MaxSim retrieval
loop (f-2) and the gradient-accumulation slice mode of ColClipLoss (f-4)."""
import os
import socket
from contextlib import nullcontext

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from colxlip_b200 import loss as L
from colxlip_b200 import retrieval as R
from oracle import colxlip_oracle as O
from tests.ref_kernels import TorchRefKernels


# ---- f-1 ClipLoss ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("scale", [1 / 0.07, 100.0])
@pytest.mark.parametrize("local_loss", [False, True])     # one process: both flags give the same loss
def test_cliploss_single_process_matches_oracle(scale, local_loss):
    img, txt, _, _ = O.synth_inputs(11, 3, 3, 32, seed=21)
    xs = [t.clone().requires_grad_() for t in (img, txt)]
    ls = torch.tensor(scale, requires_grad=True)
    out = L.ClipLoss(local_loss=local_loss, kernels=TorchRefKernels())(*xs, ls, output_dict=True)
    assert set(out) == {"total_loss"} and out["total_loss"].dim() == 0
    out["total_loss"].backward()
    ys = [t.clone().requires_grad_() for t in (img, txt)]
    ls2 = torch.tensor(scale, requires_grad=True)
    ref = O.clip_loss(*ys, ls2)
    ref.backward()
    assert abs(out["total_loss"].item() - ref.item()) < 1e-5
    for x, y in zip(xs, ys):
        assert (x.grad - y.grad).abs().max() <= 1e-5 * y.grad.abs().max() + 1e-8
    assert abs(ls.grad.item() - ls2.grad.item()) <= 1e-5 * max(1e-3, abs(ls2.grad.item()))


def test_cliploss_api_contract():
    img, txt, _, _ = O.synth_inputs(5, 3, 3, 16, seed=22)
    crit = L.ClipLoss(cache_labels=True, kernels=TorchRefKernels())
    t = crit(img, txt, 10.0)                                        # python-float scale, tensor out
    assert t.dim() == 0 and abs(t.item() - O.clip_loss(img, txt, torch.tensor(10.0)).item()) < 1e-5
    # logit_bias / precomputed logits are accepted and ignored, as in the reference's forward (loss.py:172)
    t2 = crit(img, txt, torch.tensor(10.0), logit_bias=torch.tensor(3.0), logits_per_image=img, foo=1)
    assert abs(t2.item() - t.item()) < 1e-7
    lab = crit.get_ground_truth(torch.device("cpu"), 5)
    assert lab.tolist() == [0, 1, 2, 3, 4] and crit.get_ground_truth(torch.device("cpu"), 5) is lab
    lpi, lpt = crit.get_logits(img, txt, torch.tensor(10.0), logit_bias=torch.tensor(0.5))
    torch.testing.assert_close(lpi, 10.0 * img @ txt.T + 0.5)
    torch.testing.assert_close(lpt, 10.0 * txt @ img.T + 0.5)
    with pytest.raises(RuntimeError):
        L.ClipLoss(use_horovod=True, world_size=2, kernels=TorchRefKernels())(img, txt, torch.tensor(1.0))
    with pytest.raises(ValueError):
        crit(img, txt[:3], torch.tensor(1.0))
    with pytest.raises(RuntimeError):                               # product backend: no CPU fallback
        L.ClipLoss()(img, txt, torch.tensor(1.0))


# ---- f-2 retrieval loop ----------------------------------------------------------------------------------
class _Loader(list):
    num_samples = 0


class _ToyModel(torch.nn.Module):
    """encode_* return (global, tokens) like ColXLIP (model.py:532-609); features are looked up by id."""

    def __init__(self, img_tok, txt_tok):
        super().__init__()
        self.img_tok, self.txt_tok = img_tok, txt_tok
        self.logit_scale = torch.nn.Parameter(torch.tensor(np.log(20.0), dtype=torch.float32))

    def encode_image(self, images, normalize=False):
        return None, self.img_tok[images.long().view(-1)]

    def encode_text(self, texts, normalize=False):
        return None, self.txt_tok[texts.long().view(-1)]


def test_retrieval_on_split_matches_oracle():
    n_img, caps_per = 9, 2
    _, _, ti, tt = O.synth_inputs(n_img, 7, 5, 16, seed=31, pad_text=True, Bt=n_img * caps_per)
    img_ids = torch.tensor([40, 7, 13, 99, 5, 61, 22, 80, 3])        # arbitrary gallery ids, loader order
    img2txt = {int(img_ids[i]): [2 * i, 2 * i + 1] for i in range(n_img)}
    txt2img = {c: [int(img_ids[c // 2])] for c in range(n_img * caps_per)}
    img_loader = _Loader([(torch.arange(0, 5).float(), img_ids[:5]), (torch.arange(5, 9).float(), img_ids[5:])])
    txt_loader = _Loader([(torch.arange(0, 10), torch.arange(0, 10)), (torch.arange(10, 18), torch.arange(10, 18))])
    img_loader.num_samples, txt_loader.num_samples = n_img, n_img * caps_per
    model = _ToyModel(ti, tt)
    metrics = R.retrieval_on_split("docci", model, txt_loader, img_loader, img2txt, txt2img, None, 3, {},
                                   torch.device("cpu"), torch.float32, nullcontext, kernels=TorchRefKernels())
    S = O.maxsim_scores(ti, tt, O.MEAN)                              # [txt, img]
    ref = O.compute_retrieval(20.0 * S.T, [c // 2 for c in range(18)], [[2 * i, 2 * i + 1] for i in range(n_img)])
    assert metrics["epoch"] == 3 and metrics["docci_num_text_samples"] == 18 and metrics["docci_num_image_samples"] == 9
    for k, v in ref.items():
        assert abs(metrics["docci_" + k] - v) < 1e-9, k
    # block-wise score matrix == one-shot scores
    M = R.maxsim_score_matrix(ti, tt, image_chunk=4, text_chunk=5, kernels=TorchRefKernels())
    torch.testing.assert_close(M, S.T, rtol=1e-6, atol=1e-7)


# ---- f-4 gradient-accumulation slice ---------------------------------------------------------------------
@pytest.mark.parametrize("lo,hi", [(0, 4), (4, 8), (2, 7)])
def test_grad_rows_equals_cat_of_cached_and_live(lo, hi):
    """train.py:166-183: inputs = cat(cached[:j] + [live] + cached[j+1:]); with grad_rows the live slice's
    gradients are those of the full backward, the rest is zero."""
    feats = O.synth_inputs(8, 5, 4, 16, seed=41, pad_text=True)
    live = [t[lo:hi].clone().requires_grad_() for t in feats]
    cat = [torch.cat([t[:lo], l, t[hi:]]) for t, l in zip(feats, live)]
    ls = torch.tensor(20.0, requires_grad=True)
    out = L.ColClipLoss(alpha=0.3, kernels=TorchRefKernels())(*cat, ls, output_dict=True, grad_rows=(lo, hi))
    out["total_loss"].backward()
    losses, g = O.loss_and_grads(*feats, 20.0, 0.3)
    assert abs(out["total_loss"].item() - losses["total_loss"].item()) < 1e-5
    for l, n in zip(live, ("image_features", "text_features", "token_image_features", "token_text_features")):
        torch.testing.assert_close(l.grad, g[n][lo:hi], rtol=1e-4, atol=1e-7)
    assert abs(ls.grad.item() - g["logit_scale"].item()) <= 1e-5 * abs(g["logit_scale"].item())
    # token gradients outside the slice are zeros (not computed)
    full = [t.clone().requires_grad_() for t in feats]
    L.ColClipLoss(alpha=0.3, kernels=TorchRefKernels())(*full, torch.tensor(20.0), grad_rows=(lo, hi)).backward()
    mask = torch.ones(8, dtype=torch.bool)
    mask[lo:hi] = False
    assert full[2].grad[mask].abs().max() == 0 and full[3].grad[mask].abs().max() == 0
    with pytest.raises(ValueError):
        L.ColClipLoss(kernels=TorchRefKernels())(*feats, torch.tensor(20.0), grad_rows=(3, 9))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, gwg, q, shared=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    try:
        b, lo, hi = 4, 1, 3
        feats = O.synth_inputs(b * world, 6, 5, 16, seed=42, pad_text=True)
        sl = slice(rank * b, (rank + 1) * b)
        xs = [t[sl].clone().requires_grad_() for t in feats]
        ls = torch.tensor(12.0, requires_grad=True)
        crit = L.ColClipLoss(alpha=0.4, gather_with_grad=gwg, rank=rank, world_size=world, kernels=TorchRefKernels())
        if shared:      # ColClipLoss.accumulation: live slice tensors only
            state = crit.accumulation(*[x.detach() for x in xs], ls)
            live = [x.detach()[lo:hi].clone().requires_grad_() for x in xs]
            state.slice_loss(lo, hi, *live, ls).backward()
            grads = []
            for x, l in zip(xs, live):
                g = torch.zeros_like(x)
                g[lo:hi] = l.grad
                grads.append(g.numpy().copy())
            q.put((rank, grads))
        else:
            crit(*xs, ls, grad_rows=(lo, hi)).backward()
            q.put((rank, [x.grad.numpy().copy() for x in xs]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("gwg,shared", [(False, False), (True, False), (False, True), (True, True)])
def test_grad_rows_world2_gloo(gwg, shared):
    world, b, lo, hi = 2, 4, 1, 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, gwg, q, shared)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    feats = O.synth_inputs(b * world, 6, 5, 16, seed=42, pad_text=True)
    _, g = O.loss_and_grads(*feats, 12.0, 0.4)
    mult = O.gather_scaling(gwg, world)
    for rank, grads in res:
        for gr, n in zip(grads[2:], ("token_image_features", "token_text_features")):
            ref = g[n][rank * b + lo:rank * b + hi] * mult
            gr = torch.from_numpy(gr)
            assert (gr[lo:hi] - ref).abs().max() <= 1e-5 * ref.abs().max() + 1e-8, (rank, n)
            assert gr[:lo].abs().max() == 0 and gr[hi:].abs().max() == 0


# ---- f-3 normalize_tokens (autograd wiring; the kernels themselves are checked on the GPU) ---------------
@pytest.mark.parametrize("out_dtype", [None, torch.float32, torch.bfloat16])
def test_normalize_tokens_wiring(out_dtype):
    import torch.nn.functional as F
    from colxlip_b200.model import normalize_tokens
    g = torch.Generator().manual_seed(5)
    x0 = torch.randn(4, 7, 16, generator=g) * 2
    x0[0, 1] = 0
    grad = torch.randn(4, 7, 16, generator=g)
    x = x0.clone().requires_grad_()
    y = normalize_tokens(x, out_dtype=out_dtype, kernels=TorchRefKernels())
    assert y.dtype == (out_dtype or torch.float32)
    y.backward(grad.to(y.dtype))
    xr = x0.clone().requires_grad_()
    F.normalize(xr, dim=-1).backward(grad.to(y.dtype).float())
    tol = 1e-6 if y.dtype == torch.float32 else 1e-2
    assert (y.float() - F.normalize(x0, dim=-1)).abs().max() <= max(tol, 4e-3 if y.dtype == torch.bfloat16 else 0)
    assert (x.grad - xr.grad).abs().max() <= tol * xr.grad.abs().max()
    assert x.grad.dtype == torch.float32
    # no grad requested: nothing saved, plain tensor out
    assert not normalize_tokens(x0, kernels=TorchRefKernels()).requires_grad
    with pytest.raises(ValueError):
        normalize_tokens(torch.randn(3, 12), kernels=TorchRefKernels())
    with pytest.raises(RuntimeError):           # product backend refuses CPU tensors: no CPU fallback
        normalize_tokens(x0)


# ---- f-4, shared forward: one K1 for the whole accumulation step ------------------------------------------
def test_accumulation_shared_forward_equals_reference_loop():
    """The reference's loop (train.py:166-183) on the oracle vs ColClipLoss.accumulation + slice_loss."""
    B, acc = 8, 4
    sub = B // acc
    feats = O.synth_inputs(B, 5, 4, 16, seed=51, pad_text=True)
    # reference-style loop on the oracle: accum_freq full losses, each back-propagating into its live slice
    ls_ref = torch.tensor(15.0, requires_grad=True)
    ref_grads = [torch.zeros_like(t) for t in feats]
    for j in range(acc):
        lo, hi = j * sub, (j + 1) * sub
        live = [t[lo:hi].clone().requires_grad_() for t in feats]
        cat = [torch.cat([t[:lo], l, t[hi:]]) for t, l in zip(feats, live)]
        total = O.ColClipLossOracle(0.6)(*cat, ls_ref)
        total.backward()
        for g, l in zip(ref_grads, live):
            g[lo:hi] = l.grad
    # shared forward
    crit = L.ColClipLoss(alpha=0.6, kernels=TorchRefKernels())
    ls = torch.tensor(15.0, requires_grad=True)
    state = crit.accumulation(*feats, ls)
    assert isinstance(state, L.ColClipAccumulation)
    for j in range(acc):
        lo, hi = j * sub, (j + 1) * sub
        live = [t[lo:hi].clone().requires_grad_() for t in feats]
        out = state.slice_loss(lo, hi, *live, ls, output_dict=True, verify=True)
        assert set(out) == {"global_contrastive_loss", "token_contrastive_loss", "total_loss"}
        assert abs(out["total_loss"].item() - total.item()) < 1e-5
        out["total_loss"].backward()
        for l, g, n in zip(live, ref_grads, ("img", "txt", "tok_img", "tok_txt")):
            assert l.grad.shape == l.shape
            torch.testing.assert_close(l.grad, g[lo:hi], rtol=1e-4, atol=1e-7, msg=n)
    assert abs(ls.grad.item() - ls_ref.grad.item()) <= 1e-5 * abs(ls_ref.grad.item())      # accum_freq x d scale
    # a model that is NOT deterministic between passes is caught by verify=True
    bad = [t[0:sub].clone() for t in feats]
    bad[3] = bad[3] * 0.5
    with pytest.raises(RuntimeError):
        state.slice_loss(0, sub, *bad, ls, verify=True)
    with pytest.raises(ValueError):
        state.slice_loss(0, sub + 1, *[t[0:sub] for t in feats], ls)


def test_token_head_is_state_dict_compatible_and_masks_like_the_reference():
    """TokenHead = the reference's nn.Sequential(LayerNorm, Linear, GELU, LayerNorm) (model.py:518-530): same keys, same
    CPU results; encode_tokens zeroes the tokens at and after the EOT position BEFORE the head (model.py:580-593)."""
    import torch.nn as nn
    from colxlip_b200 import TokenHead, encode_tokens, eot_valid_len
    torch.manual_seed(0)
    ref = nn.Sequential(nn.LayerNorm(64), nn.Linear(64, 32), nn.GELU(), nn.LayerNorm(32))
    head = TokenHead(64, 32)
    assert list(head.state_dict().keys()) == list(ref.state_dict().keys())
    head.load_state_dict(ref.state_dict())
    x = torch.randn(3, 7, 64)
    torch.testing.assert_close(head(x), ref(x))
    text = torch.tensor([[5, 9, 49407, 0, 0, 0, 0], [3, 49407, 0, 0, 0, 0, 0], [1, 2, 3, 4, 5, 6, 49407]])
    vl = eot_valid_len(text)
    assert vl.tolist() == [2, 1, 6]
    # the reference's masking (model.py:583-593)
    seq = torch.arange(7)[None, :].expand(3, -1)
    mask = seq < text.argmax(dim=-1)[:, None]
    masked = torch.zeros_like(x)
    masked[mask] = x[mask]
    want = torch.nn.functional.normalize(ref(masked), dim=-1)
    got = encode_tokens(x, head, valid_len=vl, normalize=True)
    torch.testing.assert_close(got, want)
    assert torch.allclose(got[0, 2], got[1, 5])                  # masked positions all carry the head's image of zero
