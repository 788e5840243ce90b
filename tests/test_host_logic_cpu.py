"""Host-side logic of colxlip_b200.loss on CPU with a stand-in kernel backend (no CUDA needed):
autograd wiring, alpha/coef handling, output contract, error behaviour, and the world_size-2
row-sharded path over gloo (SURVEY 5.8 scaling rules)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from colxlip_b200 import loss as L
from colxlip_b200.retrieval import compute_retrieval
from oracle import colxlip_oracle as O
from tests.ref_kernels import TorchRefKernels


def _inputs(B, P, T, D, seed, pad=True):
    img, txt, ti, tt = O.synth_inputs(B, P, T, D, seed=seed, pad_text=pad)
    return img, txt, ti, tt


@pytest.mark.parametrize("alpha,scale", [(0.5, 1 / 0.07), (0.7, 30.0), (0.0, 5.0), (1.0, 100.0)])
def test_single_process_matches_oracle(alpha, scale):
    img, txt, ti, tt = _inputs(9, 11, 6, 16, 3)
    ref_losses, ref_grads = O.loss_and_grads(img, txt, ti, tt, scale, alpha)
    xs = [t.clone().requires_grad_() for t in (img, txt, ti, tt)]
    ls = torch.tensor(scale, requires_grad=True)
    crit = L.ColClipLoss(alpha=alpha, kernels=TorchRefKernels())
    out = crit(*xs, ls, output_dict=True)
    assert set(out) == {"global_contrastive_loss", "token_contrastive_loss", "total_loss"}
    for k in out:
        assert out[k].dim() == 0
        assert abs(out[k].item() - ref_losses[k].item()) < 1e-5
    out["total_loss"].backward()
    names = ["image_features", "text_features", "token_image_features", "token_text_features"]
    for x, n in zip(xs, names):
        tol = 1e-5 * ref_grads[n].abs().max().item() + 1e-8
        assert (x.grad - ref_grads[n]).abs().max().item() <= tol, n
    assert abs(ls.grad.item() - ref_grads["logit_scale"].item()) <= 1e-5 * max(1e-3, abs(ref_grads["logit_scale"].item()))


def test_individual_losses_are_differentiable():
    img, txt, ti, tt = _inputs(6, 7, 5, 16, 4)
    xs = [t.clone().requires_grad_() for t in (img, txt, ti, tt)]
    out = L.ColClipLoss(alpha=0.3, kernels=TorchRefKernels())(*xs, torch.tensor(10.0), output_dict=True)
    out["token_contrastive_loss"].backward()
    ys = [t.clone().requires_grad_() for t in (img, txt, ti, tt)]
    ref = O.ColClipLossOracle(0.3)(*ys, torch.tensor(10.0), output_dict=True)
    ref["token_contrastive_loss"].backward()
    assert xs[0].grad.abs().max() == 0 and xs[1].grad.abs().max() == 0
    torch.testing.assert_close(xs[2].grad, ys[2].grad, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(xs[3].grad, ys[3].grad, rtol=1e-4, atol=1e-7)


def test_no_backward_means_no_saved_argmax():
    """Validation-style calls (torch.no_grad, or inputs that do not require grad) must not ask K1 for the argmax:
    at cfg-2 that is an 82 GB buffer and extra store traffic for nothing."""
    class Spy(TorchRefKernels):
        def __init__(self):
            self.saved = []

        def maxsim_fwd(self, txt_op, img_op, T, P, prec, mean_mode, save_argmax):
            self.saved.append(bool(save_argmax))
            return super().maxsim_fwd(txt_op, img_op, T, P, prec, mean_mode, save_argmax)

    img, txt, ti, tt = _inputs(5, 4, 3, 16, 6)
    spy = Spy()
    crit = L.ColClipLoss(kernels=spy)
    a = crit(img, txt, ti, tt, torch.tensor(14.0))                        # nothing requires grad
    xs = [t.clone().requires_grad_() for t in (img, txt, ti, tt)]
    with torch.no_grad():
        b = crit(*xs, torch.tensor(14.0))                                 # grad mode off
    c = crit(*xs, torch.tensor(14.0))                                     # training call
    c.backward()
    assert spy.saved == [False, False, True]
    assert abs(a.item() - c.item()) < 1e-6 and abs(b.item() - c.item()) < 1e-6
    ls = torch.tensor(14.0, requires_grad=True)                           # only logit_scale trains: K3 grad needs S, cnt
    spy.saved.clear()
    crit(img, txt, ti, tt, ls).backward()
    assert spy.saved == [True] and ls.grad is not None
    # same for the bare scorer
    spy.saved.clear()
    with torch.no_grad():
        L._maxsim(xs[2], xs[3], 0, spy)
    L._maxsim(xs[2], xs[3], 0, spy).sum().backward()
    assert spy.saved == [False, True]


def test_precision_selection_follows_the_operand_dtype_of_the_reference_einsum():
    """fp16 inputs (or fp16 autocast, the reference's `--precision amp`) select the fp16 operand path: they are not
    silently computed with bf16 operands (VERDICT r1 missing #2)."""
    h, bf, f = (torch.zeros(2, 3, 64, dtype=d) for d in (torch.float16, torch.bfloat16, torch.float32))
    assert L._select_precision(h, h) == "fp16"
    assert L._select_precision(bf, bf) == "bf16"
    assert L._select_precision(f, f) == "fp32" and L._select_precision(f, bf) == "fp32"
    assert L._select_precision(h, bf) == "bf16"
    assert L._select_precision(torch.zeros(2, 3, 48, dtype=torch.float16), h[..., :48]) == "fp32"   # D % 64 != 0
    ws = L._w_scale_from_bound(torch.tensor(3e-7))
    assert ws.shape == (2,) and ws[0].item() == 2.0 ** 36 and ws[0].item() * 3e-7 <= 2.0 ** 15 < 2 * ws[0].item() * 3e-7 * 2
    L.set_precision("fp16")
    try:
        assert L._select_precision(f, f) == "fp16"
    finally:
        L.set_precision("auto")
    # stand-in backend: fp16 rounding of the operands is what distinguishes the path
    img, txt, ti, tt = _inputs(5, 4, 3, 64, 7)
    a = L.ColClipLoss(kernels=TorchRefKernels())(img.half(), txt.half(), ti.half(), tt.half(), torch.tensor(14.0))
    ref = O.ColClipLossOracle(0.5)(img.half().float(), txt.half().float(), ti.half().float(), tt.half().float(), torch.tensor(14.0))
    assert abs(a.item() - ref.item()) < 1e-5


def test_total_loss_only_and_scalar_scale():
    img, txt, ti, tt = _inputs(5, 4, 3, 16, 5)
    t = L.ColClipLoss(kernels=TorchRefKernels())(img, txt, ti, tt, torch.tensor(14.0))
    assert t.dim() == 0
    ref = O.ColClipLossOracle(0.5)(img, txt, ti, tt, torch.tensor(14.0))
    assert abs(t.item() - ref.item()) < 1e-5


def test_accum_freq_pattern_partial_requires_grad():
    """train.py:176-180: cat of cached (no grad) and fresh (grad) features."""
    img, txt, ti, tt = _inputs(8, 5, 4, 16, 6)
    fresh = [t[:4].clone().requires_grad_() for t in (img, txt, ti, tt)]
    cat = [torch.cat([f, t[4:]]) for f, t in zip(fresh, (img, txt, ti, tt))]
    L.ColClipLoss(kernels=TorchRefKernels())(*cat, torch.tensor(20.0)).backward()
    _, g = O.loss_and_grads(img, txt, ti, tt, 20.0, 0.5)
    torch.testing.assert_close(fresh[2].grad, g["token_image_features"][:4], rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(fresh[3].grad, g["token_text_features"][:4], rtol=1e-4, atol=1e-7)


def test_get_logits_and_bias_quirk(golden_dir):
    z = np.load(os.path.join(golden_dir, "logits_bias_b6.npz"))
    t = lambda k: torch.from_numpy(z[k])
    out = L.ColClipLoss(kernels=TorchRefKernels()).get_logits(
        t("image_features"), t("text_features"), t("token_image_features"), t("token_text_features"),
        torch.tensor(10.0), logit_bias=torch.tensor(0.25))
    for k in ("logits_per_image", "logits_per_text", "logits_per_text_token"):
        torch.testing.assert_close(out[k], t(k), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(out["logits_per_image_token"], t("logits_per_text_token").T, rtol=1e-5, atol=1e-5)


def test_error_behaviour():
    img, txt, ti, tt = _inputs(4, 3, 3, 16, 7)
    with pytest.raises(NotImplementedError):
        L.ColClipLoss(local_loss=True, world_size=2, kernels=TorchRefKernels())(img, txt, ti, tt, torch.tensor(1.0))
    with pytest.raises(RuntimeError):
        L.ColClipLoss(use_horovod=True, world_size=2, kernels=TorchRefKernels())(img, txt, ti, tt, torch.tensor(1.0))
    with pytest.raises(RuntimeError):   # product backend refuses CPU tensors: no CPU fallback
        L.compute_colbert_similarity(ti, tt)
    crit = L.ColClipLoss(cache_labels=True)
    lab = crit.get_ground_truth(torch.device("cpu"), 5)
    assert lab.tolist() == [0, 1, 2, 3, 4] and crit.get_ground_truth(torch.device("cpu"), 5) is lab


def test_compute_retrieval_matches_oracle():
    g = torch.Generator().manual_seed(1)
    s = torch.randn(15, 30, generator=g)
    s[3, 7] = s[3, 9]   # a tie
    txt2img = [i // 2 for i in range(30)]
    img2txt = [[2 * i, 2 * i + 1] for i in range(15)]
    ours = compute_retrieval(s, txt2img, img2txt, kernels=TorchRefKernels())
    ref = O.compute_retrieval(s, txt2img, img2txt)
    assert ours.keys() == ref.keys()
    for k in ref:
        assert abs(ours[k] - ref[k]) < 1e-9, k


def test_compute_retrieval_takes_the_position_keyed_dicts_of_remap_indices():
    """The reference's retrieval_on_split passes the dicts remap_indices returns and indexes them by position
    (train.py:469,489); shuffled gallery ids, several captions per image, one image without captions."""
    from colxlip_b200.retrieval import remap_indices
    g = torch.Generator().manual_seed(3)
    n_img, caps = 9, 2
    s = torch.randn(n_img, n_img * caps, generator=g)
    gallery_ids = torch.tensor([17, 3, 99, 42, 8, 23, 5, 61, 70])          # loader order -> original image id
    img2txt = {int(i): [caps * j, caps * j + 1] for j, i in enumerate(gallery_ids.tolist())}
    txt2img = {c: [int(gallery_ids[c // caps])] for c in range(n_img * caps)}
    new_i2t, new_t2i = remap_indices(gallery_ids, torch.arange(n_img * caps), img2txt, txt2img)
    assert isinstance(new_i2t, dict) and isinstance(new_t2i, dict)
    ours = compute_retrieval(s, new_t2i, new_i2t, kernels=TorchRefKernels())
    ref = O.compute_retrieval(s, new_t2i, new_i2t)
    as_lists = compute_retrieval(s, [new_t2i[c] for c in range(n_img * caps)], [new_i2t[i] for i in range(n_img)],
                                 kernels=TorchRefKernels())
    for k in ref:
        assert abs(ours[k] - ref[k]) < 1e-9 and ours[k] == as_lists[k], k
    # an image without captions keeps the reference's sentinel rank (train.py:485) instead of raising
    new_i2t[4] = []
    ours = compute_retrieval(s, new_t2i, new_i2t, kernels=TorchRefKernels())
    ref = O.compute_retrieval(s, new_t2i, new_i2t)
    for k in ref:
        assert abs(ours[k] - ref[k]) <= 1e-9 * max(1.0, abs(ref[k])), k
    assert ours["image_to_text_mean_rank"] > 1e8


def test_compute_retrieval_rejects_bad_indices():
    s = torch.randn(4, 8, generator=torch.Generator().manual_seed(0))
    good_t, good_i = [i // 2 for i in range(8)], [[2 * i, 2 * i + 1] for i in range(4)]
    with pytest.raises(IndexError):
        compute_retrieval(s, [0, 1, 2, 3, 4, 0, 0, 0], good_i, kernels=TorchRefKernels())      # image 4 of 4
    with pytest.raises(IndexError):
        compute_retrieval(s, good_t, [[0], [-1], [2], [3]], kernels=TorchRefKernels())         # stale -1 sentinel
    with pytest.raises(ValueError):
        compute_retrieval(s, good_t[:-1], good_i, kernels=TorchRefKernels())
    with pytest.raises(KeyError):
        compute_retrieval(s, {i + 100: v for i, v in enumerate(good_t)}, good_i, kernels=TorchRefKernels())   # caption-id keys


# ---------------------------------------------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, gather_with_grad, q, shard_dimg=False):
    if shard_dimg:
        os.environ["CXL_DIMG_REDUCE_MB"] = "0.0005"     # dImg partials go home in pieces of one or two images (cfg-5 sized runs use bounded pieces)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    try:
        b = 4
        img, txt, ti, tt = _inputs(b * world, 6, 5, 16, 9)
        sl = slice(rank * b, (rank + 1) * b)
        xs = [t[sl].clone().requires_grad_() for t in (img, txt, ti, tt)]
        ls = torch.tensor(12.0, requires_grad=True)
        crit = L.ColClipLoss(alpha=0.4, gather_with_grad=gather_with_grad, rank=rank, world_size=world,
                             kernels=TorchRefKernels())
        out = crit(*xs, ls, output_dict=True)
        out["total_loss"].backward()
        q.put((rank, {k: v.item() for k, v in out.items()}, [x.grad.numpy().copy() for x in xs], ls.grad.item()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("gather_with_grad,shard_dimg", [(False, False), (True, False), (False, True), (True, True)])
def test_world2_gloo_matches_global_batch(gather_with_grad, shard_dimg):
    world, b = 2, 4
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, gather_with_grad, q, shard_dimg)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    img, txt, ti, tt = _inputs(b * world, 6, 5, 16, 9)
    ref_losses, ref_grads = O.loss_and_grads(img, txt, ti, tt, 12.0, 0.4)
    mult = O.gather_scaling(gather_with_grad, world)
    names = ["image_features", "text_features", "token_image_features", "token_text_features"]
    for rank, losses, grads, ls_grad in res:
        for k in losses:   # every rank reports the global-batch loss
            assert abs(losses[k] - ref_losses[k].item()) < 1e-5
        sl = slice(rank * b, (rank + 1) * b)
        for g, n in zip(grads, names):
            ref = ref_grads[n][sl] * mult
            g = torch.from_numpy(g)
            assert (g - ref).abs().max().item() <= 1e-5 * ref.abs().max().item() + 1e-8, (rank, n)
        assert abs(ls_grad - ref_grads["logit_scale"].item()) < 1e-5 * max(1e-3, abs(ref_grads["logit_scale"].item()))


def test_strict_autocast_switch_reproduces_the_reference_under_cpu_autocast():
    """set_strict_autocast: K1 is asked for 16-bit row maxima (mean_mode | ROUND16) exactly when the operands are 16-bit.
    With the stand-in kernels the scores then equal the reference's own arithmetic under torch.autocast (einsum output
    in bf16 before torch.max, loss.py:34-35; sum and division in fp32) up to fp32 summation order."""
    from colxlip_b200.kernels import ROUND16
    assert L._fwd_mode(0, "bf16") == 0 and L._fwd_mode(0, "fp32") == 0
    _, _, ti, tt = _inputs(6, 10, 7, 64, 3)
    ti, tt = ti.to(torch.bfloat16).float(), tt.to(torch.bfloat16).float()
    k = TorchRefKernels()
    try:
        L.set_strict_autocast(True)
        assert L.get_strict_autocast()
        assert L._fwd_mode(0, "bf16") == ROUND16 and L._fwd_mode(1, "fp16") == (1 | ROUND16) and L._fwd_mode(0, "fp32") == 0
        L.set_precision("bf16")
        S_strict = L._maxsim(ti, tt, 0, k)
    finally:
        L.set_strict_autocast(False)
        L.set_precision("auto")
    L.set_precision("bf16")
    try:
        S_default = L._maxsim(ti, tt, 0, k)
    finally:
        L.set_precision("auto")
    with torch.autocast("cpu", dtype=torch.bfloat16):
        sim = torch.einsum('mnd,kqd->mknq', tt, ti)
        assert sim.dtype == torch.bfloat16
        mx = sim.max(dim=3)[0]
        ref = torch.sum(mx.float(), dim=2) / (torch.sum((mx != 0).float(), dim=2) + 1e-8)
    assert (S_strict - ref).abs().max().item() <= 2e-4 * ref.abs().max().item()
    assert (S_default - ref).abs().max().item() > (S_strict - ref).abs().max().item()
