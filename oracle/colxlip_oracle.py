"""CPU oracle for the ColXLIP MaxSim-loss hot path.  TEST INFRASTRUCTURE ONLY.

This module is a plain-PyTorch fp32 restatement of the reference algorithm.  It is
the *checker* the CUDA path is compared against; it is never the thing shipped or
measured.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it.  The product package
``colxlip_b200`` must never import anything from ``oracle/``.

Parity status: **pinned by execution of the reference itself** (the reference
ships no tests, golden vectors or fixtures -- SURVEY.md section 4).  The script
``tests/golden/make_golden.py`` imports ``/root/reference/src/colxlip/loss.py`` by
path in the build container, runs it on seeded inputs and commits the resulting
vectors under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks every
function below against those vectors.

Reference lines followed (all relative to /root/reference/src/colxlip/):
  maxsim_rowmax / maxsim_scores(masked)  loss.py:34-44   (einsum, max, !=0 mask, +1e-8)
  maxsim_scores(mean)                    model.py:626-628 (einsum, max, plain mean)
  gather_scaling                         loss.py:48-92    (semantics probed)
  ColClipLossOracle.get_logits           loss.py:223-269
  ColClipLossOracle.forward              loss.py:271-296
  compute_retrieval                      train.py:457-508
  clip_loss / clip_loss_local            loss.py:95-182   (ClipLoss; the multi-rank cases are pinned by
                                                           tests/golden/dist_w2.npz = the reference under 2 gloo ranks)
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

MASKED = 0  # loss.py:38-44  (count of non-zero maxima + 1e-8)
MEAN = 1    # model.py:628   (plain mean over all T tokens)


def maxsim_rowmax(img_tok: torch.Tensor, txt_tok: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per text token: best patch similarity and its (first) index.

    img_tok [Bi,P,D], txt_tok [Bt,T,D] -> (mx [Bt,Bi,T], arg [Bt,Bi,T] int64).
    Follows loss.py:34-35: the contraction is 'mnd,kqd->mknq' and the max runs over
    the patch axis q; ``torch.max`` returns the first maximal index on ties.
    Done one text at a time so the [Bt,Bi,T,P] tensor is never materialised.
    """
    Bi, P, D = img_tok.shape
    Bt, T, _ = txt_tok.shape
    flat_img = img_tok.reshape(Bi * P, D)
    mx = torch.empty(Bt, Bi, T, dtype=img_tok.dtype, device=img_tok.device)
    arg = torch.empty(Bt, Bi, T, dtype=torch.int64, device=img_tok.device)
    for m in range(Bt):
        sim = (txt_tok[m] @ flat_img.T).view(T, Bi, P)           # [T,Bi,P]
        v, i = torch.max(sim, dim=2)                              # [T,Bi]
        mx[m] = v.T
        arg[m] = i.T
    return mx, arg


def maxsim_scores(img_tok: torch.Tensor, txt_tok: torch.Tensor, mode: int = MASKED) -> torch.Tensor:
    """S[Bt,Bi].  mode=MASKED follows loss.py:38-44, mode=MEAN follows model.py:628.

    Differentiable (autograd flows through the selected maxima exactly as it does
    through ``torch.max`` in the reference).
    """
    sim = torch.einsum("mnd,kqd->mknq", txt_tok, img_tok)
    mx = sim.max(dim=3).values
    if mode == MEAN:
        return mx.mean(dim=-1)
    nz = (mx != 0).to(torch.float32)
    return mx.sum(dim=2) / (nz.sum(dim=2) + 1e-8)


def maxsim_scores_blockwise(img_tok, txt_tok, mode: int = MASKED, block: int = 64) -> torch.Tensor:
    """Same values as :func:`maxsim_scores` without the [Bt,Bi,T,P] tensor."""
    Bt, Bi = txt_tok.shape[0], img_tok.shape[0]
    out = torch.empty(Bt, Bi, dtype=torch.float32, device=img_tok.device)
    for m0 in range(0, Bt, block):
        for k0 in range(0, Bi, block):
            out[m0:m0 + block, k0:k0 + block] = maxsim_scores(
                img_tok[k0:k0 + block].float(), txt_tok[m0:m0 + block].float(), mode)
    return out


def top2_gap(img_tok: torch.Tensor, txt_tok: torch.Tensor) -> torch.Tensor:
    """Gap between best and second-best patch similarity, [Bt,Bi,T].

    Used by the tests to classify argmax mismatches: a mismatch is only tolerated
    where the oracle's own top-2 scores tie within tolerance (north_star).
    """
    Bi, P, D = img_tok.shape
    Bt, T, _ = txt_tok.shape
    gap = torch.empty(Bt, Bi, T, device=img_tok.device)
    flat_img = img_tok.reshape(Bi * P, D)
    for m in range(Bt):
        sim = (txt_tok[m] @ flat_img.T).view(T, Bi, P)
        if P == 1:
            gap[m] = float("inf")
            continue
        top = sim.topk(2, dim=2).values
        gap[m] = (top[..., 0] - top[..., 1]).T
    return gap


def infonce_pair(logits: torch.Tensor) -> torch.Tensor:
    """(CE(rows) + CE(cols)) / 2 against the diagonal -- loss.py:285-293."""
    labels = torch.arange(logits.shape[0], device=logits.device)
    return (F.cross_entropy(logits, labels) + F.cross_entropy(logits.T, labels)) / 2


class ColClipLossOracle:
    """Single-process restatement of ColClipLoss (loss.py:184-296), world_size == 1."""

    def __init__(self, alpha: float = 0.5):
        self.alpha = alpha

    def get_logits(self, image_features, text_features, token_image_features,
                   token_text_features, logit_scale, logit_bias=None) -> Dict[str, torch.Tensor]:
        logits_per_image = logit_scale * image_features @ text_features.T            # loss.py:259
        logits_per_text = logits_per_image.T
        logits_per_text_token = logit_scale * maxsim_scores(token_image_features,    # loss.py:262
                                                            token_text_features, MASKED)
        logits_per_image_token = logits_per_text_token.T
        if logit_bias is not None:
            # loss.py:265-267 adds the bias in place to logits_per_image AND to
            # logits_per_text, which is a transposed *view* of the same storage, so the
            # reference's global logits end up with 2x the bias (pinned by golden vector).
            logits_per_image = logits_per_image + 2 * logit_bias
            logits_per_text = logits_per_image.T
        return {"logits_per_image": logits_per_image, "logits_per_text": logits_per_text,
                "logits_per_image_token": logits_per_image_token,
                "logits_per_text_token": logits_per_text_token}

    def __call__(self, image_features, text_features, token_image_features,
                 token_text_features, logit_scale, output_dict: bool = False):
        out = self.get_logits(image_features, text_features, token_image_features,
                              token_text_features, logit_scale)
        g = infonce_pair(out["logits_per_image"])
        t = infonce_pair(out["logits_per_text_token"])
        total = self.alpha * g + (1 - self.alpha) * t
        if output_dict:
            return {"global_contrastive_loss": g, "token_contrastive_loss": t, "total_loss": total}
        return total


def clip_loss(image_features, text_features, logit_scale) -> torch.Tensor:
    """ClipLoss.forward on one process / on the gathered global batch (loss.py:147-148 or 150-151, 172-181)."""
    logits_per_image = logit_scale * image_features @ text_features.T
    logits_per_text = logit_scale * text_features @ image_features.T
    labels = torch.arange(logits_per_image.shape[0], device=logits_per_image.device)
    return (F.cross_entropy(logits_per_image, labels) + F.cross_entropy(logits_per_text, labels)) / 2


def clip_loss_local(image_features, text_features, logit_scale, rank: int, world_size: int,
                    all_image_features=None, all_text_features=None) -> torch.Tensor:
    """ClipLoss with local_loss=True as rank `rank` of `world_size` sees it (loss.py:144-146, 118-119): its own
    b images against all texts and its own b texts against all images, labels offset by b*rank.  The
    gathered operands default to the inputs themselves (single-process emulation: slices of the same leaf,
    which reproduces gather_with_grad=True when the per-rank losses are summed); pass detached copies for
    gather_with_grad=False."""
    B = image_features.shape[0]
    b = B // world_size
    sl = slice(rank * b, (rank + 1) * b)
    all_img = image_features if all_image_features is None else all_image_features
    all_txt = text_features if all_text_features is None else all_text_features
    logits_per_image = logit_scale * image_features[sl] @ all_txt.T
    logits_per_text = logit_scale * text_features[sl] @ all_img.T
    labels = torch.arange(b, device=image_features.device) + b * rank
    return (F.cross_entropy(logits_per_image, labels) + F.cross_entropy(logits_per_text, labels)) / 2


def loss_and_grads(image_features, text_features, token_image_features, token_text_features,
                   logit_scale: float, alpha: float = 0.5):
    """fp32 loss values and gradients w.r.t. the five inputs (autograd of the restatement)."""
    xs = [t.detach().clone().float().requires_grad_() for t in
          (image_features, text_features, token_image_features, token_text_features)]
    ls = torch.tensor(float(logit_scale), requires_grad=True)
    out = ColClipLossOracle(alpha)(*xs, ls, output_dict=True)
    out["total_loss"].backward()
    return ({k: v.detach() for k, v in out.items()},
            {"image_features": xs[0].grad, "text_features": xs[1].grad,
             "token_image_features": xs[2].grad, "token_text_features": xs[3].grad,
             "logit_scale": ls.grad})


def maxsim_closed_form_grads(img_tok, txt_tok, dS, mode: int = MASKED, round_w_bf16: bool = False,
                             arg_override: Optional[torch.Tensor] = None):
    """Closed-form backward of maxsim_scores: no autograd, numpy-style loops.

    round_w_bf16: round the pair weight w to bf16 first -- what the reference's own backward bmm sees
    under bf16 autocast (dS arrives in bf16), and what the fast CUDA backward kernels compute with.

    cnt = nnz + 1e-8 (or T), w = dS / cnt,
    dTxt[m,n] = sum_k w[m,k] img[k,q*],  dImg[k,q] = sum_{m,n:q*=q} w[m,k] txt[m,n].
    """
    mx, arg = maxsim_rowmax(img_tok.float(), txt_tok.float())
    if arg_override is not None:
        # backward-only checks: scatter through a given argmax [Bt,Bi,T] (near-tied maxima may legitimately
        # resolve differently between accumulate orders; argmax parity itself is tested on the forward)
        arg = arg_override
    Bt, Bi, T = mx.shape
    if mode == MEAN:
        cnt = torch.full((Bt, Bi), float(T), device=img_tok.device)
    else:
        cnt = (mx != 0).float().sum(dim=2) + 1e-8
    w = dS / cnt
    if round_w_bf16:
        w = w.to(torch.bfloat16).float()
    dtxt = torch.zeros_like(txt_tok, dtype=torch.float32)
    dimg = torch.zeros_like(img_tok, dtype=torch.float32)
    for m in range(Bt):
        for k in range(Bi):
            q = arg[m, k]                                    # [T]
            dtxt[m] += w[m, k] * img_tok[k, q].float()
            dimg[k].index_add_(0, q, w[m, k] * txt_tok[m].float())
    return dtxt, dimg


def maxsim_closed_form_grads_by_text(img_tok, txt_tok, w, arg):
    """Same sums as maxsim_closed_form_grads for a GIVEN pair weight w [Bt,Bi] and argmax [Bt,Bi,T], one text at
    a time with vectorised gathers / index_add (runs on whatever device the inputs live on; used by the GPU
    tests at sizes where the double Python loop would take hours).  fp32 throughout."""
    Bi, P, D = img_tok.shape
    Bt, T, _ = txt_tok.shape
    img32, txt32 = img_tok.float(), txt_tok.float()
    flat_img = img32.reshape(Bi * P, D)
    dtxt = torch.zeros_like(txt32)
    dimg = torch.zeros(Bi * P, D, dtype=torch.float32, device=img_tok.device)
    base = (torch.arange(Bi, device=img_tok.device) * P)[:, None]                  # [Bi,1]
    for m in range(Bt):
        idx = (base + arg[m]).reshape(-1)                                          # [Bi*T] flat patch rows
        picked = flat_img[idx].view(Bi, T, D)
        dtxt[m] = torch.einsum("k,ktd->td", w[m], picked)                          # sum_k w[m,k] img[k,q*]
        contrib = (w[m][:, None, None] * txt32[m][None]).reshape(Bi * T, D)        # w[m,k] txt[m,n]
        dimg.index_add_(0, idx, contrib)
    return dtxt, dimg.view(Bi, P, D)


def loss_and_grads_blockwise(image_features, text_features, token_image_features, token_text_features,
                             logit_scale: float, alpha: float = 0.5, round_w_bf16: bool = False):
    """loss_and_grads (loss.py:271-296 + autograd) WITHOUT the [Bt,Bi,T,P] tensor: the token scores come from
    maxsim_rowmax, the InfoNCE halves go through autograd on the [B,B] matrices, and the token-feature gradients
    through the closed form.  Device-agnostic: the GPU parity tests evaluate it on the B200 at B = 1024, where
    the plain restatement would need 63 GB per copy of the similarity tensor.
    Returns (losses, grads, extras) with extras = {S, cnt, arg, dS}."""
    dev = token_image_features.device
    img = image_features.detach().clone().float().requires_grad_()
    txt = text_features.detach().clone().float().requires_grad_()
    ti, tt = token_image_features.detach().float(), token_text_features.detach().float()
    ls = torch.tensor(float(logit_scale), device=dev, requires_grad=True)
    mx, arg = maxsim_rowmax(ti, tt)
    nz = (mx != 0).float().sum(dim=2)
    S = (mx.sum(dim=2) / (nz + 1e-8)).requires_grad_()
    g = infonce_pair(ls * img @ txt.T)
    t = infonce_pair(ls * S)
    total = alpha * g + (1 - alpha) * t
    total.backward()
    w = S.grad / (nz + 1e-8)
    if round_w_bf16:
        w = w.to(torch.bfloat16).float()
    dtxt, dimg = maxsim_closed_form_grads_by_text(ti, tt, w, arg)
    losses = {"global_contrastive_loss": g.detach(), "token_contrastive_loss": t.detach(), "total_loss": total.detach()}
    grads = {"image_features": img.grad, "text_features": txt.grad, "token_image_features": dimg,
             "token_text_features": dtxt, "logit_scale": ls.grad}
    return losses, grads, {"S": S.detach(), "cnt": nz, "arg": arg, "dS": S.grad}


def gather_scaling(gather_with_grad: bool, world_size: int) -> float:
    """Local-feature gradient = this factor x the single-process slice."""
    return float(world_size) if gather_with_grad else 1.0


def compute_retrieval(similarity_scores: torch.Tensor, txt2img: Sequence[int],
                      img2txt: Sequence[Sequence[int]]) -> Dict[str, float]:
    """Recall@k / rank statistics; restates train.py:457-508 with vectorised rank search.

    similarity_scores is [N_img, N_txt] (i2t); its transpose is used for t2i.
    """
    i2t = similarity_scores.float()
    t2i = i2t.T
    t2i_ranks = np.zeros(t2i.shape[0])
    for t, row in enumerate(t2i):
        order = torch.argsort(row, descending=True)
        t2i_ranks[t] = int((order == txt2img[t]).nonzero()[0, 0])
    i2t_ranks = np.zeros(i2t.shape[0])
    for i, row in enumerate(i2t):
        order = torch.argsort(row, descending=True)
        pos = torch.empty_like(order)
        pos[order] = torch.arange(order.numel())
        i2t_ranks[i] = min((int(pos[j]) for j in img2txt[i]), default=1e10)      # train.py:485: rank = 1e10

    def rep(r, name):
        return {f"{name}_R@1": float((r < 1).mean()), f"{name}_R@5": float((r < 5).mean()),
                f"{name}_R@10": float((r < 10).mean()), f"{name}_mean_rank": float(r.mean() + 1),
                f"{name}_median_rank": float(math.floor(np.median(r)) + 1)}

    return {**rep(t2i_ranks, "text_to_image"), **rep(i2t_ranks, "image_to_text")}


def synth_inputs(B: int, P: int, T: int, D: int, seed: int = 1234, pad_text: bool = False,
                 Bt: Optional[int] = None):
    """SURVEY 8(d) synthetic inputs: L2-normalised Gaussian rows, optional zero-padded tails."""
    g = torch.Generator().manual_seed(seed)
    Bt = B if Bt is None else Bt
    tok_img = F.normalize(torch.randn(B, P, D, generator=g), dim=-1)
    tok_txt = F.normalize(torch.randn(Bt, T, D, generator=g), dim=-1)
    img = F.normalize(torch.randn(B, D, generator=g), dim=-1)
    txt = F.normalize(torch.randn(Bt, D, generator=g), dim=-1)
    if pad_text:
        lens = torch.randint(min(4, T), T + 1, (Bt,), generator=g)
        mask = torch.arange(T)[None, :] < lens[:, None]
        tok_txt = tok_txt * mask[..., None]
    return img, txt, tok_img, tok_txt
