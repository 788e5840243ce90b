"""Retrieval ranking on the device (reference train.py:457-508, `compute_retrieval`).

Same inputs and report keys as the reference; the per-row `argsort` + `where` loop is replaced
by one rank-of-target kernel launch per direction (cxl_rank_of_target).
"""
from __future__ import annotations

import math
from typing import Dict

import numpy as np
import torch

from .kernels import default_kernels


def _report(ranks: np.ndarray, name: str) -> Dict[str, float]:
    n = len(ranks)
    return {
        f"{name}_R@1": float((ranks < 1).sum()) / n,
        f"{name}_R@5": float((ranks < 5).sum()) / n,
        f"{name}_R@10": float((ranks < 10).sum()) / n,
        f"{name}_mean_rank": float(ranks.mean()) + 1,
        f"{name}_median_rank": float(math.floor(np.median(ranks))) + 1,
    }


def _by_position(x, n: int, what: str):
    """The reference indexes txt2img / img2txt by POSITION (`txt2img[index]`, train.py:469,489): dicts keyed
    0..n-1 (what remap_indices returns) and sequences are both accepted."""
    if isinstance(x, dict):
        try:
            return [x[i] for i in range(n)]
        except KeyError as e:
            raise KeyError(f"compute_retrieval: {what} has no entry for position {e.args[0]} (expected keys 0..{n - 1})") from e
    x = list(x)
    if len(x) != n:
        raise ValueError(f"compute_retrieval: {what} has {len(x)} entries for {n} score rows")
    return x


def compute_retrieval(similarity_scores, txt2img, img2txt, kernels=None):
    """similarity_scores: [N_img, N_txt] (i2t) or a tuple (i2t, t2i) as in the reference (train.py:457-508).
    txt2img: position -> gallery image index; img2txt: position -> list of caption indices (dicts keyed by
    position or sequences).  An image without captions keeps the reference's sentinel rank 1e10; an index
    outside the score matrix raises IndexError (the reference's `torch.where(...)[0][0]` does)."""
    k = kernels if kernels is not None else default_kernels()
    if isinstance(similarity_scores, tuple):
        i2t, t2i = similarity_scores
    else:
        i2t, t2i = similarity_scores, similarity_scores.t()
    dev = i2t.device
    n_txt, n_gal = t2i.shape
    n_img, n_cap = i2t.shape
    t_tgt = np.asarray([int(v) for v in _by_position(txt2img, n_txt, "txt2img")], dtype=np.int64)
    if t_tgt.size and (t_tgt.min() < 0 or t_tgt.max() >= n_gal):
        raise IndexError(f"compute_retrieval: txt2img holds an image index outside [0, {n_gal})")
    caps = [[int(c) for c in x] for x in _by_position(img2txt, n_img, "img2txt")]
    for x in caps:
        if any(c < 0 or c >= n_cap for c in x):
            raise IndexError(f"compute_retrieval: img2txt holds a caption index outside [0, {n_cap})")
    t2i_ranks = k.rank_of_target(t2i, torch.as_tensor(t_tgt, dtype=torch.int32, device=dev))
    # image -> text: best (lowest) rank over the image's ground-truth captions
    width = max((len(x) for x in caps), default=0)
    has_cap = np.asarray([len(x) > 0 for x in caps], dtype=bool)
    best = None
    for j in range(width):
        tgt = torch.as_tensor([x[min(j, len(x) - 1)] if x else 0 for x in caps], dtype=torch.int32, device=dev)
        r = k.rank_of_target(i2t, tgt)
        best = r if best is None else torch.minimum(best, r)
    t2i_np = t2i_ranks.cpu().numpy().astype(np.float64)
    i2t_np = best.cpu().numpy().astype(np.float64) if best is not None else np.zeros(n_img)
    i2t_np = np.where(has_cap, i2t_np, 1e10)                       # train.py:485: `rank = 1e10` when the loop is empty
    return {**_report(t2i_np, "text_to_image"), **_report(i2t_np, "image_to_text")}


# ---------------------------------------------------------------------------------------------------
# MaxSim retrieval loop.  The reference's retrieval_on_split (train.py:510-585) is a leftover
# of its FLAIR ancestry: it scores global embeddings only and calls model methods ColXLIP does not define.
# The functions below keep its call shape (loaders + index dicts -> metrics dict with the same keys) and
# score with the model's own token-level scorer (model.py:611-629) on the K1 kernel.
# ---------------------------------------------------------------------------------------------------
def _mean_scores(token_image_features, token_text_features, kernels=None):
    from .loss import _maxsim
    from .kernels import MEAN
    return _maxsim(token_image_features, token_text_features, MEAN, kernels)


def maxsim_score_matrix(token_image_features, token_text_features, image_chunk: int = 8192, text_chunk: int = 8192,
                        kernels=None):
    """i2t score matrix [N_img, N_txt] = compute_colbert_similarity(...).T (model.py:626-628), built block by
    block so that neither operand copy nor the [chunk, chunk] score block grows with the gallery."""
    n_img, n_txt = token_image_features.shape[0], token_text_features.shape[0]
    out = torch.empty((n_img, n_txt), dtype=torch.float32, device=token_image_features.device)
    with torch.no_grad():
        for t0 in range(0, n_txt, text_chunk):
            txt = token_text_features[t0:t0 + text_chunk]
            for i0 in range(0, n_img, image_chunk):
                s = _mean_scores(token_image_features[i0:i0 + image_chunk], txt, kernels)   # [txt, img]
                out[i0:i0 + image_chunk, t0:t0 + text_chunk] = s.t()
    return out


def _unwrap_model(model):
    return model.module if hasattr(model, "module") else model            # train.py:61-65


def remap_indices(merged_img_ids, cap_ids, img2txt_dict, txt2img_dict):
    """Gallery order -> positions (train.py:430-454): image ids are renumbered by their position in the
    image loader's output; caption ids are expected to be 0..N-1 already."""
    mapping = {int(old): new for new, old in enumerate(merged_img_ids)}
    new_img2txt = {mapping[i]: list(caps) for i, caps in img2txt_dict.items()}
    new_txt2img = {c: mapping[txt2img_dict[c][0]] for c in txt2img_dict.keys()}
    return new_img2txt, new_txt2img


def compute_similarity_scores_maxsim(model, img_loader, all_text_token_features, device, input_dtype, autocast,
                                     kernels=None):
    """Token-level counterpart of compute_similarity_scores_original_clip (train.py:587-614): stream the image
    loader, score every image batch against ALL text tokens, return (scores [M, N] on `device`, img_ids)."""
    blocks, all_img_ids = [], []
    m = _unwrap_model(model)
    for images, img_id in img_loader:
        images = images.to(device=device, dtype=input_dtype, non_blocking=True)
        all_img_ids.append(img_id.detach().cpu())
        with autocast():
            _, image_tokens = m.encode_image(images, normalize=True)
            logit_scale = m.logit_scale.exp()
            s = _mean_scores(image_tokens, all_text_token_features, kernels)                         # [N, b]
        blocks.append((logit_scale.float() * s.t()).detach())
    return torch.cat(blocks), torch.cat(all_img_ids)


def retrieval_on_split(keyword, model, txt_loader, img_loader, img2txt_dict, txt2img_dict, args, epoch, metrics,
                       device, input_dtype, autocast, kernels=None):
    """Same signature and metric keys as the reference (train.py:510-585); text tokens stay on the device."""
    num_txt_samples = txt_loader.num_samples
    num_img_samples = img_loader.num_samples
    m = _unwrap_model(model)
    text_tokens, cap_ids = [], []
    with torch.no_grad():
        for texts, cap_id in txt_loader:
            texts = texts.to(device=device, non_blocking=True)
            with autocast():
                _, tok = m.encode_text(texts, normalize=True)
            text_tokens.append(tok.detach())
            cap_ids.append(cap_id.detach().cpu())
        all_text_tokens = torch.cat(text_tokens)
        cap_ids = torch.cat(cap_ids)
        scores, img_ids = compute_similarity_scores_maxsim(m, img_loader, all_text_tokens, device, input_dtype,
                                                           autocast, kernels)
        new_img2txt, new_txt2img = remap_indices(img_ids, cap_ids, img2txt_dict, txt2img_dict)
        retrieval_metrics = compute_retrieval(scores, new_txt2img, new_img2txt, kernels=kernels)
    if keyword != '':
        keyword = keyword + '_'
        retrieval_metrics = {keyword + k: v for k, v in retrieval_metrics.items()}
    if "epoch" not in metrics:
        metrics.update({"epoch": epoch})
    metrics.update({**retrieval_metrics, f"{keyword}num_text_samples": num_txt_samples,
                    f"{keyword}num_image_samples": num_img_samples})
    return metrics
