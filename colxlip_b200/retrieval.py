"""Retrieval ranking on the device (reference train.py:457-508, `compute_retrieval`).

Same inputs and report keys as the reference; the per-row `argsort` + `where` loop is replaced
by one rank-of-target kernel launch per direction (cxl_rank_of_target).
"""
from __future__ import annotations

import math
from typing import Dict, Sequence

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


def compute_retrieval(similarity_scores, txt2img: Sequence[int], img2txt: Sequence[Sequence[int]], kernels=None):
    """similarity_scores: [N_img, N_txt] (i2t) or a tuple (i2t, t2i) as in the reference."""
    k = kernels if kernels is not None else default_kernels()
    if isinstance(similarity_scores, tuple):
        i2t, t2i = similarity_scores
    else:
        i2t, t2i = similarity_scores, similarity_scores.t()
    dev = i2t.device
    t2i_ranks = k.rank_of_target(t2i, torch.as_tensor(list(txt2img), dtype=torch.int32, device=dev))
    # image -> text: best (lowest) rank over the image's ground-truth captions
    width = max(len(x) for x in img2txt)
    best = None
    for j in range(width):
        tgt = torch.as_tensor([x[min(j, len(x) - 1)] for x in img2txt], dtype=torch.int32, device=dev)
        r = k.rank_of_target(i2t, tgt)
        best = r if best is None else torch.minimum(best, r)
    t2i_np = t2i_ranks.cpu().numpy().astype(np.float64)
    i2t_np = best.cpu().numpy().astype(np.float64)
    return {**_report(t2i_np, "text_to_image"), **_report(i2t_np, "image_to_text")}
