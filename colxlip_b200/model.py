"""Scoring method of the reference's ColXLIP model class on the B200 path.

The reference defines `ColXLIP.compute_colbert_similarity(self, token_image_features,
token_text_features)` (model.py:611-629): the same contraction + max over patches as the loss,
followed by a plain mean over all T text tokens.  The encoders themselves are out of scope
(SURVEY 2, rows 7-9); a ColXLIP model picks the method up through this mixin, or the function can be
called directly for retrieval scoring.
"""
from __future__ import annotations

import torch
from torch.autograd.function import once_differentiable

from .kernels import default_kernels
from .loss import compute_colbert_similarity_mean


class _L2NormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, out_dtype, kernels):
        xc = x.detach()
        if xc.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            xc = xc.float()
        if xc.dtype != torch.float32 and out_dtype not in (xc.dtype, torch.float32):
            xc = xc.float()                            # bf16 <-> fp16: through fp32 (no such kernel pair)
        y, inv = kernels.l2norm_fwd(xc.contiguous(), out_dtype, save_inv=x.requires_grad)
        if x.requires_grad:
            ctx.save_for_backward(y, inv)
        ctx.meta = (kernels, x.dtype)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        kernels, dt = ctx.meta
        y, inv = ctx.saved_tensors
        if g.dtype != torch.float32 and g.dtype != y.dtype:
            g = g.float()
        out_dt = dt if dt in (torch.float32, torch.bfloat16, torch.float16) else torch.float32
        if y.dtype == torch.float32 and g.dtype == torch.float16:
            g = g.float()
        return kernels.l2norm_bwd(y, g.contiguous(), inv, out_dt).to(dt), None, None


def normalize_tokens(x, out_dtype=None, kernels=None):
    """F.normalize(x, dim=-1) as the reference applies it to the token / global embeddings
    (model.py:551-554, 605-607), fused with the cast to the dtype the MaxSim kernels consume.

    out_dtype=None: the autocast dtype under CUDA autocast (bfloat16, or float16 for `--precision amp`), the
    input's dtype for 16-bit inputs (what torch.einsum at loss.py:34 would cast its operand to), else float32.
    A 16-bit result goes into ColClipLoss / compute_colbert_similarity without a
    further pack pass, and into the feature all-gather at half the fp32 bytes.  x: [..., D], D % 8 == 0."""
    if x.dim() < 1 or x.shape[-1] % 8 != 0:
        raise ValueError("normalize_tokens needs an embedding dimension that is a multiple of 8 (16-byte row accesses)")
    k = kernels if kernels is not None else default_kernels()
    if out_dtype is None:
        if x.is_cuda and torch.is_autocast_enabled("cuda"):
            out_dtype = torch.get_autocast_dtype("cuda")
        elif x.dtype in (torch.bfloat16, torch.float16):
            out_dtype = x.dtype
        else:
            out_dtype = torch.float32
    return _L2NormFn.apply(x, out_dtype, k)


class ColXLIPScoringMixin:
    """Mix into (or monkey-patch onto) the reference's ColXLIP class."""

    def compute_colbert_similarity(self, token_image_features, token_text_features):
        """[Bt,Bi] plain-mean MaxSim scores, values in [-1, 1] for normalised features."""
        return compute_colbert_similarity_mean(token_image_features, token_text_features)


    @staticmethod
    def normalize_tokens(x):
        return normalize_tokens(x)


def patch_colxlip(model_cls):
    """Install the B200 scorer on the reference's model class: patch_colxlip(colxlip.model.ColXLIP)."""
    model_cls.compute_colbert_similarity = ColXLIPScoringMixin.compute_colbert_similarity
    return model_cls
