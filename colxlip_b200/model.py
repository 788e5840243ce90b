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


def eot_valid_len(text: torch.Tensor) -> torch.Tensor:
    """Number of leading text tokens that keep their features: the reference zeroes the token at the EOT position
    (`text.argmax(dim=-1)`, text_pool_type == 'argmax') and every token after it (model.py:580-593)."""
    return text.argmax(dim=-1).to(torch.int32)


def _torch_head(x, head, valid_len, normalize):
    """The reference's own op sequence (model.py:589-593, 602-607; the nn.Sequential in functional form) -- used for
    the backward (recompute) and for the cases the fused launches do not serve.  head = (ln1.weight, ln1.bias,
    linear.weight, linear.bias, ln2.weight, ln2.bias, ln1.eps, ln2.eps)."""
    if valid_len is not None:
        pos = torch.arange(x.shape[1], device=x.device)[None, :]
        x = x * (pos < valid_len[:, None].to(x.device)).to(x.dtype)[..., None]
    ln1_w, ln1_b, lin_w, lin_b, ln2_w, ln2_b, e1, e2 = head
    y = torch.nn.functional.layer_norm(x, (x.shape[-1],), ln1_w, ln1_b, e1)
    y = torch.nn.functional.gelu(torch.nn.functional.linear(y, lin_w, lin_b))
    y = torch.nn.functional.layer_norm(y, (y.shape[-1],), ln2_w, ln2_b, e2)
    return torch.nn.functional.normalize(y, dim=-1) if normalize else y


class _TokenHeadFn(torch.autograd.Function):
    """Forward: three launches (cxl_head_ln_in -> cxl_gemm_nt -> cxl_head_out).  Backward: the head is recomputed with
    torch ops on the saved input and differentiated by autograd (activation checkpointing: nothing but the input is
    kept; the towers' gradient is the reference's own)."""

    @staticmethod
    def forward(ctx, x, valid_len, normalize, out_dtype, op_dtype, kernels, eps, *params):
        y = kernels.token_head(x.detach(), *[p.detach() for p in params], out_dtype=out_dtype, op_dtype=op_dtype,
                               valid_len=valid_len, normalize=normalize, eps1=eps[0], eps2=eps[1])
        if y is None:
            raise RuntimeError("token head widths outside the fused kernels")       # caller checks supported() first
        ctx.save_for_backward(x, *params)
        ctx.meta = (valid_len, normalize, eps, op_dtype)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        valid_len, normalize, eps, op_dtype = ctx.meta
        x, *params = ctx.saved_tensors
        with torch.enable_grad():
            xs = x.detach().requires_grad_(x.requires_grad)
            ps = [p.detach().requires_grad_(p.requires_grad) for p in params]
            with torch.autocast("cuda", dtype=op_dtype):
                y = _torch_head(xs, (*ps, eps[0], eps[1]), valid_len, normalize)
            need = [t for t in [xs] + ps if t.requires_grad]
            grads = torch.autograd.grad(y, need, g.to(y.dtype)) if need else []
        it = iter(grads)
        out = [next(it) if t.requires_grad else None for t in [xs] + ps]
        return (out[0], None, None, None, None, None, None, *out[1:])


def token_head_supported(width: int, D: int) -> bool:
    return width % 256 == 0 and width <= 1024 and D % 256 == 0 and D <= 1024


def encode_tokens(tokens, head, valid_len=None, normalize=True, out_dtype=None, kernels=None):
    """`F.normalize(head(masked(tokens)), dim=-1)` of ColXLIP.encode_image / encode_text (model.py:543-554, 589-607) with
    `head` = the reference's nn.Sequential(LayerNorm, Linear, GELU, LayerNorm) (its parameters are used as they are).
    tokens [N, T, width]; valid_len [N] (see eot_valid_len) or None; the result [N, T, D] is written in the dtype the
    MaxSim kernels consume (the autocast dtype, else the inputs' 16-bit dtype, else fp32)."""
    ln1, lin, ln2 = head[0], head[1], head[3]
    hp = (ln1.weight, ln1.bias, lin.weight, lin.bias, ln2.weight, ln2.bias, ln1.eps, ln2.eps)
    D, width = lin.weight.shape
    if out_dtype is None:
        if tokens.is_cuda and torch.is_autocast_enabled("cuda"):
            out_dtype = torch.get_autocast_dtype("cuda")
        else:
            out_dtype = tokens.dtype if tokens.dtype in (torch.bfloat16, torch.float16) else torch.float32
    if not tokens.is_cuda or not token_head_supported(width, D) or out_dtype == torch.float32:
        # CPU models, odd widths, and fp32 results (no autocast: the caller asked for fp32 arithmetic, the fused GEMM
        # multiplies 16-bit operands) run the reference's own torch ops
        return _torch_head(tokens, hp, valid_len, normalize).to(out_dtype)
    op_dtype = out_dtype if out_dtype in (torch.bfloat16, torch.float16) else torch.bfloat16
    k = kernels if kernels is not None else default_kernels()
    return _TokenHeadFn.apply(tokens, valid_len, bool(normalize), out_dtype, op_dtype, k, (ln1.eps, ln2.eps),
                              ln1.weight, ln1.bias, lin.weight, lin.bias, ln2.weight, ln2.bias)


class TokenHead(torch.nn.Sequential):
    """Drop-in for the reference's `vision_token_layer` / `text_token_layer` (model.py:518-530): same children, same
    state_dict keys; `forward(tokens, valid_len=None, normalize=False)` runs the fused launches on CUDA."""

    def __init__(self, width: int, embed_dim: int):
        super().__init__(torch.nn.LayerNorm(width), torch.nn.Linear(width, embed_dim), torch.nn.GELU(),
                         torch.nn.LayerNorm(embed_dim))

    def forward(self, tokens, valid_len=None, normalize=False, out_dtype=None):
        return encode_tokens(tokens, self, valid_len, normalize, out_dtype)


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
