"""Drop-in mirror of the reference's loss/scoring API for the MaxSim hot path.

Same names, argument meaning and error behaviour as /root/reference/src/colxlip/loss.py:
  compute_colbert_similarity  (loss.py:20-46)
  gather_features             (loss.py:48-92)
  ColClipLoss                 (loss.py:184-296)
but the arithmetic runs in hand-written sm_100a CUDA kernels behind the C-ABI of
include/colxlip_b200.h (K1 tcgen05 MaxSim forward, K2 argmax-driven backward, K3 fused
scale/softmax/cross-entropy).  The [Bt,Bi,T,P] similarity tensor is never materialised.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn as nn
from torch.autograd.function import once_differentiable

try:
    import torch.distributed.nn
    from torch import distributed as dist

    has_distributed = True
except ImportError:  # pragma: no cover
    has_distributed = False
    dist = None

from .kernels import MASKED, MEAN, ROUND16, default_kernels

# "auto": tensor-core path in the 16-bit dtype torch.einsum would use at loss.py:34 -- the autocast dtype under CUDA
# autocast (bf16, or fp16 for the reference's `--precision amp`, colxlip.sh:53 / train.py:131-134), else the
# inputs' own 16-bit dtype; fp32 SIMT path for fp32 inputs outside autocast.  "bf16"/"fp16"/"fp32" force one path.
_PRECISION = os.environ.get("COLXLIP_B200_PRECISION", "auto")


def set_precision(mode: str) -> None:
    global _PRECISION
    if mode not in ("auto", "bf16", "fp16", "fp32"):
        raise ValueError("precision must be auto|bf16|fp16|fp32")
    _PRECISION = mode


def get_precision() -> str:
    return _PRECISION


# Strict autocast arithmetic (off by default): under torch.autocast the reference's einsum (loss.py:34) returns 16-bit
# similarities, so torch.max, the != 0 mask and the sum see row maxima ROUNDED to bf16 / fp16.  With the switch on K1
# rounds every row maximum the same way (CXL_MEAN_ROUND16) and the scores agree with the reference's autocast pipeline
# to fp32 summation order instead of to the 16-bit tolerance; off, the maxima stay in fp32 (closer to the fp32 oracle).
_STRICT_AUTOCAST = os.environ.get("COLXLIP_B200_STRICT_AUTOCAST", "0") not in ("", "0")


def set_strict_autocast(on: bool) -> None:
    global _STRICT_AUTOCAST
    _STRICT_AUTOCAST = bool(on)


def get_strict_autocast() -> bool:
    return _STRICT_AUTOCAST


def _fwd_mode(mean_mode: int, prec: str) -> int:
    """mean_mode as K1 takes it: + ROUND16 when strict autocast arithmetic is on and the operands are 16-bit"""
    return (mean_mode | ROUND16) if (_STRICT_AUTOCAST and prec in ("bf16", "fp16")) else mean_mode


def _auto_precision(*tensors: torch.Tensor) -> str:
    t0 = tensors[0]
    if t0.is_cuda and torch.is_autocast_enabled("cuda"):
        return "fp16" if torch.get_autocast_dtype("cuda") == torch.float16 else "bf16"
    if all(t.dtype == torch.float16 for t in tensors):
        return "fp16"
    if all(t.dtype in (torch.bfloat16, torch.float16) for t in tensors):
        return "bf16"                                   # mixed 16-bit inputs: torch promotes bf16 x fp16 to fp32; bf16 operands here
    return "fp32"


def _select_precision(*tensors: torch.Tensor) -> str:
    D = tensors[0].shape[-1]
    mode = _PRECISION
    if mode == "auto":
        mode = _auto_precision(*tensors)
    if mode in ("bf16", "fp16") and D % 64 != 0:
        mode = "fp32"  # the tcgen05 kernel tiles D in 64-wide swizzled blocks; fp32 SIMT kernel covers the rest
    return mode


def _w_scale_from_bound(bound: torch.Tensor) -> torch.Tensor:
    """[s, 1/s] on the device, s = the power of two that puts `bound` (an upper bound of |w|, a device scalar) at
    2^15: fp16 pair weights then use the top of their range whatever the loss scale is (no host sync)."""
    e = torch.ceil(torch.log2(bound.float().abs().clamp_min(1e-30)))
    s = torch.exp2(15.0 - e).reshape(1)
    return torch.cat([s, 1.0 / s]).contiguous()


def _check_tokens(img: torch.Tensor, txt: torch.Tensor) -> None:
    if img.dim() != 3 or txt.dim() != 3 or img.shape[2] != txt.shape[2]:
        raise ValueError("expected token_image_features [Bi,P,D] and token_text_features [Bt,T,D] with equal D")
    if min(img.shape) == 0 or min(txt.shape) == 0:
        raise ValueError("empty token tensors are not supported")


class _MaxSimFn(torch.autograd.Function):
    """S[Bt,Bi] = pooled max-over-patches similarity; backward from the saved argmax (K1 + K2)."""

    @staticmethod
    def forward(ctx, img_tok, txt_tok, mean_mode, prec, kernels, grad_mode=True):
        k = kernels
        Bi, P, D = img_tok.shape
        Bt, T, _ = txt_tok.shape
        # grad_mode = torch.is_grad_enabled() at the call site (always False in here): no argmax under no_grad
        need_grad = bool(grad_mode and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]))
        txt_op = k.prepare(txt_tok, T, prec)
        img_op = k.prepare(img_tok, k.img_row_stride(P, prec), prec)
        S, cnt, argmax = k.maxsim_fwd(txt_op, img_op, T, P, prec, _fwd_mode(mean_mode, prec), need_grad)
        if need_grad:
            ctx.save_for_backward(txt_op, img_op, argmax, cnt)
        ctx.meta = (T, P, prec, mean_mode, k, img_tok.dtype, txt_tok.dtype)
        return S

    @staticmethod
    @once_differentiable
    def backward(ctx, dS):
        T, P, prec, mean_mode, k, img_dt, txt_dt = ctx.meta
        txt_op, img_op, argmax, cnt = ctx.saved_tensors
        need_img, need_txt = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        w = k.pair_weights(dS, cnt, T, mean_mode)
        if prec == "fp16":
            # fp16 pair weights: scale by the power of two that puts max|w| at 2^15, undone exactly at the store
            D = txt_op.shape[2]
            dtxt = torch.empty((txt_op.shape[0], T, D), dtype=torch.float32, device=txt_op.device) if need_txt else None
            dimg = k.maxsim_bwd_shard(txt_op, img_op, argmax, w, 0, T, P, prec, dtxt=dtxt, need_img=need_img,
                                      w_scale=_w_scale_from_bound(w.abs().amax()))
        else:
            dtxt, dimg = k.maxsim_bwd(txt_op, img_op, argmax, w, T, P, prec, need_txt=need_txt, need_img=need_img)
        return (dimg.to(img_dt) if need_img else None, dtxt.to(txt_dt) if need_txt else None, None, None, None, None)


def _maxsim(token_image_features, token_text_features, mean_mode, kernels=None):
    _check_tokens(token_image_features, token_text_features)
    k = kernels if kernels is not None else default_kernels()
    prec = _select_precision(token_image_features, token_text_features)
    return _MaxSimFn.apply(token_image_features, token_text_features, mean_mode, prec, k, torch.is_grad_enabled())


def compute_colbert_similarity(token_image_features, token_text_features):
    """Token-level late-interaction score, masked-mean variant (reference loss.py:20-46).

    Args:
        token_image_features: [batch_size_img, n_img_tokens, embed_dim]
        token_text_features:  [batch_size_txt, n_txt_tokens, embed_dim]
    Returns:
        [batch_size_txt, batch_size_img] fp32: for every text token the best patch similarity,
        summed over tokens and divided by (number of non-zero maxima + 1e-8).
    """
    return _maxsim(token_image_features, token_text_features, MASKED)


def compute_colbert_similarity_mean(token_image_features, token_text_features):
    """Plain-mean variant used for retrieval inference (reference model.py:611-629)."""
    return _maxsim(token_image_features, token_text_features, MEAN)


def gather_features(
        image_features,
        text_features,
        local_loss=False,
        gather_with_grad=False,
        rank=0,
        world_size=1,
        use_horovod=False
):
    """Concatenate every rank's features along dim 0 in rank order (reference loss.py:48-92).

    gather_with_grad=True : autograd-aware all_gather (backward = SUM reduce-scatter).
    gather_with_grad=False: plain all_gather under no_grad; unless local_loss, this rank's slot is
    replaced by the live local tensor so local gradients still flow.
    """
    assert has_distributed, 'torch.distributed did not import correctly, please use a PyTorch version with support.'
    if use_horovod:
        raise RuntimeError("horovod is not supported by colxlip_b200 (the reference asserts 'Please install horovod')")
    if gather_with_grad:
        all_image_features = torch.cat(torch.distributed.nn.all_gather(image_features), dim=0)
        all_text_features = torch.cat(torch.distributed.nn.all_gather(text_features), dim=0)
        return all_image_features, all_text_features
    with torch.no_grad():
        img_slots = [torch.zeros_like(image_features) for _ in range(world_size)]
        txt_slots = [torch.zeros_like(text_features) for _ in range(world_size)]
        dist.all_gather(img_slots, image_features.contiguous())
        dist.all_gather(txt_slots, text_features.contiguous())
    if not local_loss:
        img_slots[rank] = image_features
        txt_slots[rank] = text_features
    return torch.cat(img_slots, dim=0), torch.cat(txt_slots, dim=0)


# ---------------------------------------------------------------------------------------------------
# Fused loss: rows of the B x B logit matrices are sharded over ranks, columns are global.
# ---------------------------------------------------------------------------------------------------
def _all_gather_cat(x: torch.Tensor, world_size: int, group) -> torch.Tensor:
    if world_size == 1:
        return x
    x = x.contiguous()
    out = torch.empty((world_size * x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x, group=group)
    return out


def _reduce_scatter_rows(x_all: torch.Tensor, world_size: int, group) -> torch.Tensor:
    """[W*b, ...] partial sums on every rank -> [b, ...] fully summed rows of this rank."""
    if world_size == 1:
        return x_all
    b = x_all.shape[0] // world_size
    x_all = x_all.contiguous()
    if dist.get_backend(group) == "gloo":  # CPU test harness: gloo has no reduce_scatter
        dist.all_reduce(x_all, group=group)
        r = dist.get_rank(group)
        return x_all[r * b:(r + 1) * b].clone()
    out = torch.empty((b,) + tuple(x_all.shape[1:]), dtype=x_all.dtype, device=x_all.device)
    dist.reduce_scatter_tensor(out, x_all, op=dist.ReduceOp.SUM, group=group)
    return out


def _dimg_chunk_images(b: int, P: int, D: int) -> int:
    """Images per dImg partial that goes home in one SUM-reduce (multi-GPU backward).  The fp32 partial of a whole
    home-rank shard is 14.5 GB at cfg-5; CXL_DIMG_REDUCE_MB (default 4096) bounds it, two partials are in flight."""
    limit = max(1, int(float(os.environ.get("CXL_DIMG_REDUCE_MB", "4096")) * (1 << 20)))
    n_chunks = max(1, -(-(b * P * D * 4) // limit))
    return -(-b // n_chunks)


_COMM_STREAMS = {}

# ---- per-phase CUDA-event split of a step (bench.py reports it as "phases"; off unless a recorder is installed) ----
_PHASE_REC = None


def phase_recording(on: bool = True):
    """Install / remove the phase recorder; `phase_times_ms()` after a synchronize() returns [(name, ms since the
    previous mark)] of the marks recorded on the current stream since the recorder was installed."""
    global _PHASE_REC
    _PHASE_REC = [] if on else None


# NVTX ranges over the same phase boundaries (tracing): visible in Nsight Systems / torch profiler timelines,
# a few hundred nanoseconds per call otherwise.  COLXLIP_B200_NVTX=0 turns them off.  The phase marks name the interval
# that just ENDED; the table names the range that STARTS at each mark (None: nothing of this path follows).
_NVTX = os.environ.get("COLXLIP_B200_NVTX", "1") not in ("", "0")
_NVTX_NEXT = {
    "step_start": "cxl: fwd shard gather + K1 (MaxSim)",
    "fwd: gather issue + K1 (all shards)": "cxl: fwd global GEMM + K3 statistics",
    "fwd: global GEMM + K3 statistics + column-LSE exchange": None,
    "(between forward and backward)": "cxl: bwd K3 gradient + global gradient GEMMs",
    "bwd: K3 gradient + global gradient GEMMs + their reduce-scatter": "cxl: bwd K2 (dTxt, dImg) + dImg reductions",
    "bwd: K2 over all shards (dImg reductions overlapped on the side stream)": "cxl: bwd wait for the last dImg reduction",
    "bwd: wait for the last dImg reduction": None,
}
_nvtx_open = False


def _mark(name: str, dev) -> None:
    global _nvtx_open
    if _NVTX and dev.type == "cuda":
        if _nvtx_open:
            torch.cuda.nvtx.range_pop()
            _nvtx_open = False
        nxt = _NVTX_NEXT.get(name)
        if nxt is not None:
            torch.cuda.nvtx.range_push(nxt)
            _nvtx_open = True
    if _PHASE_REC is not None and dev.type == "cuda":
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(dev))
        _PHASE_REC.append((name, ev))


def phase_times_ms():
    rec = _PHASE_REC or []
    return [(rec[i][0], rec[i - 1][1].elapsed_time(rec[i][1])) for i in range(1, len(rec))]


def _comm_stream(dev: torch.device):
    """Side stream for the shard broadcasts / dImg reductions, so they overlap K1 / K2 on the current stream."""
    key = (dev.type, dev.index)
    if key not in _COMM_STREAMS:
        _COMM_STREAMS[key] = torch.cuda.Stream(device=dev)
    return _COMM_STREAMS[key]


def _global_rank(group, r: int) -> int:
    return dist.get_global_rank(group, r) if group is not None else r


# K1 is a persistent kernel that fills every SM (one CTA with ~220 KB of shared memory each), so an NCCL kernel launched
# while a K1 runs is not scheduled until that K1 drains -- on THIS rank, which stalls the same broadcast on every other
# rank: with K1 started on the local shard while the remote ones were still in flight the ranks fell into a convoy
# (8 x B200, cfg-2: forward 1867-2193 ms for 1410 ms of K1 launches, another 330-550 ms of skew at the next collective;
# profiles/r4m_bench_cfg2_8gpu_overlapped_gather.json).  The broadcasts move 5.8 GB per rank at cfg-2 (50 GB at cfg-5):
# tens of milliseconds over NVSwitch against 1.4 s (32 s) of K1, so they now finish before the first K1 launch.
# CXL_GATHER_OVERLAP=1 restores the overlapped schedule (for A/B).
_OVERLAP_GATHER = os.environ.get("CXL_GATHER_OVERLAP", "0") not in ("", "0")


def _gather_shards(x: torch.Tensor, world_size: int, rank: int, group):
    """Every rank's `x` as SEPARATE tensors (shard r = rank r's x; this rank's own is `x` itself, no copy), fetched by
    one broadcast per home rank on a side stream, so there is no [W*b, ...] concatenation and the backward can free
    the shards one by one.  ready[r] is the CUDA event the consumer of shard r has to wait for (None: nothing pending
    -- the default, the current stream already waits for all of them; see _OVERLAP_GATHER).  Replaces all_gather of
    the reference's gather_features (loss.py:70-86) for the token operand."""
    shards = [x if r == rank else torch.empty_like(x) for r in range(world_size)]
    ready = [None] * world_size
    if world_size == 1:
        return shards, ready
    if x.is_cuda:
        main, comm = torch.cuda.current_stream(x.device), _comm_stream(x.device)
        comm.wait_stream(main)                       # the local operand (and the fresh buffers' memory) are ready
        with torch.cuda.stream(comm):
            for r in range(world_size):
                dist.broadcast(shards[r], src=_global_rank(group, r), group=group)
                if r != rank and _OVERLAP_GATHER:
                    ready[r] = torch.cuda.Event()
                    ready[r].record(comm)
        if _OVERLAP_GATHER:
            x.record_stream(comm)
        else:
            main.wait_stream(comm)                   # all shards have landed before the first K1 launch (and before
                                                     # anything on the current stream can free them)
    else:
        for r in range(world_size):
            dist.broadcast(shards[r], src=_global_rank(group, r), group=group)
    return shards, ready


class _FwdState:
    """What the fused forward keeps for its backward.  Held on the autograd ctx as a plain object (not through
    save_for_backward) because the backward releases the image shards and their argmax one home rank at a time."""
    __slots__ = ("txt_op", "shards", "args", "cnt", "S", "G", "g_txt", "g_img", "scale", "rl_g", "cl_g", "rl_t",
                 "cl_t", "consumed")


_OP_DTYPE = {"bf16": torch.bfloat16, "fp16": torch.float16}


def _round_like_gemm_input(x: torch.Tensor, prec: str) -> torch.Tensor:
    """The global features as the logits GEMM sees them: a 16-bit copy on the 16-bit paths (what the reference's
    autocast matmul multiplies, loss.py:252-253 -- and half the bytes in the feature all-gather), fp32 otherwise."""
    x = x.detach()
    if prec in _OP_DTYPE:
        return x.to(_OP_DTYPE[prec]).contiguous()
    return x.float()


def _logits_gemm(k, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """a [M,D] @ b [N,D]^T -> fp32 [M,N].  16-bit operands run on the tcgen05 GEMM (cxl_gemm_nt); fp32 operands (the
    fp32 parity mode) and shapes the kernel does not take (D % 64 != 0) on the library GEMM."""
    if a.dtype != torch.float32:
        c = k.gemm_nt(a, b)
        if c is not None:
            return c
        return a.float() @ b.float().T
    return a @ b.T


def _grad_gemms(k, dX: torch.Tensor, rows_op: Optional[torch.Tensor], cols_op: Optional[torch.Tensor]):
    """(dX @ cols_op, dX^T @ rows_op) for dX [m,n] fp32 = dL/dlogits: the two gradient GEMMs of a logits block
    X = rows_op [m,D] @ cols_op [n,D]^T (either operand None: that product is skipped).  On the 16-bit paths dX is
    cast to the operand dtype -- the reference's autocast backward multiplies a 16-bit dlogits too -- in one
    cast+transpose pass, and both products run on the tcgen05 GEMM."""
    op = cols_op if cols_op is not None else rows_op
    if op is None:
        return None, None
    if op.dtype == torch.float32:
        return (dX @ cols_op if cols_op is not None else None), (dX.T @ rows_op if rows_op is not None else None)
    sc = None
    if op.dtype == torch.float16:
        # fp16 copy of dL/dlogits: a power of two puts its largest entry at 2^15 (small entries stay out of the
        # subnormal range whatever the loss scale is); undone exactly on the [.,D] results
        sc = _w_scale_from_bound(dX.abs().amax())
    d16, d16t = k.cast_transpose(dX, op.dtype, want_plain=cols_op is not None, want_t=rows_op is not None,
                                 scale=None if sc is None else sc[0:1])
    g_rows = g_cols = None
    if cols_op is not None:                               # [m,n] x [n,D]: contraction over n
        g_rows = k.gemm_nt(d16, cols_op.t().contiguous())
        if g_rows is None:
            g_rows = d16.float() @ cols_op.float()
    if rows_op is not None:                               # [n,m] x [m,D]: contraction over m
        g_cols = k.gemm_nt(d16t, rows_op.t().contiguous())
        if g_cols is None:
            g_cols = d16t.float() @ rows_op.float()
    if sc is not None:
        g_rows = g_rows * sc[1] if g_rows is not None else None
        g_cols = g_cols * sc[1] if g_cols is not None else None
    return g_rows, g_cols


class _ColClipLossFn(torch.autograd.Function):
    """losses[3] = (global, token, total) of ColClipLoss.forward (reference loss.py:271-296)."""

    @staticmethod
    def forward(ctx, image_features, text_features, token_image_features, token_text_features, logit_scale, cfg):
        # precision is decided by cfg["prec"]; the few torch ops in here must not be re-cast by an ambient autocast
        # validation / no_grad calls save nothing: no argmax buffer (82 GB at cfg-2) and no argmax stores in K1
        cfg = dict(cfg, need_bwd=bool(cfg.get("grad_mode", True) and any(ctx.needs_input_grad[:5])))
        with torch.autocast(token_image_features.device.type, enabled=False):
            return _ColClipLossFn._forward(ctx, image_features, text_features, token_image_features,
                                           token_text_features, logit_scale, cfg)

    @staticmethod
    def _forward(ctx, image_features, text_features, token_image_features, token_text_features, logit_scale, cfg):
        k = cfg["kernels"]
        W, rank, group = cfg["world_size"], cfg["rank"], cfg.get("group")
        alpha, prec = float(cfg["alpha"]), cfg["prec"]
        b, P, D = token_image_features.shape
        _, T, _ = token_text_features.shape
        B = b * W
        row0 = rank * b
        dev = token_image_features.device

        # global features first (small collectives), then the token operand: one broadcast per home rank (K4, forward
        # half), all landed before K1 starts (see _OVERLAP_GATHER)
        _mark("step_start", dev)
        g_txt = _round_like_gemm_input(text_features, prec)
        g_img = _all_gather_cat(_round_like_gemm_input(image_features, prec), W, group)
        txt_op = k.prepare(token_text_features, T, prec)
        shards, ready = _gather_shards(k.prepare(token_image_features, k.img_row_stride(P, prec), prec), W, rank, group)

        # K1: token-level scores of this rank's texts against all images, one home rank's shard at a time
        need_bwd = bool(cfg.get("need_bwd", True))
        args = [None] * W
        if W == 1:
            S, cnt, args[0] = k.maxsim_fwd(txt_op, shards[0], T, P, prec, _fwd_mode(MASKED, prec), need_bwd)
        else:
            S = torch.empty((b, B), dtype=torch.float32, device=dev)
            cnt = torch.empty((b, B), dtype=torch.float32, device=dev)
            for i in range(W):
                r = (rank + i) % W                       # own shard first: nothing to wait for
                if ready[r] is not None:
                    torch.cuda.current_stream(dev).wait_event(ready[r])
                S_r, cnt_r, args[r] = k.maxsim_fwd(txt_op, shards[r], T, P, prec, _fwd_mode(MASKED, prec), need_bwd)
                S[:, r * b:(r + 1) * b] = S_r
                cnt[:, r * b:(r + 1) * b] = cnt_r
                del S_r, cnt_r
        _mark("fwd: gather issue + K1 (all shards)", dev)
        # global CLIP scores: tcgen05 GEMM on the 16-bit paths (cxl_gemm_nt), fp32 accumulate -- SURVEY 8 f-1
        G = _logits_gemm(k, g_txt, g_img)

        scale = logit_scale.detach().float().reshape(1).to(dev)
        stats = []
        for X in (G, S):
            row_lse, col_max, col_sum, diag = k.infonce_stats(X, scale, row0)
            col_lse = col_max + torch.log(col_sum)
            part = row_lse.sum() - 2.0 * diag.sum()
            if W > 1:
                col_lse = torch.logsumexp(_all_gather_cat(col_lse.unsqueeze(0), W, group), dim=0)
                dist.all_reduce(part, group=group)
            loss = (part + col_lse.sum()) / (2.0 * B)
            stats.append((row_lse, col_lse, loss))
        loss_g, loss_t = stats[0][2], stats[1][2]
        total = alpha * loss_g + (1.0 - alpha) * loss_t
        _mark("fwd: global GEMM + K3 statistics + column-LSE exchange", dev)

        if need_bwd:
            st = _FwdState()
            st.txt_op, st.shards, st.args, st.cnt, st.S, st.G = txt_op, shards, args, cnt, S, G
            st.g_txt, st.g_img, st.scale = g_txt, g_img, scale
            st.rl_g, st.cl_g, st.rl_t, st.cl_t = stats[0][0], stats[0][1], stats[1][0], stats[1][1]
            st.consumed = False
            ctx.state = st
        ctx.cfg = cfg
        ctx.meta = (T, P, b, row0, image_features.dtype, text_features.dtype, token_image_features.dtype,
                    token_text_features.dtype, logit_scale.dtype, logit_scale.shape)
        return torch.stack([loss_g, loss_t, total])

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        # consume=True: w / dG overwrite S / G and every shard is released as soon as K2 is done with it
        return _colclip_backward(ctx.state, ctx.cfg, ctx.meta, g, ctx.needs_input_grad, consume=True) + (None,)


def _send_home(part, home, lo, hi, out_ref, cfg, shape, dtype, mult, inflight):
    """SUM-reduce one dImg partial (images lo..hi-1 of home rank `home`'s shard) to its home rank, on the side stream
    so that K2 goes on with the next images; the home rank casts the sum into rows lo..hi-1 of its output."""
    W, rank, group = cfg["world_size"], cfg["rank"], cfg.get("group")
    if part.is_cuda:
        main, comm = torch.cuda.current_stream(part.device), _comm_stream(part.device)
        comm.wait_stream(main)
        with torch.cuda.stream(comm):
            dist.reduce(part, dst=_global_rank(group, home), group=group)
            if rank == home:
                if out_ref[0] is None:
                    out_ref[0] = torch.empty(shape, dtype=dtype, device=part.device)
                out_ref[0][lo:hi].copy_(part if mult == 1.0 else part * mult)
            ev = torch.cuda.Event()
            ev.record(comm)
        part.record_stream(comm)
        inflight.append(ev)
        if len(inflight) > 1:                       # at most two partials alive: wait for the one before the last
            main.wait_event(inflight.pop(0))
    else:
        dist.reduce(part, dst=_global_rank(group, home), group=group)
        if rank == home:
            if out_ref[0] is None:
                out_ref[0] = torch.empty(shape, dtype=dtype)
            out_ref[0][lo:hi].copy_(part if mult == 1.0 else part * mult)


def _colclip_backward(st, cfg, meta, g, need, consume=False):
    """Gradients of losses[3] w.r.t. (image, text, token image, token text features, logit_scale) from the forward's
    state.  cfg["grad_rows"] restricts the token gradients to one slice of the local batch (zeros elsewhere, or just
    the slices of all four feature gradients when cfg["slice_out"] is set).  consume: this is the only backward of
    that forward -- the pair weights / dG are written over S / G and shards are freed as K2 finishes them."""
    k = cfg["kernels"]
    W, rank, group = cfg["world_size"], cfg["rank"], cfg.get("group")
    alpha, prec = float(cfg["alpha"]), cfg["prec"]
    T, P, b, row0, dt_img, dt_txt, dt_timg, dt_ttxt, dt_ls, ls_shape = meta
    if st.consumed:
        raise RuntimeError("colxlip_b200: this ColClipLoss forward was already back-propagated (its buffers were "
                           "reused in place); retain_graph / double backward are not supported by the fused loss")
    txt_op, cnt, S, G, scale = st.txt_op, st.cnt, st.S, st.G, st.scale
    D = txt_op.shape[2]
    dev = txt_op.device
    _mark("(between forward and backward)", dev)
    g = g.float()
    coef_g = (g[0] + alpha * g[2]).reshape(1).contiguous()
    coef_t = (g[1] + (1.0 - alpha) * g[2]).reshape(1).contiguous()

    # K3 gradient pass (also yields d logit_scale partials of this rank's rows)
    w, dscale_t = k.infonce_grad(S, scale, coef_t, row0, st.rl_t, st.cl_t, cnt=cnt, out=S if consume else None)
    dG, dscale_g = k.infonce_grad(G, scale, coef_g, row0, st.rl_g, st.cl_g, out=G if consume else None)
    if consume:
        st.consumed = True
        st.S = st.G = st.cnt = None
    del cnt, S, G
    dscale = dscale_t + dscale_g
    if W > 1:
        dist.all_reduce(dscale, group=group)
    # fp16 operands: |w| <= 2 |coef_t| scale / 2B (softmax terms <= 1, cnt >= 1): that bound goes to 2^15 (see header)
    w_scale = _w_scale_from_bound(coef_t * scale / float(b * W)) if prec == "fp16" else None

    mult = float(W) if (cfg["gather_with_grad"] and W > 1) else 1.0
    d_txt, d_img_all = _grad_gemms(k, dG, st.g_txt if need[0] else None, st.g_img if need[1] else None)
    d_img = _reduce_scatter_rows(d_img_all, W, group) if need[0] else None
    del dG, d_img_all

    _mark("bwd: K3 gradient + global gradient GEMMs + their reduce-scatter", dev)
    # K2: token gradients from the saved argmax, one home rank's shard at a time in rank order (the order fixes the
    # chain of fp32 additions of dTxt: same bits as one call over all images).  K4 (backward half): each dImg
    # partial is SUM-reduced to its home rank while K2 works on the next one; gather_with_grad=True reproduces the
    # reference's x world_size (every rank back-propagates the same full loss).
    rows = cfg.get("grad_rows")
    slice_out = bool(cfg.get("slice_out")) and rows is not None
    lo_t, hi_t = rows if rows is not None else (0, b)
    n_t = hi_t - lo_t
    dtxt_tok = None
    if need[3]:
        dtxt_tok = torch.empty((n_t, T, D), dtype=torch.float32, device=dev)
    out_ref, inflight = [None], []
    chunk = _dimg_chunk_images(n_t, P, D) if W > 1 else n_t
    for r in range(W):
        img_sh, arg_sh = st.shards[r], st.args[r]
        if rows is None:
            first = True
            for lo in range(0, b, chunk):
                hi = min(b, lo + chunk)
                part = k.maxsim_bwd_shard(txt_op, img_sh[lo:hi], k.argmax_images(arg_sh, lo, hi), w, r * b + lo, T, P,
                                          prec, dtxt=dtxt_tok, accumulate_txt=not (r == 0 and first), need_img=need[2],
                                          w_scale=w_scale)
                first = False
                if need[2]:
                    if W == 1:
                        out_ref[0] = part if (lo == 0 and hi == b) else _place(out_ref, part, lo, hi, (b, P, D))
                    else:
                        _send_home(part, r, lo, hi, out_ref, cfg, (b, P, D), dt_timg, mult, inflight)
                del part
        else:
            # gradient-accumulation call (train.py:166-183): only samples lo_t..hi_t-1 of every rank's batch carry
            # gradient, so K2 runs on that slice alone -- 1/accum_freq of the items
            if need[3]:
                k.maxsim_bwd_dtxt_rows_shard(img_sh, arg_sh, w, r * b, b, lo_t, n_t, T, P, prec, dtxt_tok, r > 0,
                                             w_scale=w_scale)
            if need[2]:
                for lo in range(lo_t, hi_t, chunk):
                    hi = min(hi_t, lo + chunk)
                    part = k.maxsim_bwd_shard(txt_op, img_sh[lo:hi], k.argmax_images(arg_sh, lo, hi), w, r * b + lo,
                                              T, P, prec, need_img=True, w_scale=w_scale)
                    if W == 1:
                        out_ref[0] = part if (lo == lo_t and hi == hi_t) else _place(out_ref, part, lo - lo_t, hi - lo_t,
                                                                                     (n_t, P, D))
                    else:
                        _send_home(part, r, lo - lo_t, hi - lo_t, out_ref, cfg, (n_t, P, D), dt_timg, mult, inflight)
                    del part
        del img_sh, arg_sh
        if consume and W > 1:
            st.shards[r] = None                     # this rank is done with home rank r's images: 1/W of the gathered
            st.args[r] = None                       # operand and of the argmax go back to the allocator now
    _mark("bwd: K2 over all shards (dImg reductions overlapped on the side stream)", dev)
    if inflight:
        torch.cuda.current_stream(dev).wait_stream(_comm_stream(dev))
    _mark("bwd: wait for the last dImg reduction", dev)
    dimg_tok = out_ref[0] if need[2] else None
    if consume:
        st.shards = st.args = None

    if rows is not None and not slice_out:          # zeros outside the slice (torch.cat's backward drops them anyway)
        if dtxt_tok is not None:
            full = torch.zeros((b, T, D), dtype=torch.float32, device=dev)
            full[lo_t:hi_t] = dtxt_tok
            dtxt_tok = full
        if dimg_tok is not None:
            full = torch.zeros((b, P, D), dtype=dimg_tok.dtype, device=dev)
            full[lo_t:hi_t] = dimg_tok
            dimg_tok = full
    if slice_out:
        d_img = d_img[lo_t:hi_t] if d_img is not None else None
        d_txt = d_txt[lo_t:hi_t] if d_txt is not None else None

    def fin(x, dt, scaled=False):
        if x is None:
            return None
        if mult != 1.0 and not scaled:
            x = x * mult
        return x.to(dt)

    d_ls = dscale.reshape(ls_shape).to(dt_ls) if need[4] else None
    return (fin(d_img, dt_img), fin(d_txt, dt_txt), fin(dimg_tok, dt_timg, scaled=W > 1), fin(dtxt_tok, dt_ttxt), d_ls)


def _place(out_ref, part, lo, hi, shape):
    """single-GPU chunked dImg (not used by the default plan: one chunk): collect the chunks in one fp32 buffer"""
    if out_ref[0] is None or out_ref[0].shape != tuple(shape):
        out_ref[0] = torch.empty(shape, dtype=torch.float32, device=part.device)
    out_ref[0][lo:hi] = part
    return out_ref[0]


class _SavedForward:
    """Stand-in for an autograd ctx: lets `_ColClipLossFn._forward` run outside autograd and keeps its state."""
    state = None


class _AccumSliceFn(torch.autograd.Function):
    """losses[3] of the shared forward, differentiable w.r.t. ONE sub-batch's live features and logit_scale."""

    @staticmethod
    def forward(ctx, image_features, text_features, token_image_features, token_text_features, logit_scale, state,
                lo, hi):
        ctx.state, ctx.rows = state, (lo, hi)
        ctx.dtypes = (image_features.dtype, text_features.dtype, token_image_features.dtype,
                      token_text_features.dtype, logit_scale.dtype, logit_scale.shape)
        return state.losses.clone()

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        st = ctx.state
        cfg = dict(st.fwd.cfg, grad_rows=ctx.rows, slice_out=True)
        meta = st.fwd.meta[:4] + ctx.dtypes
        return _colclip_backward(st.fwd.state, cfg, meta, g, ctx.needs_input_grad, consume=False) + (None, None, None)


class ColClipAccumulation:
    """Shared forward of one gradient-accumulation step (reference train.py:138-185).

    The reference re-runs the loss on `cat(cached[:j] + [live_j] + cached[j+1:])` for every sub-batch j: accum_freq
    full forwards (K1!) and backwards of which only sub-batch j's gradients survive.  When the model is deterministic
    between the caching pass and the live pass (no dropout / stochastic depth -- the CLIP towers of the reference have
    none), the live features equal the cached ones, so all accum_freq calls see the SAME logits: this object runs
    K1 + K3 statistics ONCE on the cached features and `slice_loss(lo, hi, **model_out)` returns the loss dict for
    sub-batch lo..hi-1 whose backward runs K3-grad + K2 on that slice only.  Over the whole step: one forward and
    one backward's worth of K2 instead of accum_freq of each.  logit_scale receives the full-batch gradient in every
    slice call, exactly as in the reference's loop.  Built by `ColClipLoss.accumulation(...)`.
    """

    def __init__(self, fwd: _SavedForward, losses: torch.Tensor, local_batch: int):
        self.fwd, self.losses, self.local_batch = fwd, losses, local_batch

    def slice_loss(self, lo, hi, image_features=None, text_features=None, token_image_features=None,
                   token_text_features=None, logit_scale=None, logit_bias=None, output_dict=False, verify=False,
                   **kwargs):
        lo, hi = int(lo), int(hi)
        if not (0 <= lo < hi <= self.local_batch):
            raise ValueError("slice must satisfy 0 <= lo < hi <= local batch of the shared forward")
        feats = (image_features, text_features, token_image_features, token_text_features)
        if any(f.shape[0] != hi - lo for f in feats):
            raise ValueError("live features must hold exactly hi - lo samples")
        if not torch.is_tensor(logit_scale):
            logit_scale = torch.tensor(float(logit_scale), device=image_features.device)
        if verify:   # debugging aid: the live token features must reproduce the cached operands bit for bit
            T, P = self.fwd.meta[0], self.fwd.meta[1]
            k, prec = self.fwd.cfg["kernels"], self.fwd.cfg["prec"]
            live = k.prepare(token_text_features, T, prec)
            if not torch.equal(live, self.fwd.state.txt_op[lo:hi]):
                raise RuntimeError("live text-token features differ from the cached ones: the model is not "
                                   "deterministic between passes; use ColClipLoss.forward(grad_rows=...) instead")
        losses = _AccumSliceFn.apply(*feats, logit_scale, self, lo, hi)
        if output_dict:
            return {"global_contrastive_loss": losses[0], "token_contrastive_loss": losses[1], "total_loss": losses[2]}
        return losses[2]


class ColClipLoss(nn.Module):
    """alpha * global CLIP InfoNCE + (1 - alpha) * token-level MaxSim InfoNCE (reference loss.py:184-296).

    Constructor and call signatures are the reference's; `forward` is one fused autograd node
    (K1 -> K3 stats | K3 grad -> K2) with text rows sharded over ranks when world_size > 1.
    """

    def __init__(
            self,
            local_loss=False,
            gather_with_grad=False,
            cache_labels=False,
            rank=0,
            world_size=1,
            use_horovod=False,
            alpha=0.5,
            **kwargs
    ):
        super().__init__()
        self.local_loss = local_loss
        self.gather_with_grad = gather_with_grad
        self.cache_labels = cache_labels
        self.rank = rank
        self.world_size = world_size
        self.use_horovod = use_horovod
        self.alpha = alpha
        # cache state (kept for API compatibility; the fused kernels use the implicit diagonal)
        self.prev_num_logits = 0
        self.labels = {}
        self._kernels = kwargs.get("kernels", None)   # test hook: stand-in backend for CPU host-logic tests
        self._group = kwargs.get("group", None)

    def _k(self):
        return self._kernels if self._kernels is not None else default_kernels()

    def get_ground_truth(self, device, num_logits) -> torch.Tensor:
        """arange labels, cached per device when cache_labels (reference loss.py:209-221)."""
        if self.prev_num_logits != num_logits or device not in self.labels:
            labels = torch.arange(num_logits, device=device, dtype=torch.long)
            if self.world_size > 1 and self.local_loss:
                labels = labels + num_logits * self.rank
            if self.cache_labels:
                self.labels[device] = labels
                self.prev_num_logits = num_logits
        else:
            labels = self.labels[device]
        return labels

    def get_logits(self, image_features, text_features, token_image_features, token_text_features, logit_scale,
                   logit_bias=None):
        """The four materialised [B,B] logit matrices (reference loss.py:223-269)."""
        if self.world_size > 1:
            if self.use_horovod:
                raise RuntimeError("horovod is not supported by colxlip_b200")
            image_features, text_features = gather_features(
                image_features, text_features, local_loss=self.local_loss, gather_with_grad=self.gather_with_grad,
                rank=self.rank, world_size=self.world_size, use_horovod=False)
            token_image_features, token_text_features = gather_features(
                token_image_features, token_text_features, local_loss=self.local_loss,
                gather_with_grad=self.gather_with_grad, rank=self.rank, world_size=self.world_size, use_horovod=False)
            if self.local_loss:
                raise NotImplementedError  # same as the reference (loss.py:246-248)
        logits_per_image = logit_scale * image_features @ text_features.T
        logits_per_text_token = logit_scale * _maxsim(token_image_features, token_text_features, MASKED, self._k())
        if logit_bias is not None:
            # the reference adds the bias in place to logits_per_image and again through its transposed
            # view (loss.py:265-267): both global matrices end up with 2 * logit_bias
            logits_per_image = logits_per_image + 2 * logit_bias
        return {"logits_per_image": logits_per_image, "logits_per_text": logits_per_image.T,
                "logits_per_image_token": logits_per_text_token.T, "logits_per_text_token": logits_per_text_token}

    def accumulation(self, image_features, text_features, token_image_features, token_text_features, logit_scale,
                     **kwargs) -> "ColClipAccumulation":
        """Shared no-grad forward over the CACHED features of all accum_freq sub-batches (this rank's, concatenated in
        order): see ColClipAccumulation.  logit_scale is read by value."""
        if self.world_size > 1:
            if self.use_horovod:
                raise RuntimeError("horovod is not supported by colxlip_b200")
            if self.local_loss:
                raise NotImplementedError
        _check_tokens(token_image_features, token_text_features)
        if token_image_features.shape[0] != token_text_features.shape[0]:
            raise ValueError("ColClipLoss needs as many texts as images (labels are the diagonal)")
        if not torch.is_tensor(logit_scale):
            logit_scale = torch.tensor(float(logit_scale), device=image_features.device)
        cfg = {
            "kernels": self._k(), "world_size": self.world_size, "rank": self.rank, "group": self._group,
            "alpha": self.alpha, "gather_with_grad": self.gather_with_grad,
            "prec": _select_precision(token_image_features, token_text_features),
        }
        fwd = _SavedForward()
        with torch.no_grad(), torch.autocast(token_image_features.device.type, enabled=False):
            losses = _ColClipLossFn._forward(fwd, image_features.detach(), text_features.detach(),
                                             token_image_features.detach(), token_text_features.detach(),
                                             logit_scale.detach(), cfg)
        return ColClipAccumulation(fwd, losses, token_text_features.shape[0])

    def forward(
            self,
            image_features=None,
            text_features=None,
            token_image_features=None,
            token_text_features=None,
            logit_scale=None,
            logit_bias=None,
            output_dict=False,
            **kwargs,
    ):
        # logit_bias is accepted and ignored exactly as in the reference (loss.py:283 does not forward it)
        if self.world_size > 1:
            if self.use_horovod:
                raise RuntimeError("horovod is not supported by colxlip_b200")
            if self.local_loss:
                raise NotImplementedError
        _check_tokens(token_image_features, token_text_features)
        if token_image_features.shape[0] != token_text_features.shape[0]:
            raise ValueError("ColClipLoss needs as many texts as images (labels are the diagonal)")
        if not torch.is_tensor(logit_scale):
            logit_scale = torch.tensor(float(logit_scale), device=image_features.device)
        cfg = {
            "kernels": self._k(), "world_size": self.world_size, "rank": self.rank, "group": self._group,
            "alpha": self.alpha, "gather_with_grad": self.gather_with_grad,
            "prec": _select_precision(token_image_features, token_text_features),
        }
        # Extension for the reference's gradient-accumulation loop (train.py:166-183): grad_rows=(lo, hi) declares
        # that only samples lo..hi-1 of this rank's batch carry gradient (the live sub-batch inside the cat of cached
        # features).  Token-feature gradients outside the slice are returned as zeros -- torch.cat's backward drops
        # them anyway -- and K2 does 1/accum_freq of the work.  The reference's forward swallows unknown kwargs, so
        # the same call is valid against it.
        rows = kwargs.get("grad_rows")
        if rows is not None:
            lo, hi = int(rows[0]), int(rows[1])
            if not (0 <= lo < hi <= token_text_features.shape[0]):
                raise ValueError("grad_rows must satisfy 0 <= lo < hi <= local batch")
            cfg["grad_rows"] = (lo, hi)
        cfg["grad_mode"] = torch.is_grad_enabled()         # autograd.Function.forward itself always runs with grad off
        losses = _ColClipLossFn.apply(image_features, text_features, token_image_features, token_text_features,
                                      logit_scale, cfg)
        global_contrastive_loss, token_contrastive_loss, total_loss = losses[0], losses[1], losses[2]
        if output_dict:
            return {"global_contrastive_loss": global_contrastive_loss,
                    "token_contrastive_loss": token_contrastive_loss, "total_loss": total_loss}
        return total_loss


# ---------------------------------------------------------------------------------------------------
# ClipLoss (reference loss.py:95-182): the global-embedding InfoNCE alone, for the non-colxlip models the
# reference's factory also serves.  Same K3 kernels; the [b,B] logit block is a library GEMM.
# ---------------------------------------------------------------------------------------------------
def _select_global_precision(*tensors: torch.Tensor) -> str:
    mode = _PRECISION
    if mode == "auto":
        mode = _auto_precision(*tensors)
    return mode


class _ClipLossFn(torch.autograd.Function):
    """total_loss of ClipLoss.forward.  Two layouts:
    symmetric (local_loss=False, or one rank): this rank's text rows against ALL images, row + column CE;
    local     (local_loss=True, W > 1): the reference's two one-directional [b,B] blocks (loss.py:144-146)."""

    @staticmethod
    def forward(ctx, image_features, text_features, logit_scale, cfg):
        with torch.autocast(image_features.device.type, enabled=False):
            return _ClipLossFn._forward(ctx, image_features, text_features, logit_scale, cfg)

    @staticmethod
    def _forward(ctx, image_features, text_features, logit_scale, cfg):
        k = cfg["kernels"]
        W, rank, group, prec = cfg["world_size"], cfg["rank"], cfg.get("group"), cfg["prec"]
        b = image_features.shape[0]
        B, row0 = b * W, rank * b
        dev = image_features.device
        local = bool(cfg["local_loss"]) and W > 1
        a_img = _round_like_gemm_input(image_features, prec)
        a_txt = _round_like_gemm_input(text_features, prec)
        all_img = _all_gather_cat(a_img, W, group)
        scale = logit_scale.detach().float().reshape(1).to(dev)
        if not local:
            G = _logits_gemm(k, a_txt, all_img)                       # rows = this rank's texts
            row_lse, col_max, col_sum, diag = k.infonce_stats(G, scale, row0)
            col_lse = col_max + torch.log(col_sum)
            part = row_lse.sum() - 2.0 * diag.sum()
            if W > 1:
                col_lse = torch.logsumexp(_all_gather_cat(col_lse.unsqueeze(0), W, group), dim=0)
                dist.all_reduce(part, group=group)
            loss = (part + col_lse.sum()) / (2.0 * B)
            ctx.save_for_backward(G, a_txt, all_img, scale, row_lse, col_lse)
        else:
            all_txt = _all_gather_cat(a_txt, W, group)
            X_i = _logits_gemm(k, a_img, all_txt)                     # logits_per_image of this rank's images
            X_t = _logits_gemm(k, a_txt, all_img)                     # logits_per_text of this rank's texts
            rl_i, _, _, dg_i = k.infonce_stats(X_i, scale, row0, rows_only=True)
            rl_t, _, _, dg_t = k.infonce_stats(X_t, scale, row0, rows_only=True)
            loss = ((rl_i - dg_i).sum() + (rl_t - dg_t).sum()) / (2.0 * b)
            ctx.save_for_backward(X_i, X_t, a_img, a_txt, all_img, all_txt, scale, rl_i, rl_t)
        ctx.cfg, ctx.local = cfg, local
        ctx.meta = (b, B, row0, image_features.dtype, text_features.dtype, logit_scale.dtype, logit_scale.shape)
        return loss

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        cfg = ctx.cfg
        k = cfg["kernels"]
        W, group = cfg["world_size"], cfg.get("group")
        gwg = bool(cfg["gather_with_grad"]) and W > 1
        b, B, row0, dt_img, dt_txt, dt_ls, ls_shape = ctx.meta
        need = ctx.needs_input_grad
        g = g.float().reshape(1).contiguous()
        if not ctx.local:
            G, a_txt, all_img, scale, row_lse, col_lse = ctx.saved_tensors
            dG, dscale = k.infonce_grad(G, scale, g, row0, row_lse, col_lse)
            if W > 1:
                dist.all_reduce(dscale, group=group)              # every rank holds the full-batch loss
            d_txt, d_img_all = _grad_gemms(k, dG, a_txt if need[0] else None, all_img if need[1] else None)
            d_img = _reduce_scatter_rows(d_img_all, W, group) if need[0] else None
            mult = float(W) if gwg else 1.0                       # autograd all_gather sums W identical losses
            if mult != 1.0:
                d_txt = d_txt * mult if d_txt is not None else None
                d_img = d_img * mult if d_img is not None else None
        else:
            X_i, X_t, a_img, a_txt, all_img, all_txt, scale, rl_i, rl_t = ctx.saved_tensors
            coef = g * (float(B) / float(b))                      # the kernel divides by 2B; the loss by 2b
            dX_i, ds_i = k.infonce_grad(X_i, scale, coef, row0, rl_i, None)
            dX_t, ds_t = k.infonce_grad(X_t, scale, coef, row0, rl_t, None)
            dscale = ds_i + ds_t                                  # this rank's own loss: no reduction
            # X_i = a_img x all_txt^T: rows grad -> d_img, gathered-operand grad -> all texts; X_t likewise
            d_img, d_alltxt = _grad_gemms(k, dX_i, a_img if (gwg and need[1]) else None, all_txt if need[0] else None)
            d_txt, d_allimg = _grad_gemms(k, dX_t, a_txt if (gwg and need[0]) else None, all_img if need[1] else None)
            if gwg:                                               # gathered operands carry grad home (SUM)
                if need[1]:
                    d_txt = d_txt + _reduce_scatter_rows(d_alltxt, W, group)
                if need[0]:
                    d_img = d_img + _reduce_scatter_rows(d_allimg, W, group)
        d_ls = dscale.reshape(ls_shape).to(dt_ls) if need[2] else None
        return (d_img.to(dt_img) if d_img is not None else None,
                d_txt.to(dt_txt) if d_txt is not None else None, d_ls, None)


class ClipLoss(nn.Module):
    """Symmetric InfoNCE over global image/text embeddings (reference loss.py:95-182), same constructor and call
    signatures; `forward` is one autograd node over the K3 kernels with logit rows sharded over ranks."""

    def __init__(
            self,
            local_loss=False,
            gather_with_grad=False,
            cache_labels=False,
            rank=0,
            world_size=1,
            use_horovod=False,
            **kwargs
    ):
        super().__init__()
        self.local_loss = local_loss
        self.gather_with_grad = gather_with_grad
        self.cache_labels = cache_labels
        self.rank = rank
        self.world_size = world_size
        self.use_horovod = use_horovod
        self.prev_num_logits = 0
        self.labels = {}
        self._kernels = kwargs.get("kernels", None)
        self._group = kwargs.get("group", None)

    _k = ColClipLoss._k
    get_ground_truth = ColClipLoss.get_ground_truth

    def get_logits(self, image_features, text_features, logit_scale, logit_bias=None):
        """(logits_per_image, logits_per_text), materialised with torch ops as in the reference (loss.py:132-160)."""
        if self.world_size > 1:
            all_image_features, all_text_features = gather_features(
                image_features, text_features, local_loss=self.local_loss, gather_with_grad=self.gather_with_grad,
                rank=self.rank, world_size=self.world_size, use_horovod=self.use_horovod)
            if self.local_loss:
                logits_per_image = logit_scale * image_features @ all_text_features.T
                logits_per_text = logit_scale * text_features @ all_image_features.T
            else:
                logits_per_image = logit_scale * all_image_features @ all_text_features.T
                logits_per_text = logits_per_image.T
        else:
            logits_per_image = logit_scale * image_features @ text_features.T
            logits_per_text = logit_scale * text_features @ image_features.T
        if logit_bias is not None:
            if self.world_size > 1 and not self.local_loss:
                # in-place add through the transposed view as well (loss.py:157-158): 2 x bias on both
                logits_per_image = logits_per_image + 2 * logit_bias
                logits_per_text = logits_per_image.T
            else:
                logits_per_image = logits_per_image + logit_bias
                logits_per_text = logits_per_text + logit_bias
        return logits_per_image, logits_per_text

    def forward(
            self,
            image_features=None,
            text_features=None,
            logit_scale=None,
            logit_bias=None,
            output_dict=False,
            logits_per_image=None,
            logits_per_text=None,
            **kwargs,
    ):
        # logit_bias / logits_per_* are accepted and ignored exactly as in the reference (loss.py:172)
        if self.world_size > 1 and self.use_horovod:
            raise RuntimeError("horovod is not supported by colxlip_b200")
        if image_features.dim() != 2 or text_features.dim() != 2 or image_features.shape != text_features.shape:
            raise ValueError("ClipLoss expects image_features and text_features of the same [B, D] shape")
        if not torch.is_tensor(logit_scale):
            logit_scale = torch.tensor(float(logit_scale), device=image_features.device)
        cfg = {"kernels": self._k(), "world_size": self.world_size, "rank": self.rank, "group": self._group,
               "gather_with_grad": self.gather_with_grad, "local_loss": self.local_loss,
               "prec": _select_global_precision(image_features, text_features)}
        total_loss = _ClipLossFn.apply(image_features, text_features, logit_scale, cfg)
        return {"total_loss": total_loss} if output_dict else total_loss
