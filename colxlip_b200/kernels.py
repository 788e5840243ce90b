"""Torch-tensor front end of the C-ABI: allocates outputs, passes raw device pointers + stream.

`CudaKernels` is the only compute backend the product has.  The autograd functions in
`colxlip_b200.loss` take a backend object so that the *host-side* logic (sharding, collectives,
gradient scaling) can be exercised by CPU tests with a stand-in backend that lives under tests/;
nothing in this package falls back to CPU.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import (CXL_DTYPE_BF16, CXL_DTYPE_F16, CXL_DTYPE_F32, CXL_MEAN_MASKED, CXL_MEAN_PLAIN, CXL_MEAN_ROUND16,
                   check)

MASKED = CXL_MEAN_MASKED
MEAN = CXL_MEAN_PLAIN
ROUND16 = CXL_MEAN_ROUND16
# scores summed in a fixed order (cxl_maxsim_fwd_ws) unless CXL_S_ATOMIC=1 asks for the atomicAdd path of cxl_maxsim_fwd
DETERMINISTIC_SCORES = os.environ.get("CXL_S_ATOMIC", "0") in ("", "0")


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


_PREC_DT = {"bf16": CXL_DTYPE_BF16, "fp16": CXL_DTYPE_F16, "fp32": CXL_DTYPE_F32}
_PREC_TORCH = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}
_TORCH_DT = {torch.float32: CXL_DTYPE_F32, torch.bfloat16: CXL_DTYPE_BF16, torch.float16: CXL_DTYPE_F16}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(*ts: torch.Tensor) -> torch.device:
    dev = None
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("colxlip_b200 runs on CUDA (sm_100a) tensors only; there is no CPU fallback")
        dev = t.device if dev is None else dev
        if t.device != dev:
            raise RuntimeError("colxlip_b200: all tensors must live on the same CUDA device")
    return dev


class CudaKernels:
    """Thin wrappers; every method enqueues on the current CUDA stream of the tensors' device."""

    name = "cuda"

    def __init__(self):
        self.lib = _lib.load()

    def launches(self) -> int:
        """CUDA kernels launched by the library in this process so far (bench.py reports the per-step delta)."""
        return int(self.lib.cxl_kernel_launches())

    # ---- layout helpers ---------------------------------------------------------------------
    def img_row_stride(self, P: int, prec: str) -> int:
        return int(self.lib.cxl_img_row_stride(P)) if prec != "fp32" else P

    def prepare(self, x: torch.Tensor, n_pad: int, prec: str) -> torch.Tensor:
        """[N,n,D] features -> kernel operand: bf16 / fp16 [N,n_pad,D] (tensor-core path) or fp32 [N,n,D]."""
        dev = _require_cuda(x)
        N, n, D = x.shape
        x = x.detach()
        if prec == "fp32":
            return x.float().contiguous()
        op_dt = _PREC_TORCH[prec]
        if x.dtype == op_dt and x.is_contiguous() and n_pad == n:
            return x
        if x.dtype not in _TORCH_DT:
            x = x.float()
        if x.stride(2) != 1:
            x = x.contiguous()
        out = torch.empty((N, n_pad, D), dtype=op_dt, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cxl_pack_operand(_ptr(x), _TORCH_DT[x.dtype], x.stride(0), x.stride(1), N, n, n_pad, D,
                                            _ptr(out), _PREC_DT[prec], _stream()), "cxl_pack_operand")
        return out

    def _alloc_argmax(self, Bt: int, Bi: int, T: int, P: int, dev) -> torch.Tensor:
        """Image-major saved argmax [Bi, cxl_argmax_ld(Bt,T)] (u8 for P <= 256, else 16-bit)."""
        ld = int(self.lib.cxl_argmax_ld(Bt, T))
        return torch.empty((Bi, ld), dtype=torch.uint8 if P <= 256 else torch.int16, device=dev)

    @staticmethod
    def argmax_images(argmax: torch.Tensor, lo: int, hi: int) -> torch.Tensor:
        """images lo..hi-1 of a saved argmax (image-major: a row range)"""
        return argmax if (lo == 0 and hi == argmax.shape[0]) else argmax[lo:hi]

    @staticmethod
    def argmax_as_text_major(argmax: torch.Tensor, Bt: int, T: int) -> torch.Tensor:
        """[Bi, ld] image-major -> [Bt, Bi, T] int64 (for checks against the oracle)."""
        Bi = argmax.shape[0]
        a = argmax[:, :Bt * T].reshape(Bi, Bt, T).permute(1, 0, 2)
        return a.long() & 0xFFFF

    # ---- K1 --------------------------------------------------------------------------------------
    def maxsim_fwd(self, txt_op: torch.Tensor, img_op: torch.Tensor, T: int, P: int, prec: str, mean_mode: int,
                   save_argmax: bool, deterministic: Optional[bool] = None
                   ) -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor]]:
        """deterministic (default: on unless CXL_S_ATOMIC=1): fixed-order sums of a text's row maxima through
        cxl_maxsim_fwd_ws (a transient workspace of ((T+30)//32) * Bt * Bi floats) instead of atomicAdd."""
        dev = _require_cuda(txt_op, img_op)
        Bt, Bi, D = txt_op.shape[0], img_op.shape[0], txt_op.shape[2]
        S = torch.empty((Bt, Bi), dtype=torch.float32, device=dev)
        cnt = torch.empty((Bt, Bi), dtype=torch.float32, device=dev) if (mean_mode & ~ROUND16) == MASKED else None
        argmax = self._alloc_argmax(Bt, Bi, T, P, dev) if save_argmax else None
        if deterministic is None:
            deterministic = DETERMINISTIC_SCORES
        with torch.cuda.device(dev):
            if deterministic:
                nws = int(self.lib.cxl_maxsim_fwd_workspace_bytes(Bt, Bi, T, _PREC_DT[prec]))
                ws = torch.empty(nws, dtype=torch.uint8, device=dev)
                check(self.lib.cxl_maxsim_fwd_ws(_ptr(txt_op), _ptr(img_op), Bt, Bi, T, P, D, _PREC_DT[prec], mean_mode,
                                                 _ptr(S), _ptr(cnt), _ptr(argmax), _ptr(ws), nws, _stream()),
                      "cxl_maxsim_fwd_ws")
            else:
                check(self.lib.cxl_maxsim_fwd(_ptr(txt_op), _ptr(img_op), Bt, Bi, T, P, D, _PREC_DT[prec], mean_mode,
                                              _ptr(S), _ptr(cnt), _ptr(argmax), _stream()), "cxl_maxsim_fwd")
        return S, cnt, argmax

    # ---- K2 --------------------------------------------------------------------------------------
    def pair_weights(self, dS: torch.Tensor, cnt: Optional[torch.Tensor], T: int, mean_mode: int) -> torch.Tensor:
        """w = dS / (cnt + 1e-8) (masked mean) or dS / T (plain mean), fp32 [Bt,Bi]."""
        dev = _require_cuda(dS, cnt)
        dS = dS.float().contiguous()
        w = torch.empty_like(dS)
        with torch.cuda.device(dev):
            check(self.lib.cxl_pair_weights(_ptr(dS), _ptr(cnt), None, mean_mode, T, dS.numel(), _ptr(w), _stream()),
                  "cxl_pair_weights")
        return w

    def maxsim_bwd(self, txt_op, img_op, argmax, w, T: int, P: int, prec: str,
                   need_txt: bool = True, need_img: bool = True, k_off: int = 0, n_img: Optional[int] = None,
                   generic: bool = False, library_scratch: bool = False):
        """w: [Bt, ld_pairs] fp32 pair weights (pair_weights / infonce_grad with cnt)."""
        dev = _require_cuda(txt_op, img_op, argmax, w)
        Bt, D = txt_op.shape[0], txt_op.shape[2]
        ld = w.shape[1]
        n_img = img_op.shape[0] - k_off if n_img is None else n_img
        assert w.is_contiguous() and w.dtype == torch.float32
        dtxt = torch.empty((Bt, T, D), dtype=torch.float32, device=dev) if need_txt else None
        dimg = torch.empty((n_img, P, D), dtype=torch.float32, device=dev) if need_img else None
        img_shard = img_op[k_off:k_off + n_img]
        dt = _PREC_DT[prec]
        with torch.cuda.device(dev):
            if generic:
                check(self.lib.cxl_maxsim_bwd_generic(_ptr(txt_op), _ptr(img_shard), _ptr(argmax), _ptr(w), Bt, n_img,
                                                      T, P, D, dt, ld, k_off, _ptr(dtxt), _ptr(dimg), 0, _stream()),
                      "cxl_maxsim_bwd_generic")
            elif library_scratch:
                # plain entry point: the library takes the dImg scratch from the device's default CUDA memory pool
                check(self.lib.cxl_maxsim_bwd(_ptr(txt_op), _ptr(img_shard), _ptr(argmax), _ptr(w), Bt, n_img, T, P, D,
                                              dt, ld, k_off, _ptr(dtxt), _ptr(dimg), 0, _stream()), "cxl_maxsim_bwd")
            else:
                # scratch for the dImg row lists comes from torch's caching allocator (no second memory pool)
                nws = int(self.lib.cxl_maxsim_bwd_workspace_bytes(Bt, n_img, T, P, D, dt)) if need_img else 0
                ws = torch.empty(nws, dtype=torch.uint8, device=dev) if nws > 0 else None
                check(self.lib.cxl_maxsim_bwd_ws(_ptr(txt_op), _ptr(img_shard), _ptr(argmax), _ptr(w), Bt, n_img, T, P,
                                                 D, dt, ld, k_off, _ptr(dtxt), _ptr(dimg), 0, _ptr(ws), nws, _stream()),
                      "cxl_maxsim_bwd_ws")
        return dtxt, dimg

    def maxsim_bwd_shard(self, txt_op, img_sh, arg_sh, w, k_off: int, T: int, P: int, prec: str,
                         dtxt: Optional[torch.Tensor] = None, accumulate_txt: bool = False, need_img: bool = True,
                         w_scale: Optional[torch.Tensor] = None):
        """K2 for ONE SHARD of the images with shard-local operands (cxl_maxsim_bwd_shard): img_sh [n,p_stride,D] and
        arg_sh [n, ld] are the shard's own buffers, w [Bt, ld_pairs] covers all images and the shard is its columns
        [k_off, k_off+n).  `dtxt` [Bt,T,D] fp32 (optional) is overwritten, or added to when accumulate_txt.
        w_scale: optional device pair [s, 1/s] (power of two) applied to the 16-bit copy of w inside the kernels.
        Returns the shard's dImg partial [n,P,D] fp32 (None unless need_img)."""
        dev = _require_cuda(txt_op, img_sh, arg_sh, w, dtxt)
        Bt, D = txt_op.shape[0], txt_op.shape[2]
        n = img_sh.shape[0]
        assert w.is_contiguous() and w.dtype == torch.float32 and img_sh.is_contiguous() and arg_sh.is_contiguous()
        assert dtxt is None or (dtxt.is_contiguous() and dtxt.dtype == torch.float32)
        dimg = torch.empty((n, P, D), dtype=torch.float32, device=dev) if need_img else None
        dt = _PREC_DT[prec]
        with torch.cuda.device(dev):
            nws = int(self.lib.cxl_maxsim_bwd_workspace_bytes(Bt, n, T, P, D, dt)) if need_img else 0
            ws = torch.empty(nws, dtype=torch.uint8, device=dev) if nws > 0 else None
            check(self.lib.cxl_maxsim_bwd_shard(_ptr(txt_op), _ptr(img_sh), _ptr(arg_sh), _ptr(w), Bt, n, T, P, D, dt,
                                                w.shape[1], k_off, _ptr(dtxt), 1 if accumulate_txt else 0, _ptr(dimg),
                                                0, _ptr(w_scale), _ptr(ws), nws, _stream()), "cxl_maxsim_bwd_shard")
        return dimg

    def maxsim_bwd_dtxt_rows_shard(self, img_sh, arg_sh, w, k_off: int, Bt_total: int, m0: int, m_cnt: int, T: int,
                                   P: int, prec: str, dtxt: torch.Tensor, accumulate_txt: bool,
                                   w_scale: Optional[torch.Tensor] = None):
        """dTxt of texts m0..m0+m_cnt-1 against one shard of the images (shard-local img / argmax), into dtxt
        [m_cnt,T,D] fp32 (overwritten or accumulated)."""
        dev = _require_cuda(img_sh, arg_sh, w, dtxt)
        D = img_sh.shape[2]
        assert w.is_contiguous() and w.dtype == torch.float32 and w.shape[0] == Bt_total
        assert dtxt.is_contiguous() and dtxt.dtype == torch.float32
        dt = _PREC_DT[prec]
        with torch.cuda.device(dev):
            check(self.lib.cxl_maxsim_bwd_dtxt_rows_shard(_ptr(img_sh), _ptr(arg_sh), _ptr(w), Bt_total, m0, m_cnt,
                                                          img_sh.shape[0], T, P, D, dt, w.shape[1], k_off, _ptr(dtxt),
                                                          1 if accumulate_txt else 0, _ptr(w_scale), _stream()),
                  "cxl_maxsim_bwd_dtxt_rows_shard")
        return dtxt

    def maxsim_bwd_dtxt_rows(self, img_op, argmax, w, Bt_total: int, m0: int, m_cnt: int, T: int, P: int, prec: str,
                             generic: bool = False):
        """dTxt [m_cnt,T,D] of texts m0..m0+m_cnt-1 only (all images); argmax / w are the forward's full buffers.
        The sub-batch backward of the gradient-accumulation loop (reference train.py:166-183)."""
        dev = _require_cuda(img_op, argmax, w)
        D = img_op.shape[2]
        assert w.is_contiguous() and w.dtype == torch.float32 and w.shape[0] == Bt_total
        dtxt = torch.empty((m_cnt, T, D), dtype=torch.float32, device=dev)
        dt = _PREC_DT[prec]
        with torch.cuda.device(dev):
            check(self.lib.cxl_maxsim_bwd_dtxt_rows(_ptr(img_op), _ptr(argmax), _ptr(w), Bt_total, m0, m_cnt,
                                                    img_op.shape[0], T, P, D, dt, w.shape[1], 0, _ptr(dtxt),
                                                    1 if generic else 0, _stream()), "cxl_maxsim_bwd_dtxt_rows")
        return dtxt

    # ---- K3 --------------------------------------------------------------------------------------
    def infonce_stats(self, X: torch.Tensor, scale: torch.Tensor, row0: int, rows_only: bool = False):
        """rows_only: just (row_lse, None, None, diag) -- the one-directional CE of ClipLoss(local_loss=True)."""
        dev = _require_cuda(X, scale)
        B_loc, B = X.shape
        row_lse = torch.empty(B_loc, dtype=torch.float32, device=dev)
        diag = torch.empty(B_loc, dtype=torch.float32, device=dev)
        if rows_only:
            with torch.cuda.device(dev):
                check(self.lib.cxl_infonce_stats(_ptr(X), _ptr(scale), B_loc, B, row0, _ptr(row_lse), None, None,
                                                 _ptr(diag), None, 0, _stream()), "cxl_infonce_stats")
            return row_lse, None, None, diag
        col_max = torch.empty(B, dtype=torch.float32, device=dev)
        col_sum = torch.empty(B, dtype=torch.float32, device=dev)
        nbytes = int(self.lib.cxl_infonce_workspace_bytes(B_loc, B))
        ws = torch.empty(max(nbytes, 4), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cxl_infonce_stats(_ptr(X), _ptr(scale), B_loc, B, row0, _ptr(row_lse), _ptr(col_max),
                                             _ptr(col_sum), _ptr(diag), _ptr(ws), nbytes, _stream()),
                  "cxl_infonce_stats")
        return row_lse, col_max, col_sum, diag

    def infonce_grad(self, X, scale, coef, row0: int, row_lse, col_lse, cnt=None, out=None):
        """dX (or, with cnt, the MaxSim pair weights dX / (cnt + 1e-8)); `out` may alias X.
        col_lse=None: rows-only gradient coef/(2B) * scale * (softmax_row - delta)."""
        dev = _require_cuda(X, scale, coef, row_lse, col_lse)
        B_loc, B = X.shape
        dX = torch.empty_like(X) if out is None else out
        dscale = torch.zeros(1, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            if DETERMINISTIC_SCORES:                      # d logit_scale summed in a fixed order (per-block partials in ws)
                nws = int(self.lib.cxl_infonce_grad_workspace_bytes(B_loc, B))
                ws = torch.empty(max(nws, 4), dtype=torch.uint8, device=dev)
                check(self.lib.cxl_infonce_grad_ws(_ptr(X), _ptr(scale), _ptr(coef), B_loc, B, row0, _ptr(row_lse),
                                                   _ptr(col_lse), _ptr(cnt), _ptr(dX), _ptr(dscale), _ptr(ws), nws,
                                                   _stream()), "cxl_infonce_grad_ws")
            else:
                check(self.lib.cxl_infonce_grad(_ptr(X), _ptr(scale), _ptr(coef), B_loc, B, row0, _ptr(row_lse),
                                                _ptr(col_lse), _ptr(cnt), _ptr(dX), _ptr(dscale), _stream()),
                      "cxl_infonce_grad")
        return dX, dscale

    # ---- producer side: L2 normalise + operand cast -----------------------------------
    @staticmethod
    def _dt(t: torch.Tensor) -> int:
        if t.dtype in _TORCH_DT:
            return _TORCH_DT[t.dtype]
        raise RuntimeError(f"colxlip_b200: unsupported dtype {t.dtype} (float32 / bfloat16 / float16 only)")

    def l2norm_fwd(self, x: torch.Tensor, out_dtype: torch.dtype, save_inv: bool = True):
        """x [..., D] (f32|bf16, contiguous) -> (y [..., D] out_dtype, inv_norm [rows] f32 or None)."""
        dev = _require_cuda(x)
        D = x.shape[-1]
        rows = x.numel() // D
        y = torch.empty(x.shape, dtype=out_dtype, device=dev)
        inv = torch.empty(rows, dtype=torch.float32, device=dev) if save_inv else None
        with torch.cuda.device(dev):
            check(self.lib.cxl_l2norm_fwd(_ptr(x), self._dt(x), rows, D, _ptr(y), self._dt(y), _ptr(inv), _stream()),
                  "cxl_l2norm_fwd")
        return y, inv

    def l2norm_bwd(self, y: torch.Tensor, g: torch.Tensor, inv: torch.Tensor, out_dtype: torch.dtype):
        dev = _require_cuda(y, g, inv)
        D = y.shape[-1]
        rows = y.numel() // D
        dx = torch.empty(y.shape, dtype=out_dtype, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cxl_l2norm_bwd(_ptr(y), self._dt(y), _ptr(g), self._dt(g), _ptr(inv), rows, D, _ptr(dx),
                                          self._dt(dx), _stream()), "cxl_l2norm_bwd")
        return dx

    # ---- global-embedding GEMMs --------------------------------------------------------
    def gemm_nt(self, a: torch.Tensor, b: torch.Tensor) -> Optional[torch.Tensor]:
        """C [M,N] f32 = a [M,K] @ b [N,K]^T on the tcgen05 pipeline; a, b contiguous bf16 or fp16 of one dtype.
        None when the shape is outside the kernel (K % 64 != 0): the caller falls back to its library GEMM."""
        dev = _require_cuda(a, b)
        M, K = a.shape
        N = b.shape[0]
        if K % 64 != 0 or a.dtype != b.dtype or a.dtype not in (torch.bfloat16, torch.float16):
            return None
        a, b = a.contiguous(), b.contiguous()
        c = torch.empty((M, N), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = self.lib.cxl_gemm_nt(_ptr(a), _ptr(b), M, N, K, _TORCH_DT[a.dtype], _ptr(c), N, _stream())
        if rc == -2:
            return None
        check(rc, "cxl_gemm_nt")
        return c

    def cast_transpose(self, x: torch.Tensor, dtype: torch.dtype, want_plain: bool = True, want_t: bool = True,
                       scale: Optional[torch.Tensor] = None):
        """x [R,C] f32 -> (x as `dtype` [R,C] or None, x^T as `dtype` [C,R] or None) in one pass; `scale` (device
        scalar, optional) is multiplied in before the cast."""
        dev = _require_cuda(x)
        x = x.contiguous()
        R, C = x.shape
        y = torch.empty((R, C), dtype=dtype, device=dev) if want_plain else None
        yt = torch.empty((C, R), dtype=dtype, device=dev) if want_t else None
        with torch.cuda.device(dev):
            check(self.lib.cxl_cast_transpose(_ptr(x), R, C, _ptr(y), _ptr(yt), _TORCH_DT[dtype], _ptr(scale), _stream()),
                  "cxl_cast_transpose")
        return y, yt

    # ---- token heads -----------------------------------------------------------------------
    def token_head(self, x: torch.Tensor, ln1_w, ln1_b, lin_w, lin_b, ln2_w, ln2_b, out_dtype: torch.dtype,
                   op_dtype: torch.dtype = torch.bfloat16, valid_len: Optional[torch.Tensor] = None,
                   normalize: bool = True, eps1: float = 1e-5, eps2: float = 1e-5) -> Optional[torch.Tensor]:
        """[N, T, width] -> [N, T, D]: LayerNorm -> Linear -> GELU -> LayerNorm (-> L2 normalise) in three launches
        (cxl_head_ln_in, cxl_gemm_nt, cxl_head_out).  valid_len [N] int32: tokens at positions >= valid_len[n] enter the
        head as zeros (EOT masking).  None when the widths are outside the kernels (caller keeps its torch modules)."""
        dev = _require_cuda(x, lin_w)
        N, T, width = x.shape
        D = lin_w.shape[0]
        if width % 256 or width > 1024 or D % 256 or D > 1024 or x.dtype not in _TORCH_DT:
            return None
        if x.dtype != torch.float32 and x.dtype != op_dtype:
            x = x.float()
        x = x.contiguous()
        rows = N * T
        f32 = lambda t: t.detach().float().contiguous()                                                   # noqa: E731
        g1, b1, g2, b2, lb = f32(ln1_w), f32(ln1_b), f32(ln2_w), f32(ln2_b), f32(lin_b)
        w16 = lin_w.detach().to(op_dtype).contiguous()
        a16 = torch.empty((rows, width), dtype=op_dtype, device=dev)
        vl = None if valid_len is None else valid_len.to(device=dev, dtype=torch.int32).contiguous()
        with torch.cuda.device(dev):
            check(self.lib.cxl_head_ln_in(_ptr(x), _TORCH_DT[x.dtype], rows, width, _ptr(g1), _ptr(b1), eps1, _ptr(vl), T,
                                          _ptr(a16), _TORCH_DT[op_dtype], _stream()), "cxl_head_ln_in")
        h = self.gemm_nt(a16, w16)
        if h is None:
            return None
        y = torch.empty((N, T, D), dtype=out_dtype, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cxl_head_out(_ptr(h), rows, D, _ptr(lb), _ptr(g2), _ptr(b2), eps2, 1 if normalize else 0,
                                        _ptr(y), _TORCH_DT[out_dtype], _stream()), "cxl_head_out")
        return y

    # ---- retrieval ---------------------------------------------------------------------------------
    def rank_of_target(self, scores: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """scores [R,C] fp32 (any strides), target [R] int32 -> rank [R] int32 (0 = best; -1 where the target is
        outside [0, C): the kernel never dereferences such an index, callers turn it into IndexError)."""
        dev = _require_cuda(scores, target)
        if scores.dtype != torch.float32:
            scores = scores.float()
        target = target.to(device=dev, dtype=torch.int32).contiguous()
        R, C = scores.shape
        out = torch.empty(R, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cxl_rank_of_target(_ptr(scores), scores.stride(0), scores.stride(1), R, C, _ptr(target),
                                              _ptr(out), _stream()), "cxl_rank_of_target")
        return out

    # ---- debug / test hooks -------------------------------------------------------------------------
    def debug_tile(self, txt_bf16: torch.Tensor, img_bf16: torch.Tensor) -> torch.Tensor:
        """Raw accumulators of one 128 x n_tile tensor-core tile: txt [<=128, D], img [129..256, D]."""
        dev = _require_cuda(txt_bf16, img_bf16)
        ra, D = txt_bf16.shape
        rb = img_bf16.shape[0]
        n_tile = (rb + 15) // 16 * 16
        out = torch.zeros((128, n_tile), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(self.lib.cxl_debug_tile(_ptr(txt_bf16.contiguous()), _ptr(img_bf16.contiguous()), ra, rb, D, n_tile,
                                          _ptr(out), _stream()), "cxl_debug_tile")
        return out

    def maxsim_fwd_tc_ex(self, txt_op, img_op, T, P, mean_mode, save_argmax, force_stream_a=0, force_grid=0):
        dev = _require_cuda(txt_op, img_op)
        Bt, Bi, D = txt_op.shape[0], img_op.shape[0], txt_op.shape[2]
        S = torch.empty((Bt, Bi), dtype=torch.float32, device=dev)
        cnt = torch.empty((Bt, Bi), dtype=torch.float32, device=dev)
        argmax = self._alloc_argmax(Bt, Bi, T, P, dev) if save_argmax else None
        with torch.cuda.device(dev):
            check(self.lib.cxl_maxsim_fwd_tc_ex(_ptr(txt_op), _ptr(img_op), Bt, Bi, T, P, D, mean_mode, _ptr(S),
                                                _ptr(cnt), _ptr(argmax), force_stream_a, force_grid, _stream()),
                  "cxl_maxsim_fwd_tc_ex")
        return S, cnt, argmax


_default: Optional[CudaKernels] = None


def default_kernels() -> CudaKernels:
    global _default
    if _default is None:
        _default = CudaKernels()
    return _default
