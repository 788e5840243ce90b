"""CPU stand-in for colxlip_b200.kernels.CudaKernels -- TEST INFRASTRUCTURE ONLY.

Implements the same backend interface with plain torch (via the oracle) so the host-side logic of
colxlip_b200.loss (autograd wiring, row sharding, collectives, gradient scaling) can be tested on CPU
with gloo.  The product never imports this module.
"""
import torch

from oracle import colxlip_oracle as O

MASKED, MEAN, ROUND16 = 0, 1, 0x10


class TorchRefKernels:
    name = "torch-ref"

    def img_row_stride(self, P, prec):
        return P

    def prepare(self, x, n_pad, prec):
        x = x.detach()
        if prec == "bf16":
            return x.to(torch.bfloat16).float().contiguous()
        if prec == "fp16":
            return x.to(torch.float16).float().contiguous()
        return x.float().contiguous()

    def maxsim_fwd(self, txt_op, img_op, T, P, prec, mean_mode, save_argmax, deterministic=None):
        mx, arg = O.maxsim_rowmax(img_op, txt_op)
        if mean_mode & ROUND16:                                  # strict autocast arithmetic: 16-bit row maxima
            mx = mx.to(torch.float16 if prec == "fp16" else torch.bfloat16).float()
            mean_mode &= ~ROUND16
        if mean_mode == MASKED:
            cnt = (mx != 0).float().sum(dim=2)
            S = mx.sum(dim=2) / (cnt + 1e-8)
        else:
            cnt = None
            S = mx.mean(dim=2)
        return S, cnt, (arg if save_argmax else None)

    @staticmethod
    def argmax_images(argmax, lo, hi):
        return argmax[:, lo:hi]                                  # the stand-in keeps the argmax text-major [Bt,Bi,T]

    def pair_weights(self, dS, cnt, T, mean_mode):
        return dS.float() / ((cnt + 1e-8) if mean_mode == MASKED else float(T))

    def maxsim_bwd(self, txt_op, img_op, argmax, w, T, P, prec, need_txt=True, need_img=True, k_off=0, n_img=None,
                   generic=False):
        Bt, _, D = txt_op.shape
        n_img = img_op.shape[0] - k_off if n_img is None else n_img
        sl = slice(k_off, k_off + n_img)
        w = w[:, sl]                                                                      # [Bt,n]
        arg = argmax[:, sl]                                                               # [Bt,n,T]
        img = img_op[sl]
        dtxt = dimg = None
        if need_txt:
            dtxt = torch.zeros(Bt, T, D)
            for k in range(n_img):
                dtxt += w[:, k, None, None] * img[k][arg[:, k]]                           # [Bt,T,D]
        if need_img:
            dimg = torch.zeros(n_img, P, D)
            for k in range(n_img):
                contrib = (w[:, k, None, None] * txt_op).reshape(Bt * T, D)
                dimg[k].index_add_(0, arg[:, k].reshape(-1), contrib)
        return dtxt, dimg

    def maxsim_bwd_shard(self, txt_op, img_sh, arg_sh, w, k_off, T, P, prec, dtxt=None, accumulate_txt=False,
                         need_img=True, w_scale=None):
        """arg_sh: the shard's argmax [Bt, n, T] (what maxsim_fwd(txt_op, img_sh) returned)."""
        Bt, _, D = txt_op.shape
        n = img_sh.shape[0]
        ws = w[:, k_off:k_off + n]
        if dtxt is not None:
            if not accumulate_txt:
                dtxt.zero_()
            for k in range(n):
                dtxt += ws[:, k, None, None] * img_sh[k][arg_sh[:, k]]
        dimg = None
        if need_img:
            dimg = torch.zeros(n, P, D)
            for k in range(n):
                contrib = (ws[:, k, None, None] * txt_op).reshape(Bt * T, D)
                dimg[k].index_add_(0, arg_sh[:, k].reshape(-1), contrib)
        return dimg

    def maxsim_bwd_dtxt_rows_shard(self, img_sh, arg_sh, w, k_off, Bt_total, m0, m_cnt, T, P, prec, dtxt, accumulate_txt,
                                   w_scale=None):
        if not accumulate_txt:
            dtxt.zero_()
        for k in range(img_sh.shape[0]):
            dtxt += w[m0:m0 + m_cnt, k_off + k, None, None] * img_sh[k][arg_sh[m0:m0 + m_cnt, k]]
        return dtxt

    def maxsim_bwd_dtxt_rows(self, img_op, argmax, w, Bt_total, m0, m_cnt, T, P, prec, generic=False):
        D = img_op.shape[2]
        dtxt = torch.zeros(m_cnt, T, D)
        for k in range(img_op.shape[0]):
            dtxt += w[m0:m0 + m_cnt, k, None, None] * img_op[k][argmax[m0:m0 + m_cnt, k]]
        return dtxt

    def infonce_stats(self, X, scale, row0, rows_only=False):
        Z = X * scale
        row_lse = torch.logsumexp(Z, dim=1)
        if rows_only:
            idx = torch.arange(X.shape[0])
            return row_lse, None, None, Z[idx, row0 + idx]
        col_max = Z.max(dim=0).values
        col_sum = torch.exp(Z - col_max).sum(dim=0)
        idx = torch.arange(X.shape[0])
        diag = Z[idx, row0 + idx]
        return row_lse, col_max, col_sum, diag

    def infonce_grad(self, X, scale, coef, row0, row_lse, col_lse, cnt=None, out=None):
        B_loc, B = X.shape
        Z = X * scale
        t = torch.exp(Z - row_lse[:, None])
        if col_lse is not None:
            t = t + torch.exp(Z - col_lse[None, :])
        idx = torch.arange(B_loc)
        t[idx, row0 + idx] -= 2.0 if col_lse is not None else 1.0
        t = t * (coef / (2.0 * B))
        dX = t * scale
        if cnt is not None:
            dX = dX / (cnt + 1e-8)
        return dX, (t * X).sum().reshape(1)

    def gemm_nt(self, a, b):
        return a.float() @ b.float().T

    def cast_transpose(self, x, dtype, want_plain=True, want_t=True, scale=None):
        y = (x if scale is None else x * scale).to(dtype)
        return (y if want_plain else None), (y.t().contiguous() if want_t else None)

    def l2norm_fwd(self, x, out_dtype, save_inv=True):
        xf = x.float()
        inv = 1.0 / xf.norm(dim=-1).clamp_min(1e-12)
        return (xf * inv[..., None]).to(out_dtype), (inv.reshape(-1) if save_inv else None)

    def l2norm_bwd(self, y, g, inv, out_dtype):
        yf, gf = y.float(), g.float()
        dot = (yf * gf).sum(dim=-1, keepdim=True)
        return ((gf - yf * dot) * inv.reshape(y.shape[:-1])[..., None]).to(out_dtype)

    def rank_of_target(self, scores, target):
        scores = scores.float()
        R, C = scores.shape
        tv = scores[torch.arange(R), target.long()]
        cols = torch.arange(C)[None, :]
        gt = scores > tv[:, None]
        eq = (scores == tv[:, None]) & (cols < target.long()[:, None])
        return (gt | eq).sum(dim=1).to(torch.int32)
