"""Timings of the SURVEY 8(f) rows on one B200 (CUDA events, after warm-up): normalize_tokens GB/s, ClipLoss,
the gradient-accumulation slice backward vs the full one, and the cfg-4 retrieval scoring.  One JSON line each."""
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from colxlip_b200 import ClipLoss, ColClipLoss, compute_retrieval, maxsim_score_matrix, normalize_tokens


def timed(fn, warm=2, it=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(it):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def main():
    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda").manual_seed(0)
    out = []
    # f-3: one rank's image tokens at cfg-2 (b=4096 x 196 x 512), fp32 head output -> bf16 operand
    x = torch.randn(4096, 196, 512, device="cuda", generator=g)
    rows, D = 4096 * 196, 512
    ms = timed(lambda: normalize_tokens(x, out_dtype=torch.bfloat16))
    out.append({"what": "normalize_tokens fwd f32->bf16 [4096,196,512]", "ms": ms, "GBps": rows * D * 6 / ms / 1e6})
    ms_t = timed(lambda: F.normalize(x, dim=-1).to(torch.bfloat16))
    out.append({"what": "torch F.normalize + .to(bf16), same tensor", "ms": ms_t, "GBps": rows * D * 6 / ms_t / 1e6})
    xr = x.clone().requires_grad_()
    y = normalize_tokens(xr, out_dtype=torch.bfloat16)
    gy = torch.randn_like(y)
    ms = timed(lambda: torch.autograd.grad(y, xr, gy, retain_graph=True))
    out.append({"what": "normalize_tokens bwd (y bf16, g bf16 -> dx f32)", "ms": ms, "GBps": rows * D * 8 / ms / 1e6})
    del x, xr, y, gy
    # f-1: ClipLoss fwd+bwd at the cfg-2 global batch
    B = 32768
    img = F.normalize(torch.randn(B, 512, device="cuda", generator=g), dim=-1).requires_grad_()
    txt = F.normalize(torch.randn(B, 512, device="cuda", generator=g), dim=-1).requires_grad_()
    ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    crit = ClipLoss()

    def clip_step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = crit(img, txt, ls)
        loss.backward()

    ms = timed(clip_step)
    out.append({"what": "ClipLoss fwd+bwd B=32768 D=512 (bf16 autocast): cxl_gemm_nt x3 + K3", "ms": ms, "pairs_per_s": B / ms * 1e3})

    def eager_clip_step():      # the reference's ClipLoss.forward as eager PyTorch on the same GPU (loss.py:132-181)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            lpi = ls * img @ txt.T
            lpt = ls * txt @ img.T
            lab = torch.arange(B, device="cuda")
            loss = (F.cross_entropy(lpi, lab) + F.cross_entropy(lpt, lab)) / 2
        loss.backward()

    ms_e = timed(eager_clip_step)
    out.append({"what": "reference ClipLoss algorithm, eager PyTorch bf16 autocast, same GPU", "ms": ms_e, "ours_speedup": ms_e / ms})
    from colxlip_b200.kernels import default_kernels
    Kk = default_kernels()
    a16, b16 = img.detach().to(torch.bfloat16), txt.detach().to(torch.bfloat16)
    ms_g = timed(lambda: Kk.gemm_nt(a16, b16))
    out.append({"what": "cxl_gemm_nt [32768,512] x [32768,512]^T -> f32 (resident A)", "ms": ms_g, "TFLOPs": 2.0 * B * B * 512 / ms_g / 1e9})
    dG = torch.randn(B, B, device="cuda", generator=g)
    d16, d16t = Kk.cast_transpose(dG, torch.bfloat16)
    ms_c = timed(lambda: Kk.cast_transpose(dG, torch.bfloat16))
    out.append({"what": "cxl_cast_transpose [32768,32768] f32 -> bf16 + bf16^T", "ms": ms_c, "GBps": B * B * 8 / ms_c / 1e6})
    bt = b16.t().contiguous()
    ms_g2 = timed(lambda: Kk.gemm_nt(d16, bt))
    out.append({"what": "cxl_gemm_nt [32768,32768] x [512,32768]^T -> f32 (streamed A, K = 32768)", "ms": ms_g2, "TFLOPs": 2.0 * B * B * 512 / ms_g2 / 1e9})
    del img, txt, dG, d16, d16t, a16, b16, bt
    # f-4: accumulation step at B=4096 (4 sub-batches of 1024), ViT-B/16 geometry
    B, P, T, D, acc = 4096, 196, 77, 512, 4
    feats = [F.normalize(torch.randn(B, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16),
             F.normalize(torch.randn(B, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16),
             F.normalize(torch.randn(B, P, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16),
             F.normalize(torch.randn(B, T, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)]
    col = ColClipLoss(alpha=0.5)
    sub = B // acc

    def accum_step(use_rows):
        for j in range(acc):
            lo, hi = j * sub, (j + 1) * sub
            live = [t[lo:hi].clone().requires_grad_() for t in feats]
            cat = [torch.cat([t[:lo], l, t[hi:]]) for t, l in zip(feats, live)]
            kw = {"grad_rows": (lo, hi)} if use_rows else {}
            col(*cat, ls, **kw).backward()

    def accum_shared():
        state = col.accumulation(*feats, ls)
        for j in range(acc):
            lo, hi = j * sub, (j + 1) * sub
            live = [t[lo:hi].clone().requires_grad_() for t in feats]
            state.slice_loss(lo, hi, *live, ls).backward()

    ms_full = timed(lambda: accum_step(False), warm=1, it=3)
    ms_rows = timed(lambda: accum_step(True), warm=1, it=3)
    out.append({"what": f"accum_freq={acc} loop at B={B} (train.py:166-183), full backward each call", "ms": ms_full})
    out.append({"what": f"same loop with grad_rows (slice backward)", "ms": ms_rows, "speedup": ms_full / ms_rows})
    ms_sh = timed(accum_shared, warm=1, it=3)
    out.append({"what": "same step with ColClipLoss.accumulation (one shared forward + slice backwards)", "ms": ms_sh,
                "speedup": ms_full / ms_sh})
    del feats
    # f-2 / cfg-4: 1000 images x 1000 captions of 248 tokens, plain-mean scorer + recall
    ti = F.normalize(torch.randn(1000, 196, 512, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
    tt = F.normalize(torch.randn(1000, 248, 512, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
    ms = timed(lambda: maxsim_score_matrix(ti, tt))
    flops = 2.0 * 1000 * 1000 * 248 * 196 * 512
    out.append({"what": "cfg-4 MaxSim score matrix 1000x1000, T=248 P=196 D=512", "ms": ms, "TFLOPs": flops / ms / 1e9})
    M = maxsim_score_matrix(ti, tt)
    t2i = list(range(1000))
    i2t = [[i] for i in range(1000)]
    ms = timed(lambda: compute_retrieval(M, t2i, i2t))
    out.append({"what": "compute_retrieval on the 1000x1000 matrix (2 rank kernels + host report)", "ms": ms})
    for o in out:
        print(json.dumps(o))


if __name__ == "__main__":
    main()
