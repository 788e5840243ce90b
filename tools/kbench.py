"""Per-kernel timings (CUDA events) of K1/K2/K3 at a given shape, for development under gpurun."""
import argparse
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from colxlip_b200.kernels import default_kernels

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=2048)
ap.add_argument("--P", type=int, default=196)
ap.add_argument("--T", type=int, default=77)
ap.add_argument("--D", type=int, default=512)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--only", default="")
a = ap.parse_args()
K = default_kernels()
B, P, T, D = a.B, a.P, a.T, a.D
g = torch.Generator(device="cuda").manual_seed(1)
nrm = torch.nn.functional.normalize
ti = nrm(torch.randn(B, P, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
tt = nrm(torch.randn(B, T, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
txt_op = K.prepare(tt, T, "bf16")
img_op = K.prepare(ti, K.img_row_stride(P, "bf16"), "bf16")


def timeit(name, fn, flops=None, items=None):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = min(ts)
    extra = ""
    if flops:
        extra += f"  {flops / ms / 1e9:.1f} TFLOP/s"
    if items:
        extra += f"  {ms * 1e-3 * 148 * 1.9e9 / items:.2f} clk/item/SM(@1.9GHz)"
    print(f"{name:28s} {ms:10.3f} ms{extra}", flush=True)
    return out


S, cnt, arg = timeit("K1 maxsim_fwd (bf16 tc)", lambda: K.maxsim_fwd(txt_op, img_op, T, P, "bf16", 0, True),
                     flops=2.0 * B * B * T * P * D)
scale = torch.tensor([1 / 0.07], device="cuda")
coef = torch.tensor([0.5], device="cuda")
rl, cm, cs, dg = timeit("K3 infonce_stats", lambda: K.infonce_stats(S, scale, 0))
cl = cm + torch.log(cs)
w, _ = timeit("K3 infonce_grad(+cnt)", lambda: K.infonce_grad(S, scale, coef, 0, rl, cl, cnt=cnt))
items = float(B) * B * T
if a.only in ("", "dtxt"):
    timeit("K2 dTxt", lambda: K.maxsim_bwd(txt_op, img_op, arg, w, T, P, "bf16", need_img=False), items=items)
if a.only in ("", "dimg"):
    timeit("K2 dImg", lambda: K.maxsim_bwd(txt_op, img_op, arg, w, T, P, "bf16", need_txt=False), items=items)
