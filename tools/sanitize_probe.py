"""Small launches of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck), VERDICT r1 #9:
the cfg-1 shape (B=64, P=49: single-CTA K1, generic-width paths) and the (300, 11, 77, 196, 512) backward case
(CTA-pair K1 with cross-CTA commits and remote mbarrier arrives, K2 TMA rings, dImg v3 and v4), plus K3, the
allocation-free K3 entry, the plain GEMM epilogue, cast/transpose, L2 normalise and the token head.

    compute-sanitizer --tool memcheck python tools/sanitize_probe.py
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from colxlip_b200 import _lib
from colxlip_b200.kernels import default_kernels

K = default_kernels()
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(3)
nrm = torch.nn.functional.normalize


def case(Bt, Bi, T, P, D, prec="bf16"):
    dt = torch.bfloat16 if prec == "bf16" else torch.float16
    ti = nrm(torch.randn(Bi, P, D, device=dev, generator=g), dim=-1).to(dt)
    tt = nrm(torch.randn(Bt, T, D, device=dev, generator=g), dim=-1).to(dt)
    tt[:, T - 5:] = 0                                               # zero-padded text tails
    txt_op = K.prepare(tt, T, prec)
    img_op = K.prepare(ti, K.img_row_stride(P, prec), prec)
    S, cnt, arg = K.maxsim_fwd(txt_op, img_op, T, P, prec, 0, True)
    dS = torch.randn(Bt, Bi, device=dev, generator=g) * 1e-3
    w = K.pair_weights(dS, cnt, T, 0)
    dtxt = torch.empty(Bt, T, D, device=dev)
    dimg = K.maxsim_bwd_shard(txt_op, img_op, arg, w, 0, T, P, prec, dtxt=dtxt)
    torch.cuda.synchronize()
    assert torch.isfinite(S).all() and torch.isfinite(dtxt).all() and torch.isfinite(dimg).all()
    return S, cnt


S, cnt = case(64, 64, 77, 49, 512)                                  # cfg-1
case(300, 11, 77, 196, 512)                                         # pair kernel + streaming K2 (dImg v3)
os.environ["CXL_DIMG_V"] = "4"
case(300, 11, 77, 196, 512)                                         # dImg v4 (env is read per call)
os.environ.pop("CXL_DIMG_V")
case(40, 9, 77, 576, 768)                                           # cfg-5 geometry: split images, u16 argmax, v4
case(70, 9, 77, 196, 256, "fp16")
# K3 (both entries) on the cfg-1 scores
scale = torch.tensor([1 / 0.07], device=dev)
coef = torch.tensor([0.5], device=dev)
rl, cm, cs, dg = K.infonce_stats(S, scale, 0)
lib = _lib.load()
p = lambda t: ctypes.c_void_p(t.data_ptr())                         # noqa: E731
o = [torch.empty(64, device=dev) for _ in range(4)]
_lib.check(lib.cxl_infonce_stats_static(p(S), p(scale), 64, 64, 0, p(o[0]), p(o[1]), p(o[2]), p(o[3]),
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "static")
K.infonce_grad(S, scale, coef, 0, rl, cm + torch.log(cs), cnt=cnt)
# plain GEMM epilogue, cast/transpose, L2 normalise, token head, ranking
A = torch.randn(200, 128, device=dev, generator=g).to(torch.bfloat16)
Bm = torch.randn(77, 128, device=dev, generator=g).to(torch.bfloat16)
C = K.gemm_nt(A, Bm)
K.cast_transpose(torch.randn(33, 50, device=dev, generator=g), torch.bfloat16)
from colxlip_b200.model import TokenHead, normalize_tokens           # noqa: E402
normalize_tokens(torch.randn(37, 9, 64, device=dev, generator=g))
head = TokenHead(256, 256).to(dev)
with torch.autocast("cuda", dtype=torch.bfloat16):
    head(torch.randn(5, 7, 256, device=dev, generator=g))
from colxlip_b200.retrieval import compute_retrieval                 # noqa: E402
compute_retrieval(torch.randn(50, 50, device=dev, generator=g), {i: i for i in range(50)}, {i: [i] for i in range(50)})
torch.cuda.synchronize()
print("sanitize_probe: all launches finished, kernels launched:", lib.cxl_kernel_launches())
