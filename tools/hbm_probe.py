"""One pass over the HBM-bound kernels at cfg-2 sizes (for `ncu --metrics gpu__time_duration.sum,dram__bytes_*`):
K3 stats + grad on a [32768, 32768] logit block with cnt, rows-only variants, l2norm fwd/bwd on one rank's image
tokens, the rank kernel on a 32768^2 score matrix, and the bf16 pack."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from colxlip_b200.kernels import default_kernels

K = default_kernels()
torch.cuda.set_device(0)
B = int(os.environ.get("PROBE_B", 32768))
g = torch.Generator(device="cuda").manual_seed(0)
X = torch.rand(B, B, device="cuda", generator=g) * 2 - 1
cnt = torch.full((B, B), 77.0, device="cuda")
scale = torch.tensor([1 / 0.07], device="cuda")
coef = torch.tensor([0.5], device="cuda")
for _ in range(2):
    rl, cm, cs, dg = K.infonce_stats(X, scale, 0)
    cl = cm + torch.log(cs)
    w, ds = K.infonce_grad(X, scale, coef, 0, rl, cl, cnt=cnt)
    K.infonce_stats(X, scale, 0, rows_only=True)
    K.infonce_grad(X, scale, coef, 0, rl, None, out=w)
    K.rank_of_target(X, torch.arange(B, device="cuda", dtype=torch.int32))
    K.rank_of_target(X.t(), torch.arange(B, device="cuda", dtype=torch.int32))
del X, cnt, w
x = torch.randn(4096, 196, 512, device="cuda", generator=g)
for _ in range(2):
    y, inv = K.l2norm_fwd(x, torch.bfloat16)
    dx = K.l2norm_bwd(y, y, inv, torch.float32)
    K.prepare(x, 196, "bf16")
torch.cuda.synchronize()
print("probe done")
