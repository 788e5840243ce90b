"""K1 over one rank's texts x all images: one launch vs one launch per home-rank shard (what the N-GPU pipeline does)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from colxlip_b200.kernels import default_kernels
K = default_kernels()
b, B, P, T, D, W = 4096, 32768, 196, 77, 512, 8
g = torch.Generator(device="cuda").manual_seed(1)
nrm = torch.nn.functional.normalize
ti = torch.empty(B, P, D, device="cuda", dtype=torch.bfloat16)
for i in range(0, B, 4096):
    ti[i:i + 4096] = nrm(torch.randn(4096, P, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
tt = nrm(torch.randn(b, T, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
def timed(fn, reps=3):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts), sorted(ts)[len(ts) // 2]
one = timed(lambda: K.maxsim_fwd(tt, ti, T, P, "bf16", 0, True))
def split():
    S = torch.empty(b, B, device="cuda"); c = torch.empty(b, B, device="cuda")
    for r in range(W):
        s_, c_, a_ = K.maxsim_fwd(tt, ti[r * b:(r + 1) * b], T, P, "bf16", 0, True)
        S[:, r * b:(r + 1) * b] = s_; c[:, r * b:(r + 1) * b] = c_
sp = timed(split)
print(f"K1 one launch  [4096 x 32768]: best {one[0]:.1f} ms  median {one[1]:.1f} ms")
print(f"K1 8 launches  [4096 x 4096] + S/cnt column copies: best {sp[0]:.1f} ms  median {sp[1]:.1f} ms")
