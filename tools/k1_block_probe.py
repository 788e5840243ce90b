"""K1 image-block size (CXL_K1_IMG_BLOCK, read once per process): one K1 launch of --b texts x --bi images at the
cfg-2 token geometry.  Defaults = the per-rank shape of cfg-2 on 8 GPUs before the shard pipeline (4096 x 32768);
--bi 4096 is one shard of the pipeline, --b 16384/32768 the single-GPU shape.  Run once per value; prints the best."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from colxlip_b200.kernels import default_kernels

ap = argparse.ArgumentParser()
ap.add_argument("--b", type=int, default=4096)
ap.add_argument("--bi", type=int, default=32768)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
K = default_kernels()
b, B, P, T, D = a.b, a.bi, 196, 77, 512
g = torch.Generator(device="cuda").manual_seed(1)
nrm = torch.nn.functional.normalize
ti = torch.empty(B, P, D, device="cuda", dtype=torch.bfloat16)
for i in range(0, B, 4096):
    ti[i:i + 4096] = nrm(torch.randn(min(4096, B - i), P, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
tt = torch.empty(b, T, D, device="cuda", dtype=torch.bfloat16)
for i in range(0, b, 4096):
    tt[i:i + 4096] = nrm(torch.randn(min(4096, b - i), T, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
ts = []
for it in range(a.reps + 1):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = K.maxsim_fwd(tt, ti, T, P, "bf16", 0, True)
    e1.record()
    torch.cuda.synchronize()
    del out
    if it:
        ts.append(e0.elapsed_time(e1))
print(f"CXL_K1_IMG_BLOCK={os.environ.get('CXL_K1_IMG_BLOCK', '(default)')}: K1 [{b} x {B}] best {min(ts):.1f} ms "
      f"(all: {' '.join(f'{t:.1f}' for t in ts)}) = {2.0 * b * B * T * P * D / min(ts) / 1e9:.0f} TFLOP/s")
