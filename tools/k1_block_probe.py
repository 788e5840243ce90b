"""K1 image-block size (CXL_K1_IMG_BLOCK, read once per process) at the per-rank shape of cfg-2 on 8 GPUs:
b = 4096 texts x 32768 images, one launch.  Run once per value; prints the best of 3."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from colxlip_b200.kernels import default_kernels
K = default_kernels()
b, B, P, T, D = 4096, 32768, 196, 77, 512
g = torch.Generator(device="cuda").manual_seed(1)
nrm = torch.nn.functional.normalize
ti = torch.empty(B, P, D, device="cuda", dtype=torch.bfloat16)
for i in range(0, B, 4096):
    ti[i:i + 4096] = nrm(torch.randn(4096, P, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
tt = nrm(torch.randn(b, T, D, device="cuda", generator=g), dim=-1).to(torch.bfloat16)
ts = []
for it in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = K.maxsim_fwd(tt, ti, T, P, "bf16", 0, True); e1.record(); torch.cuda.synchronize()
    del out
    if it: ts.append(e0.elapsed_time(e1))
print(f"CXL_K1_IMG_BLOCK={os.environ.get('CXL_K1_IMG_BLOCK', '(default)')}: K1 [4096 x 32768] best {min(ts):.1f} ms "
      f"= {2.0 * b * B * T * P * D / min(ts) / 1e9:.0f} TFLOP/s")
