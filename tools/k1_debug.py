"""First-contact diagnostics for the tcgen05 kernel on a real B200 (run under gpurun).
Prints error patterns of the raw accumulator tile so descriptor/swizzle mistakes can be read off."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from colxlip_b200.kernels import default_kernels

K = default_kernels()
print("device", torch.cuda.get_device_name(0), torch.cuda.get_device_capability(0), flush=True)
g = torch.Generator().manual_seed(0)
for ra, rb, D in [(128, 208, 64), (128, 208, 128), (128, 256, 512)]:
    a = torch.randn(ra, D, generator=g).to(torch.bfloat16).cuda()
    b = torch.randn(rb, D, generator=g).to(torch.bfloat16).cuda()
    out = K.debug_tile(a, b)
    torch.cuda.synchronize()
    out = out[:ra, :rb].cpu()
    ref = a.float().cpu() @ b.float().cpu().T
    err = (out - ref).abs()
    print(f"tile ra={ra} rb={rb} D={D}: max err {err.max():.4e}  ref absmax {ref.abs().max():.3f}", flush=True)
    if err.max() > 1e-2:
        bad_rows = (err.max(dim=1).values > 1e-2).nonzero().flatten().tolist()
        bad_cols = (err.max(dim=0).values > 1e-2).nonzero().flatten().tolist()
        print("  bad rows", bad_rows[:40], "n=", len(bad_rows))
        print("  bad cols", bad_cols[:40], "n=", len(bad_cols))
        print("  out[0,:8]", out[0, :8].tolist())
        print("  ref[0,:8]", ref[0, :8].tolist())
        # does each 64-wide K block contribute? compare against partial sums
        for kb in range(D // 64):
            part = a.float().cpu()[:, kb * 64:(kb + 1) * 64] @ b.float().cpu()[:, kb * 64:(kb + 1) * 64].T
            print(f"  |out - only_kb{kb}| max = {(out - part).abs().max():.3e}")
        for k16 in range(min(D // 16, 8)):
            part = a.float().cpu()[:, k16 * 16:(k16 + 1) * 16] @ b.float().cpu()[:, k16 * 16:(k16 + 1) * 16].T
            print(f"  |out - only_k16_{k16}| max = {(out - part).abs().max():.3e}")
print("debug done", flush=True)
