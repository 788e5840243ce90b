"""Not a test (pytest does not collect it): the reference's ALGORITHM (oracle port of loss.py: einsum -> max ->
masked mean -> 4 cross-entropies, autograd backward) run as eager PyTorch ON the B200 under bf16 autocast, next to
this repo's kernels at the same shape -- the "honest competitor" of SURVEY 8(d).  The reference materialises the
[B,B,T,P] similarity tensor, so B is limited to about 1024 at P=196.

    python tests/perf_torch_eager_on_gpu.py [--B 1024]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import colxlip_oracle as O  # noqa: E402
from colxlip_b200 import ColClipLoss  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=1024)
ap.add_argument("--P", type=int, default=196)
ap.add_argument("--T", type=int, default=77)
ap.add_argument("--D", type=int, default=512)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
dev = torch.device("cuda:0")
feats = [t.to(dev) for t in O.synth_inputs(a.B, a.P, a.T, a.D, seed=1234)]


def timed(fn):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loss = fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), float(loss)


def eager():
    xs = [t.clone().requires_grad_() for t in feats]
    ls = torch.tensor(1 / 0.07, device=dev, requires_grad=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = O.ColClipLossOracle(0.5)(*xs, ls, output_dict=True)
    out["total_loss"].backward()
    return out["total_loss"].detach()


crit = ColClipLoss(alpha=0.5)


def ours():
    xs = [t.clone().requires_grad_() for t in feats]
    ls = torch.tensor(1 / 0.07, device=dev, requires_grad=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = crit(*xs, ls, output_dict=True)
    out["total_loss"].backward()
    return out["total_loss"].detach()


te, le = timed(eager)
peak = torch.cuda.max_memory_allocated() / 2**30
to, lo = timed(ours)
print(json.dumps({"shape": [a.B, a.P, a.T, a.D], "torch_eager_ms": te, "ours_ms": to, "speedup": te / to,
                  "eager_peak_GiB": peak, "loss_eager": le, "loss_ours": lo}))
