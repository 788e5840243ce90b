"""One rank of an N-GPU ColClipLoss step on ONE GPU: per-rank memory and kernel time of a multi-GPU workload
(e.g. BASELINE config 5: B = 65536, P = 576, D = 768 over 8 GPUs -> this rank holds b = 8192 texts against all
65536 images) without the other N-1 GPUs.  torch.distributed is replaced by a stub inside colxlip_b200.loss:
broadcasts of remote shards copy the local shard (same bytes, same work), reductions are no-ops.  The numbers are
what a real rank computes and allocates; only the NVLink traffic is missing.

    python tools/rank_emulation.py --workload cfg5 --world 8 [--steps 1] [--b 8192]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import WORKLOADS, synth_local


class _StubDist:
    """The handful of torch.distributed calls colxlip_b200.loss makes, for `world` ranks emulated by rank 0."""
    class ReduceOp:
        SUM, MAX, MIN = "sum", "max", "min"

    def __init__(self, world):
        self.world, self.local = world, None

    def broadcast(self, t, src=0, group=None):
        if src == 0:
            self.local = t
        else:
            t.copy_(self.local)

    def all_gather_into_tensor(self, out, x, group=None):
        out.view((self.world,) + tuple(x.shape)).copy_(x.unsqueeze(0).expand((self.world,) + tuple(x.shape)))

    def all_reduce(self, t, op=None, group=None):
        return None

    def reduce(self, t, dst=0, group=None):
        return None

    def reduce_scatter_tensor(self, out, x, op=None, group=None):
        out.copy_(x[:out.shape[0]])

    def get_backend(self, group=None):
        return "nccl"

    def get_rank(self, group=None):
        return 0

    def get_global_rank(self, group, r):
        return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg5", choices=sorted(WORKLOADS))
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--b", type=int, default=0, help="per-rank batch (default: global batch / world)")
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=0)
    a = ap.parse_args()
    B, P, T, D, desc = WORKLOADS[a.workload]
    b = a.b or B // a.world
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    from colxlip_b200 import ColClipLoss
    from colxlip_b200 import loss as L
    L.dist = _StubDist(a.world)
    free0, total = torch.cuda.mem_get_info()
    feats = list(synth_local(b, P, T, D, 0, dev, torch.bfloat16))
    for x in feats:
        x.requires_grad_()
    ls = torch.tensor(1 / 0.07, device=dev, requires_grad=True)
    crit = ColClipLoss(alpha=0.5, rank=0, world_size=a.world)
    times = []
    for it in range(a.warmup + a.steps):
        for x in feats:
            x.grad = None
        ls.grad = None
        torch.cuda.synchronize()
        torch.cuda.reset_peak_memory_stats()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        t0 = time.time()
        e[0].record()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = crit(*feats, ls, output_dict=True)
        e[1].record()
        out["total_loss"].backward()
        e[2].record()
        torch.cuda.synchronize()
        if it >= a.warmup:
            times.append((e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), time.time() - t0))
    fwd = sum(t[0] for t in times) / len(times)
    bwd = sum(t[1] for t in times) / len(times)
    step_flops_rank = 2.0 * b * (b * a.world) * T * P * D
    res = {"workload": desc, "emulated_world": a.world, "per_rank_batch": b, "global_batch": b * a.world,
           "fwd_ms": fwd, "bwd_ms": bwd, "step_ms": fwd + bwd,
           "pairs_per_s_all_ranks": b * a.world / ((fwd + bwd) * 1e-3),
           "fwd_tflops_rank": step_flops_rank / (fwd * 1e-3) / 1e12,
           "peak_allocated_gib": torch.cuda.max_memory_allocated() / 2 ** 30,
           "peak_reserved_gib": torch.cuda.max_memory_reserved() / 2 ** 30,
           "device_total_gib": total / 2 ** 30, "loss": out["total_loss"].item(),
           "grads_finite": all(bool(torch.isfinite(x.grad.float()).all()) for x in feats),
           "note": "one rank of the N-GPU step on one GPU; collectives stubbed (remote shards = copies of the local one)"}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
