"""Debug aid: fast backward kernels vs oracle closed form (bf16-rounded w) at one shape; prints error maps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from colxlip_b200.kernels import default_kernels
from oracle import colxlip_oracle as O
Bt, Bi, T, P, D = [int(x) for x in sys.argv[1:6]]
K = default_kernels()
_, _, ti, tt = O.synth_inputs(Bi, P, T, D, seed=77 + P, pad_text=True, Bt=Bt)
ti, tt = ti.bfloat16().float(), tt.bfloat16().float()
txt_op = K.prepare(tt.cuda(), T, "bf16"); img_op = K.prepare(ti.cuda(), K.img_row_stride(P, "bf16"), "bf16")
S, cnt, arg = K.maxsim_fwd(txt_op, img_op, T, P, "bf16", O.MASKED, True)
dS = torch.randn(Bt, Bi, generator=torch.Generator().manual_seed(5))
w = K.pair_weights(dS.cuda(), cnt, T, O.MASKED)
dtxt, dimg = K.maxsim_bwd(txt_op, img_op, arg, w, T, P, "bf16")
gtxt, gimg = K.maxsim_bwd(txt_op, img_op, arg, w, T, P, "bf16", generic=True)
torch.cuda.synchronize()
btxt, bimg = O.maxsim_closed_form_grads(ti, tt, dS, O.MASKED, round_w_bf16=True)
for name, ours, gen, ref in (("dtxt", dtxt, gtxt, btxt), ("dimg", dimg[:, :P], gimg[:, :P], bimg)):
    e = (ours.cpu() - ref).abs()
    eg = (gen.cpu() - ref).abs()
    print(name, "max|ref|", ref.abs().max().item(), "fast err", e.max().item(), "generic err", eg.max().item())
    bad = (e > 1e-4 * ref.abs().max()).nonzero()
    print("  bad elems", bad.shape[0], "of", e.numel(), "first", bad[:8].tolist())
    if bad.shape[0]:
        print("  bad dim0 uniq", bad[:, 0].unique()[:20].tolist(), "dim1 uniq", bad[:, 1].unique()[:20].tolist())
