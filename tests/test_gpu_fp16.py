"""fp16 operand path (VERDICT r1 missing #2): the reference's only ColXLIP launch script runs `--precision amp`
(colxlip.sh:53) = torch.autocast(fp16) around loss(**model_out) (train.py:131-134), so its einsum at loss.py:34 sees
fp16 operands (11-bit mantissa).  Here: CXL_DTYPE_F16 through the pack kernel, K1 (tcgen05 kind::f16 with F16 operand
formats), K2 (fma.rn.f32.f16) and the L2-normalise cast; the 16-bit pair weights are fp16 scaled by a power of two so
that a GradScaler factor neither overflows nor flushes them.  `-m gpu`."""
import pytest
import torch

from oracle import colxlip_oracle as O

pytestmark = pytest.mark.gpu
TIE_GAP = 2e-6


@pytest.fixture(scope="module")
def K():
    from colxlip_b200.kernels import default_kernels
    return default_kernels()


def _h(x):
    return x.to(torch.float16).float()


SHAPES = [(70, 9, 77, 196, 256), (10, 37, 20, 576, 128), (40, 6, 33, 49, 128), (33, 12, 77, 576, 768), (6, 6, 77, 200, 64)]


@pytest.mark.parametrize("shape", SHAPES, ids=str)
def test_forward_fp16_operands(K, shape):
    Bt, Bi, T, P, D = shape
    _, _, ti, tt = O.synth_inputs(Bi, P, T, D, seed=900 + P, pad_text=True, Bt=Bt)
    ti, tt = _h(ti).cuda(), _h(tt).cuda()
    # inputs that are exact in fp16 but NOT in bf16: the bf16 path would round them
    assert not torch.equal(ti.to(torch.bfloat16).float(), ti)
    txt_op = K.prepare(tt, T, "fp16")
    img_op = K.prepare(ti, K.img_row_stride(P, "fp16"), "fp16")
    assert txt_op.dtype == torch.float16 and img_op.dtype == torch.float16
    S, cnt, arg = K.maxsim_fwd(txt_op, img_op, T, P, "fp16", O.MASKED, True)
    torch.cuda.synchronize()
    mx, ref_arg = O.maxsim_rowmax(ti, tt)                       # fp32 on the device: fp16 products are exact in fp32
    nz = (mx != 0).float().sum(dim=2)
    assert torch.equal(cnt, nz)
    torch.testing.assert_close(S, mx.sum(dim=2) / (nz + 1e-8), rtol=1e-5, atol=2e-6)
    ours = K.argmax_as_text_major(arg, Bt, T)
    bad = ours != ref_arg
    if bad.any():
        assert (O.top2_gap(ti, tt)[bad] <= TIE_GAP).all()
    # the same operands through the bf16 path differ visibly (operands lose 3 mantissa bits): the two paths are distinct
    Sb, _, _ = K.maxsim_fwd(K.prepare(tt, T, "bf16"), K.prepare(ti, K.img_row_stride(P, "bf16"), "bf16"), T, P, "bf16",
                            O.MASKED, False)
    assert (Sb - S).abs().max().item() > 1e-5


@pytest.mark.parametrize("shape", [s for s in SHAPES if s[4] % 128 == 0], ids=str)
@pytest.mark.parametrize("loss_scale", [1.0, 65536.0, 2.0 ** -12])
def test_backward_fp16_streaming_kernels(K, shape, loss_scale):
    """K2 with fp16 operands and fp16 pair weights scaled by a power of two: against the closed form evaluated with the
    same 16-bit weights; the result must not depend on the loss scale beyond that rounding (GradScaler 65536, or a
    tiny upstream gradient that would flush unscaled fp16 weights to zero)."""
    from colxlip_b200.loss import _w_scale_from_bound
    Bt, Bi, T, P, D = shape
    _, _, ti, tt = O.synth_inputs(Bi, P, T, D, seed=77 + P, pad_text=True, Bt=Bt)
    ti, tt = _h(ti).cuda(), _h(tt).cuda()
    txt_op = K.prepare(tt, T, "fp16")
    img_op = K.prepare(ti, K.img_row_stride(P, "fp16"), "fp16")
    S, cnt, arg = K.maxsim_fwd(txt_op, img_op, T, P, "fp16", O.MASKED, True)
    dS = torch.randn(Bt, Bi, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5)) * 1e-4 * loss_scale
    w = K.pair_weights(dS, cnt, T, O.MASKED)
    ws = _w_scale_from_bound(w.abs().amax())
    dtxt = torch.empty(Bt, T, D, device="cuda")
    dimg = K.maxsim_bwd_shard(txt_op, img_op, arg, w, 0, T, P, "fp16", dtxt=dtxt, w_scale=ws)
    torch.cuda.synchronize()
    s = ws[0].item()
    assert s == 2.0 ** round(torch.log2(ws[0]).item()) and abs(ws[0].item() * ws[1].item() - 1.0) < 1e-12
    w16 = (w * s).to(torch.float16).float() / s                  # what the kernels multiply with
    assert torch.isfinite(w16).all() and (w16 != 0).float().mean().item() > 0.99
    a = K.argmax_as_text_major(arg, Bt, T)
    rtxt, rimg = O.maxsim_closed_form_grads_by_text(ti, tt, w16, a)
    for ours, ref, n in ((dtxt, rtxt, "dtxt"), (dimg[:, :P], rimg, "dimg")):
        assert (ours - ref).abs().max().item() <= 2e-5 * ref.abs().max().item(), (n, shape, loss_scale)
    # without the scale a small upstream gradient underflows fp16 (this is what the scale is for)
    if loss_scale < 1e-3:
        d0 = K.maxsim_bwd_shard(txt_op, img_op, arg, w, 0, T, P, "fp16", w_scale=None)
        assert (d0[:, :P] - rimg).abs().max().item() > 1e-2 * rimg.abs().max().item()
    # two shards with shard-local argmax + accumulate == one call (bits)
    if Bi >= 2:
        h = Bi // 2
        S0, c0, a0 = K.maxsim_fwd(txt_op, img_op[:h].contiguous(), T, P, "fp16", O.MASKED, True)
        S1, c1, a1 = K.maxsim_fwd(txt_op, img_op[h:].contiguous(), T, P, "fp16", O.MASKED, True)
        torch.testing.assert_close(torch.cat([S0, S1], dim=1), S, rtol=1e-6, atol=1e-7)   # (S sums are atomics: order-free only to fp32 rounding)
        assert torch.equal(torch.cat([a0, a1])[:, :Bt * T], arg[:, :Bt * T])
        dt2 = torch.empty_like(dtxt)
        p0 = K.maxsim_bwd_shard(txt_op, img_op[:h].contiguous(), a0, w, 0, T, P, "fp16", dtxt=dt2, w_scale=ws)
        p1 = K.maxsim_bwd_shard(txt_op, img_op[h:].contiguous(), a1, w, h, T, P, "fp16", dtxt=dt2, accumulate_txt=True,
                                w_scale=ws)
        assert torch.equal(torch.cat([p0, p1]), dimg) and torch.equal(dt2, dtxt)


@pytest.mark.parametrize("shape", [(48, 196, 77, 512), (24, 576, 77, 768), (40, 49, 12, 64)], ids=str)
@pytest.mark.parametrize("grad_scale", [1.0, 65536.0])
def test_loss_under_fp16_autocast(shape, grad_scale):
    """ColClipLoss under torch.autocast(fp16) (+ a GradScaler-style factor on the loss): losses and all five
    gradients against the oracle on fp16-rounded inputs."""
    from colxlip_b200 import ColClipLoss
    B, P, T, D = shape
    inputs = O.synth_inputs(B, P, T, D, seed=17, pad_text=True)
    xs = [t.clone().cuda().requires_grad_() for t in inputs]
    ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    with torch.autocast("cuda", dtype=torch.float16):
        out = ColClipLoss(alpha=0.5)(*xs, ls, output_dict=True)
    (out["total_loss"] * grad_scale).backward()
    torch.cuda.synchronize()
    r_losses, r_grads = O.loss_and_grads(*[_h(t) for t in inputs], 1 / 0.07, 0.5)
    for k, v in r_losses.items():
        assert abs(out[k].item() - v.item()) <= 2e-5 * max(1.0, abs(v.item())), k
    names = ["image_features", "text_features", "token_image_features", "token_text_features"]
    for x, n in zip(xs, names):
        ref = r_grads[n]
        g = x.grad.float().cpu() / grad_scale
        assert torch.isfinite(g).all()
        # token grads on the streaming kernels (D % 128 == 0) carry fp16-rounded pair weights (2^-11 per term)
        # global grads: dL/dlogits enters the gradient GEMMs in fp16 (2^-11 per term), as in the reference's autocast backward
        tol = 2e-3 if (n.startswith("token") and D % 128 == 0) else (1e-4 if n.startswith("token") else 1e-3)
        assert (g - ref).abs().max().item() <= tol * ref.abs().max().item() + 1e-9, (n, grad_scale)
    assert abs(ls.grad.item() / grad_scale - r_grads["logit_scale"].item()) <= 2e-4 * max(1e-3, abs(r_grads["logit_scale"].item()))


def test_fp16_inputs_dtype_contract_and_normalize():
    from colxlip_b200 import ColClipLoss, compute_colbert_similarity
    from colxlip_b200.model import normalize_tokens
    img, txt, ti, tt = O.synth_inputs(16, 196, 20, 128, seed=8)
    xs = [t.half().cuda().requires_grad_() for t in (img, txt, ti, tt)]
    ls = torch.tensor(20.0, device="cuda", requires_grad=True)
    out = ColClipLoss()(*xs, ls, output_dict=True)              # fp16 inputs, no autocast -> fp16 operand path
    out["total_loss"].backward()
    assert out["total_loss"].dtype == torch.float32
    assert all(x.grad.dtype == torch.float16 and x.grad.shape == x.shape for x in xs)
    S = compute_colbert_similarity(xs[2].detach(), xs[3].detach())
    ref = O.maxsim_scores_blockwise(xs[2].detach().float().cpu(), xs[3].detach().float().cpu(), O.MASKED, 16)
    torch.testing.assert_close(S.cpu(), ref, rtol=1e-4, atol=2e-6)
    # L2 normalise fused with the fp16 cast, and its backward
    x = (torch.randn(37, 9, 64, device="cuda") * 3).requires_grad_()
    with torch.autocast("cuda", dtype=torch.float16):
        y = normalize_tokens(x)
    assert y.dtype == torch.float16
    x2 = x.detach().clone().requires_grad_()
    y2 = torch.nn.functional.normalize(x2, dim=-1)
    torch.testing.assert_close(y.float(), y2, rtol=1e-3, atol=1e-3)
    g = torch.randn_like(y2)
    y.backward(g.half())
    y2.backward(g.half().float())
    torch.testing.assert_close(x.grad, x2.grad, rtol=2e-3, atol=2e-3)
