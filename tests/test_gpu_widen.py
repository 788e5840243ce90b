"""GPU parity of the SURVEY 8(f) rows (through the C-ABI): ClipLoss and the rows-only InfoNCE kernels (f-1),
the MaxSim retrieval loop (f-2), normalize_tokens (f-3) and the gradient-accumulation slice backward (f-4).
`-m gpu`.  Runs after test_gpu_parity.py (file order) so the core path is checked first."""
import os
from contextlib import nullcontext

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import colxlip_oracle as O

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
NAMES = ["image_features", "text_features", "token_image_features", "token_text_features"]


@pytest.fixture(scope="module")
def K():
    from colxlip_b200.kernels import default_kernels
    return default_kernels()


def _rounded(x):
    return x.to(torch.bfloat16).float()


# ---- f-1 ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B_loc,B,row0", [(7, 7, 0), (16, 64, 32), (130, 1030, 900)])
def test_infonce_rows_only_kernels(K, B_loc, B, row0):
    from tests.ref_kernels import TorchRefKernels
    R = TorchRefKernels()
    g = torch.Generator().manual_seed(B + 1)
    X = torch.rand(B_loc, B, generator=g) * 2 - 1
    scale, coef = torch.tensor([23.0]), torch.tensor([1.7])
    rl, cm, cs, dg = K.infonce_stats(X.cuda(), scale.cuda(), row0, rows_only=True)
    rl_r, _, _, dg_r = R.infonce_stats(X, scale, row0, rows_only=True)
    assert cm is None and cs is None
    torch.testing.assert_close(rl.cpu(), rl_r, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dg.cpu(), dg_r, rtol=1e-6, atol=1e-6)
    dX, dsc = K.infonce_grad(X.cuda(), scale.cuda(), coef.cuda(), row0, rl_r.cuda(), None)
    dX_r, dsc_r = R.infonce_grad(X, scale, coef, row0, rl_r, None)
    assert (dX.cpu() - dX_r).abs().max() <= 2e-5 * dX_r.abs().max()
    assert abs(dsc.item() - dsc_r.item()) <= 1e-4 * max(abs(dsc_r.item()), 1e-3)
    # rows of a row-softmax gradient sum to zero
    assert dX.sum(dim=1).abs().max().item() <= 1e-5 * dX_r.abs().max().item() * B ** 0.5


@pytest.mark.parametrize("autocast", [False, True])
@pytest.mark.parametrize("B,D", [(9, 32), (257, 512)])
def test_cliploss_vs_oracle(autocast, B, D):
    from colxlip_b200 import ClipLoss
    img, txt, _, _ = O.synth_inputs(B, 2, 2, D, seed=5 + B)
    xs = [t.clone().cuda().requires_grad_() for t in (img, txt)]
    ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    crit = ClipLoss(cache_labels=True)
    if autocast:
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = crit(*xs, ls, output_dict=True)
    else:
        out = crit(*xs, ls, output_dict=True)
    out["total_loss"].backward()
    rnd = _rounded if autocast else (lambda t: t)
    ys = [rnd(t).clone().requires_grad_() for t in (img, txt)]
    ls2 = torch.tensor(1 / 0.07, requires_grad=True)
    ref = O.clip_loss(*ys, ls2)
    ref.backward()
    assert abs(out["total_loss"].item() - ref.item()) <= 1e-5 * max(1.0, abs(ref.item()))
    for x, y in zip(xs, ys):
        assert x.grad.dtype == torch.float32
        # autocast: dL/dlogits enters the two gradient GEMMs in bf16, as in the reference's autocast backward
        # (one bf16 rounding, 2^-9, per term); fp32 mode: library fp32 GEMMs
        assert (x.grad.cpu() - y.grad).abs().max() <= (4e-3 if autocast else 2e-5) * y.grad.abs().max() + 1e-9
    assert abs(ls.grad.item() - ls2.grad.item()) <= 1e-4 * max(1e-3, abs(ls2.grad.item()))


def test_cliploss_vs_reference_run():
    """The reference's ClipLoss on 2 gloo ranks reports the global-batch loss; one GPU on the whole batch must
    give the same value, and gradients equal to the rank-local ones of the reference (gather_with_grad=False)."""
    from colxlip_b200 import ClipLoss
    z = np.load(os.path.join(GOLDEN, "dist_w2.npz"))
    b = int(z["b"])
    xs = [torch.from_numpy(z[n]).cuda().requires_grad_() for n in NAMES[:2]]
    ls = torch.tensor(float(z["scale"]), device="cuda", requires_grad=True)
    loss = ClipLoss()(*xs, ls)
    loss.backward()
    assert abs(loss.item() - float(z["clip_l0_g0_r0_loss"])) <= 1e-5 * abs(float(z["clip_l0_g0_r0_loss"]))
    for r in range(int(z["world"])):
        for x, n in zip(xs, NAMES[:2]):
            ref = torch.from_numpy(z[f"clip_l0_g0_r{r}_g_{n}"])
            assert (x.grad[r * b:(r + 1) * b].cpu() - ref).abs().max() <= 1e-5 * ref.abs().max() + 1e-9
    assert abs(ls.grad.item() - float(z["clip_l0_g0_r0_g_logit_scale"])) <= 2e-5 * abs(float(z["clip_l0_g0_r0_g_logit_scale"]))


# ---- f-4 ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape,m0,n,exact", [
    ((12, 9, 16, 196, 128), 4, 5, True),      # m0*T = 64: TMA-staged kernel on the slice, same bits as the full call
    ((40, 6, 32, 49, 128), 16, 24, True),     # slice spans a 480-row tile boundary
    ((9, 5, 77, 196, 128), 3, 4, False),      # m0*T = 231, not 16-byte aligned -> generic kernel (fp32 weights)
    ((6, 4, 20, 576, 128), 2, 3, False),      # u16 argmax, 80 B offset
    ((7, 8, 5, 10, 64), 2, 3, False),         # shape outside the fast path altogether
])
def test_dtxt_row_slice(K, shape, m0, n, exact):
    Bt, Bi, T, P, D = shape
    _, _, ti, tt = O.synth_inputs(Bi, P, T, D, seed=3 + P, pad_text=True, Bt=Bt)
    txt_op = K.prepare(_rounded(tt).cuda(), T, "bf16")
    img_op = K.prepare(_rounded(ti).cuda(), K.img_row_stride(P, "bf16"), "bf16")
    S, cnt, arg = K.maxsim_fwd(txt_op, img_op, T, P, "bf16", O.MASKED, True)
    w = K.pair_weights(torch.randn(Bt, Bi, generator=torch.Generator().manual_seed(2)).cuda(), cnt, T, O.MASKED)
    full, _ = K.maxsim_bwd(txt_op, img_op, arg, w, T, P, "bf16", need_img=False)
    gen, _ = K.maxsim_bwd(txt_op, img_op, arg, w, T, P, "bf16", need_img=False, generic=True)
    part = K.maxsim_bwd_dtxt_rows(img_op, arg, w, Bt, m0, n, T, P, "bf16")
    part_gen = K.maxsim_bwd_dtxt_rows(img_op, arg, w, Bt, m0, n, T, P, "bf16", generic=True)
    torch.cuda.synchronize()
    assert part.shape == (n, T, D)
    assert torch.equal(part_gen, gen[m0:m0 + n])
    if exact and (P > 256 or D % 128 == 0):
        assert torch.equal(part, full[m0:m0 + n])
    else:
        ref = gen[m0:m0 + n]
        assert (part - ref).abs().max() <= 1e-2 * ref.abs().max()
    # the whole batch as one "slice" is the plain call
    assert torch.equal(K.maxsim_bwd_dtxt_rows(img_op, arg, w, Bt, 0, Bt, T, P, "bf16"), full)


@pytest.mark.parametrize("autocast", [False, True])
def test_colclip_grad_rows_matches_full_backward(autocast):
    """train.py:166-183 pattern on the device: cat(cached, live, cached) with grad_rows == the full backward's
    slice (token grads bit-identical on the fast path), zeros elsewhere, same loss."""
    from colxlip_b200 import ColClipLoss
    B, P, T, D, lo, hi = 24, 196, 16, 128, 8, 16
    feats = [t.cuda() for t in O.synth_inputs(B, P, T, D, seed=77, pad_text=True)]
    ctx = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if autocast else nullcontext

    def run(grad_rows):
        live = [t[lo:hi].clone().requires_grad_() for t in feats]
        cat = [torch.cat([t[:lo], l, t[hi:]]) for t, l in zip(feats, live)]
        ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
        with ctx():
            kw = {"grad_rows": grad_rows} if grad_rows else {}
            out = ColClipLoss(alpha=0.5)(*cat, ls, output_dict=True, **kw)
        out["total_loss"].backward()
        return out["total_loss"].item(), [l.grad for l in live], ls.grad.item()

    l_full, g_full, s_full = run(None)
    l_part, g_part, s_part = run((lo, hi))
    # two separate forwards: K1 / K3 combine partial sums with float atomics, so compare to accumulate-order
    # noise here (bit-exactness of the slice kernels on ONE forward is test_dtxt_row_slice's job)
    assert abs(l_full - l_part) <= 1e-6 * abs(l_full) and abs(s_full - s_part) <= 1e-5 * abs(s_full)
    for a, b_, n in zip(g_part, g_full, NAMES):
        assert a.dtype == b_.dtype and a.shape == b_.shape
        assert (a.float() - b_.float()).abs().max() <= 2e-5 * b_.float().abs().max() + 1e-12, n
    # against the oracle as well
    rnd = _rounded if autocast else (lambda t: t)
    _, g = O.loss_and_grads(*[rnd(t.cpu()) for t in feats], 1 / 0.07, 0.5)
    for a, n in zip(g_part, NAMES):
        ref = g[n][lo:hi]
        assert (a.float().cpu() - ref).abs().max() <= (1e-2 if autocast else 2e-5) * ref.abs().max() + 1e-9, n


# ---- f-3 ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(5, 196, 512), (3, 77, 768), (1000, 8), (2, 3, 4, 1024), (7, 264)])
@pytest.mark.parametrize("in_dt,out_dt", [(torch.float32, torch.float32), (torch.float32, torch.bfloat16),
                                          (torch.bfloat16, torch.bfloat16), (torch.bfloat16, torch.float32)])
def test_normalize_tokens_vs_torch(shape, in_dt, out_dt):
    from colxlip_b200 import normalize_tokens
    g = torch.Generator().manual_seed(sum(shape))
    x0 = (torch.randn(*shape, generator=g) * 3).to(in_dt)
    x0.view(-1, shape[-1])[1] = 0                                   # an all-zero row (masked text token)
    grad = torch.randn(*shape, generator=g)
    x = x0.clone().cuda().requires_grad_()
    y = normalize_tokens(x, out_dtype=out_dt)
    assert y.dtype == out_dt and y.shape == x.shape
    y.backward(grad.cuda().to(out_dt))
    xr = x0.float().clone().requires_grad_()
    yr = F.normalize(xr, dim=-1)
    yr.backward(grad.to(out_dt).float())
    tol_y = 1e-6 if out_dt == torch.float32 else 4e-3              # bf16 output: half an ulp = 2^-9 relative
    assert (y.float().cpu() - yr).abs().max() <= tol_y
    assert x.grad.dtype == in_dt
    live = torch.ones(yr.shape[:-1], dtype=torch.bool)
    live.view(-1)[1] = False
    tol_g = (2e-5 if (in_dt, out_dt) == (torch.float32, torch.float32) else 1.5e-2) * xr.grad[live].abs().max()
    assert (x.grad.float().cpu()[live] - xr.grad[live]).abs().max() <= tol_g
    # zero row: y = 0 and dx = g / eps exactly as torch's clamp_min(eps) gives
    assert y.view(-1, shape[-1])[1].abs().max() == 0
    zr, zo = xr.grad.view(-1, shape[-1])[1], x.grad.float().cpu().view(-1, shape[-1])[1]
    assert (zo - zr).abs().max() <= 1e-2 * zr.abs().max()


def test_normalize_tokens_feeds_the_loss_without_repacking(K):
    """bf16 normalised tokens are the K1 operand itself (no pack pass), and the chain head -> normalise -> loss
    back-propagates to the head output like F.normalize + reference loss on the same (bf16-rounded) operands."""
    from colxlip_b200 import ColClipLoss, normalize_tokens
    B, P, T, D = 10, 196, 12, 128
    g = torch.Generator().manual_seed(9)
    raw = [torch.randn(B, D, generator=g), torch.randn(B, D, generator=g),
           torch.randn(B, P, D, generator=g) * 2, torch.randn(B, T, D, generator=g) * 0.5]
    xs = [t.clone().cuda().requires_grad_() for t in raw]
    ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ys = [normalize_tokens(x) for x in xs]
        assert all(y.dtype == torch.bfloat16 and y.is_contiguous() for y in ys)
        assert K.prepare(ys[2], K.img_row_stride(P, "bf16"), "bf16").data_ptr() == ys[2].data_ptr()
        n0 = K.launches()
        loss = ColClipLoss(alpha=0.5)(*ys, ls)
        assert K.launches() - n0 > 0
    loss.backward()
    xr = [t.clone().requires_grad_() for t in raw]
    yr = [F.normalize(x, dim=-1) for x in xr]
    # the kernels see bf16(y): straight-through rounding keeps autograd flowing to the head output
    yq = [y + (_rounded(y) - y).detach() for y in yr]
    ref = O.ColClipLossOracle(0.5)(*yq, torch.tensor(1 / 0.07))
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 2e-5 * abs(ref.item())
    for x, r, n in zip(xs, xr, NAMES):
        assert (x.grad.cpu() - r.grad).abs().max() <= 3e-2 * r.grad.abs().max(), n


# ---- f-2 ---------------------------------------------------------------------------------------------------
def test_retrieval_loop_on_device(K):
    from colxlip_b200 import maxsim_score_matrix, retrieval_on_split
    from tests.test_widen_cpu import _Loader, _ToyModel
    n_img, caps, P, T, D = 40, 2, 196, 31, 64
    _, _, ti, tt = O.synth_inputs(n_img, P, T, D, seed=13, pad_text=True, Bt=n_img * caps)
    # captions correlate with their image so recall is non-trivial
    tt = F.normalize(tt + 0.8 * ti[torch.arange(n_img * caps) // caps][:, :T] * (tt.abs().sum(-1, keepdim=True) > 0), dim=-1)
    ti, tt = _rounded(ti), _rounded(tt)
    M = maxsim_score_matrix(ti.cuda(), tt.cuda())
    Mc = maxsim_score_matrix(ti.cuda(), tt.cuda(), image_chunk=16, text_chunk=33)
    assert M.shape == (n_img, n_img * caps)
    torch.testing.assert_close(M, Mc, rtol=1e-6, atol=1e-6)   # tile alignment of a text changes the sum order
    S = O.maxsim_scores_blockwise(ti, tt, O.MEAN, block=16)
    torch.testing.assert_close(M.cpu(), S.T, rtol=1e-4, atol=2e-6)
    img_ids = torch.randperm(1000, generator=torch.Generator().manual_seed(1))[:n_img]
    img2txt = {int(img_ids[i]): [caps * i + j for j in range(caps)] for i in range(n_img)}
    txt2img = {c: [int(img_ids[c // caps])] for c in range(n_img * caps)}
    img_loader = _Loader([(torch.arange(s, min(s + 16, n_img)).float(), img_ids[s:s + 16]) for s in range(0, n_img, 16)])
    txt_loader = _Loader([(torch.arange(s, min(s + 32, n_img * caps)), torch.arange(s, min(s + 32, n_img * caps)))
                          for s in range(0, n_img * caps, 32)])
    img_loader.num_samples, txt_loader.num_samples = n_img, n_img * caps
    model = _ToyModel(ti.cuda(), tt.cuda()).cuda()
    metrics = retrieval_on_split("", model, txt_loader, img_loader, img2txt, txt2img, None, 0, {"epoch": 0},
                                 torch.device("cuda"), torch.float32, nullcontext)
    ref = O.compute_retrieval(20.0 * M.cpu(), [c // caps for c in range(n_img * caps)],
                              [[caps * i + j for j in range(caps)] for i in range(n_img)])
    for k, v in ref.items():
        assert abs(metrics[k] - v) < 1e-9, k
    assert metrics["num_text_samples"] == n_img * caps and metrics["text_to_image_R@1"] > 0.5


@pytest.mark.parametrize("R,C", [(5, 8), (33, 1000), (70, 1001), (257, 64), (40, 4096)])
def test_rank_of_target_all_layouts(K, R, C):
    """Contiguous rows (16-byte path / scalar path), the transposed view used for t2i, and a strided slice;
    ties resolved by lower index as a stable descending argsort does (train.py:467-469)."""
    from tests.ref_kernels import TorchRefKernels
    ref = TorchRefKernels()
    g = torch.Generator().manual_seed(R * C)
    s = torch.randint(0, 50, (R, C), generator=g).float() / 7.0           # many exact ties
    tgt = torch.randint(0, C, (R,), generator=g, dtype=torch.int32)
    want = ref.rank_of_target(s, tgt)
    sc = s.cuda()
    assert torch.equal(K.rank_of_target(sc, tgt.cuda()).cpu(), want)
    st = s.t().contiguous().cuda().t()                                     # same values, row_stride == 1
    assert st.stride() == (1, R) and torch.equal(K.rank_of_target(st, tgt.cuda()).cpu(), want)
    wide = torch.zeros(R, 2 * C + 3).cuda()
    wide[:, 1:2 * C + 1:2] = sc                                            # col_stride 2, misaligned base
    assert torch.equal(K.rank_of_target(wide[:, 1:2 * C + 1:2], tgt.cuda()).cpu(), want)


@pytest.mark.parametrize("autocast", [False, True])
def test_accumulation_shared_forward_on_device(autocast):
    """ColClipLoss.accumulation + slice_loss == the per-sub-batch full forward/backward of the reference's loop
    (train.py:166-183) when live features equal the cached ones; K1 runs once."""
    from colxlip_b200 import ColClipLoss
    from colxlip_b200.kernels import default_kernels
    B, P, T, D, acc = 32, 196, 16, 128, 4
    sub = B // acc
    feats = [t.cuda() for t in O.synth_inputs(B, P, T, D, seed=78, pad_text=True)]
    ctx = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if autocast else nullcontext
    crit = ColClipLoss(alpha=0.5)
    # reference-style loop through the plain forward
    ls_ref = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    want = []
    for j in range(acc):
        lo, hi = j * sub, (j + 1) * sub
        live = [t[lo:hi].clone().requires_grad_() for t in feats]
        cat = [torch.cat([t[:lo], l, t[hi:]]) for t, l in zip(feats, live)]
        with ctx():
            total = crit(*cat, ls_ref)
        total.backward()
        want.append([l.grad for l in live])
    K = default_kernels()
    n0 = K.launches()
    ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    with ctx():
        state = crit.accumulation(*feats, ls)
    fwd_launches = K.launches() - n0
    for j in range(acc):
        lo, hi = j * sub, (j + 1) * sub
        live = [t[lo:hi].clone().requires_grad_() for t in feats]
        with ctx():
            out = state.slice_loss(lo, hi, *live, ls, output_dict=True, verify=True)
        assert abs(out["total_loss"].item() - total.item()) <= 1e-6 * abs(total.item())
        out["total_loss"].backward()
        for l, w_, n in zip(live, want[j], NAMES):
            assert l.grad.dtype == w_.dtype and l.grad.shape == w_.shape
            assert (l.grad.float() - w_.float()).abs().max() <= 2e-5 * w_.float().abs().max() + 1e-12, (j, n)
    assert abs(ls.grad.item() - ls_ref.grad.item()) <= 1e-5 * abs(ls_ref.grad.item())
    assert fwd_launches > 0


# ---- f-1: the global-embedding GEMMs on the tcgen05 pipeline ------------------------------------------------------
@pytest.mark.parametrize("M,N,Kd", [(300, 700, 512), (128, 256, 64), (1000, 513, 4096), (77, 1030, 768), (2048, 512, 8192)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_gemm_nt_and_cast_transpose(K, M, N, Kd, dtype):
    g = torch.Generator(device="cuda").manual_seed(M + N)
    a = (torch.randn(M, Kd, device="cuda", generator=g) / Kd ** 0.5).to(dtype)
    b = torch.randn(N, Kd, device="cuda", generator=g).to(dtype)
    c = K.gemm_nt(a, b)
    ref = a.float() @ b.float().T                                   # fp32 accumulate of exact 16-bit products
    assert c.shape == (M, N) and c.dtype == torch.float32
    assert (c - ref).abs().max().item() <= 2e-5 * ref.abs().max().item() + 1e-6
    assert K.gemm_nt(a[:, :48].contiguous(), b[:, :48].contiguous()) is None      # K % 64 != 0: caller's library GEMM
    x = torch.randn(M, N, device="cuda", generator=g)
    y, yt = K.cast_transpose(x, dtype)
    assert torch.equal(y, x.to(dtype)) and torch.equal(yt, x.to(dtype).t().contiguous())
    only_t = K.cast_transpose(x, dtype, want_plain=False)
    assert only_t[0] is None and torch.equal(only_t[1], yt)


def test_cliploss_b4096_on_the_tensor_core_gemms():
    """ClipLoss fwd+bwd at B = 4096, D = 512 under bf16 autocast: logits and both gradient GEMMs on cxl_gemm_nt
    (K = 512 resident-A, K = 4096 streamed-A); against torch autograd of the same loss on bf16-rounded features."""
    from colxlip_b200 import ClipLoss
    B, D = 4096, 512
    img, txt, _, _ = O.synth_inputs(B, 2, 2, D, seed=44)
    xs = [t.clone().cuda().requires_grad_() for t in (img, txt)]
    ls = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = ClipLoss()(*xs, ls)
    loss.backward()
    ys = [_rounded(t).cuda().requires_grad_() for t in (img, txt)]
    ls2 = torch.tensor(1 / 0.07, device="cuda", requires_grad=True)
    ref = O.clip_loss(*ys, ls2)
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    for x, y in zip(xs, ys):
        assert (x.grad - y.grad).abs().max() <= 4e-3 * y.grad.abs().max()
        assert ((x.grad - y.grad).norm() / y.grad.norm()).item() <= 4e-3
    assert abs(ls.grad.item() - ls2.grad.item()) <= 1e-4 * abs(ls2.grad.item())


# ---- f-3: token heads + EOT masking + L2 normalise as three launches -----------------------------------------------
@pytest.mark.parametrize("N,T,width,D,dtype", [(37, 77, 512, 512, torch.bfloat16), (9, 196, 768, 512, torch.bfloat16),
                                               (5, 577, 1024, 768, torch.bfloat16), (16, 77, 768, 768, torch.float16)])
def test_fused_token_head(N, T, width, D, dtype):
    import torch.nn as nn
    from colxlip_b200 import TokenHead, encode_tokens
    torch.manual_seed(N)
    head = TokenHead(width, D).cuda()
    with torch.no_grad():
        for p_ in head.parameters():
            p_.add_(0.1 * torch.randn_like(p_))
    ref = nn.Sequential(nn.LayerNorm(width), nn.Linear(width, D), nn.GELU(), nn.LayerNorm(D)).cuda()
    ref.load_state_dict(head.state_dict())
    x = torch.randn(N, T, width, device="cuda") * 2 + 0.3
    vl = torch.randint(1, T + 1, (N,), device="cuda", dtype=torch.int32) if T == 77 else None
    xm = x if vl is None else x * (torch.arange(T, device="cuda")[None, :] < vl[:, None])[..., None]
    with torch.autocast("cuda", dtype=dtype):
        y = encode_tokens(x, head, valid_len=vl, normalize=True)
        y_ac = torch.nn.functional.normalize(ref(xm), dim=-1)        # the reference's op sequence under autocast
    assert y.dtype == dtype and y.shape == (N, T, D)
    y32 = torch.nn.functional.normalize(ref(xm), dim=-1)             # fp32 ground truth
    # operands of the Linear are rounded to 16 bits (as under autocast); everything else is fp32 here, so the fused
    # result is at least as close to fp32 as the reference's own autocast pipeline
    e_fused = (y.float() - y32).abs().max().item()
    e_ac = (y_ac.float() - y32).abs().max().item()
    assert e_fused <= 1e-2 and e_fused <= 1.5 * e_ac + 2e-3, (e_fused, e_ac)
    torch.testing.assert_close(y.float().norm(dim=-1), torch.ones(N, T, device="cuda"), rtol=0, atol=1e-2)
    # goes straight into the scorer without a pack pass
    if vl is not None:
        assert torch.equal(y[0, int(vl[0]):].float().std(dim=0), torch.zeros(D, device="cuda")) or int(vl[0]) >= T - 1
    # backward = autograd of the reference's ops on the saved input (recompute)
    xr = x.clone().requires_grad_()
    with torch.autocast("cuda", dtype=dtype):
        yf = encode_tokens(xr, head, valid_len=vl, normalize=True)
    g = torch.randn_like(yf)
    yf.backward(g)
    xr2 = x.clone().requires_grad_()
    ref.zero_grad()
    with torch.autocast("cuda", dtype=dtype):
        xm2 = xr2 if vl is None else xr2 * (torch.arange(T, device="cuda")[None, :] < vl[:, None])[..., None]
        y2 = torch.nn.functional.normalize(ref(xm2), dim=-1)
    y2.backward(g.to(y2.dtype))
    assert (xr.grad - xr2.grad).abs().max() <= 2e-2 * xr2.grad.abs().max() + 1e-6
    for (n1, p1), (n2, p2) in zip(head.named_parameters(), ref.named_parameters()):
        assert (p1.grad - p2.grad).abs().max() <= 2e-2 * p2.grad.abs().max() + 1e-6, n1


def test_nvtx_ranges_bracket_the_phases_of_a_step(monkeypatch):
    """the fused loss opens one NVTX range per phase (K1, K3 statistics, K3 gradient, K2, last reduction)
    and closes every range it opened."""
    from colxlip_b200 import ColClipLoss
    events = []
    monkeypatch.setattr(torch.cuda.nvtx, "range_push", lambda name: events.append(("push", name)))
    monkeypatch.setattr(torch.cuda.nvtx, "range_pop", lambda: events.append(("pop", None)))
    xs = [t.cuda().requires_grad_() for t in O.synth_inputs(16, 196, 20, 128, seed=3)]
    ls = torch.tensor(10.0, device="cuda", requires_grad=True)
    ColClipLoss()(*xs, ls, output_dict=True)["total_loss"].backward()
    torch.cuda.synchronize()
    names = [n for kind, n in events if kind == "push"]
    assert len(names) == 5 and all(n.startswith("cxl: ") for n in names)
    assert "K1" in names[0] and "K3 statistics" in names[1] and "K2" in names[3]
    depth = 0
    for kind, _ in events:
        depth += 1 if kind == "push" else -1
        assert 0 <= depth <= 1
    assert depth == 0
