"""The oracle restatement vs vectors produced by executing the reference (tests/golden/make_golden.py)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import colxlip_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
LOSS_CASES = sorted(glob.glob(os.path.join(GOLDEN, "loss_*.npz")))


def _t(z, k):
    return torch.from_numpy(z[k])


@pytest.mark.parametrize("path", LOSS_CASES, ids=[os.path.basename(p)[:-4] for p in LOSS_CASES])
def test_loss_and_grads_match_reference(path):
    z = np.load(path)
    losses, grads = O.loss_and_grads(_t(z, "image_features"), _t(z, "text_features"),
                                     _t(z, "token_image_features"), _t(z, "token_text_features"),
                                     float(z["scale"]), float(z["alpha"]))
    for k in ("global_contrastive_loss", "token_contrastive_loss", "total_loss"):
        assert abs(losses[k].item() - float(z[k])) <= 1e-5 * max(1.0, abs(float(z[k])))
    for k in ("image_features", "text_features", "token_image_features", "token_text_features"):
        ref = _t(z, "g_" + k)
        tol = 1e-5 * ref.abs().max().item() + 1e-9
        assert (grads[k] - ref).abs().max().item() <= tol, k
    assert abs(grads["logit_scale"].item() - float(z["g_logit_scale"])) <= 1e-5 * max(1e-3, abs(float(z["g_logit_scale"])))


@pytest.mark.parametrize("path", LOSS_CASES, ids=[os.path.basename(p)[:-4] for p in LOSS_CASES])
def test_scores_argmax_match_reference(path):
    z = np.load(path)
    img, txt = _t(z, "token_image_features"), _t(z, "token_text_features")
    S = O.maxsim_scores(img, txt, O.MASKED)
    torch.testing.assert_close(S, _t(z, "S"), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(O.maxsim_scores_blockwise(img, txt, O.MASKED, block=4), _t(z, "S"),
                               rtol=1e-5, atol=1e-6)
    mx, arg = O.maxsim_rowmax(img, txt)
    ref_arg = torch.from_numpy(z["argmax"].astype(np.int64))
    bad = arg != ref_arg
    if bad.any():  # GEMM blocking differs from einsum's: only exact-tie positions may differ
        gap = O.top2_gap(img, txt)
        assert (gap[bad] <= 1e-6).all()


@pytest.mark.parametrize("path", LOSS_CASES, ids=[os.path.basename(p)[:-4] for p in LOSS_CASES])
def test_closed_form_backward(path):
    """The closed-form backward in SURVEY 8(a) reproduces autograd of the reference."""
    z = np.load(path)
    img, txt = _t(z, "token_image_features"), _t(z, "token_text_features")
    S = O.maxsim_scores(img.clone().requires_grad_(), txt.clone().requires_grad_())
    dS = torch.randn(S.shape, generator=torch.Generator().manual_seed(3))
    xi, xt = img.clone().requires_grad_(), txt.clone().requires_grad_()
    O.maxsim_scores(xi, xt).backward(dS)
    dtxt, dimg = O.maxsim_closed_form_grads(img, txt, dS)
    assert (dtxt - xt.grad).abs().max() <= 1e-5 * xt.grad.abs().max() + 1e-9
    assert (dimg - xi.grad).abs().max() <= 1e-5 * xi.grad.abs().max() + 1e-9


def test_plain_mean_variant():
    z = np.load(os.path.join(GOLDEN, "mean_bi7_bt9_p33_t12_d64.npz"))
    S = O.maxsim_scores(_t(z, "token_image_features"), _t(z, "token_text_features"), O.MEAN)
    torch.testing.assert_close(S, _t(z, "S"), rtol=1e-5, atol=1e-6)


def test_logit_bias_quirk():
    z = np.load(os.path.join(GOLDEN, "logits_bias_b6.npz"))
    out = O.ColClipLossOracle().get_logits(_t(z, "image_features"), _t(z, "text_features"),
                                           _t(z, "token_image_features"), _t(z, "token_text_features"),
                                           torch.tensor(10.0), logit_bias=torch.tensor(0.25))
    torch.testing.assert_close(out["logits_per_image"], _t(z, "logits_per_image"), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(out["logits_per_text"], _t(z, "logits_per_text"), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(out["logits_per_text_token"], _t(z, "logits_per_text_token"), rtol=1e-5, atol=1e-5)


def test_cfg1_headline_scalar():
    """SURVEY/BASELINE: cfg-1 shape, seed 1234 -> total loss 4.294426."""
    z = np.load(os.path.join(GOLDEN, "cfg1_scalars.npz"))
    img, txt, ti, tt = O.synth_inputs(64, 49, 77, 512, seed=1234)
    out = O.ColClipLossOracle(0.5)(img, txt, ti, tt, torch.tensor(1 / 0.07), output_dict=True)
    for k in out:
        assert abs(out[k].item() - float(z[k])) < 2e-5
    assert abs(out["total_loss"].item() - 4.294426) < 2e-5


def test_compute_retrieval_identity():
    g = torch.Generator().manual_seed(0)
    s = torch.randn(20, 20, generator=g) + 5 * torch.eye(20)
    r = O.compute_retrieval(s, list(range(20)), [[i] for i in range(20)])
    assert r["text_to_image_R@1"] == 1.0 and r["image_to_text_R@1"] == 1.0
    assert r["text_to_image_mean_rank"] == 1.0


def test_blockwise_oracle_equals_the_plain_restatement():
    """loss_and_grads_blockwise (what the GPU tests evaluate on the device at B = 1024) against loss_and_grads
    (autograd of the restatement that is pinned by the reference's own vectors above)."""
    img, txt, ti, tt = O.synth_inputs(7, 9, 6, 16, seed=21, pad_text=True)
    ti[2, 3] = ti[2, 5]                                   # duplicated patch: first index must win in both
    l0, g0 = O.loss_and_grads(img, txt, ti, tt, 13.0, 0.3)
    l1, g1, ex = O.loss_and_grads_blockwise(img, txt, ti, tt, 13.0, 0.3)
    for k in l0:
        assert abs(l0[k].item() - l1[k].item()) < 1e-6, k
    for k in g0:
        assert (g0[k] - g1[k]).abs().max().item() <= 1e-6 * max(1e-3, g0[k].abs().max().item()), k
    assert torch.equal(ex["arg"], O.maxsim_rowmax(ti, tt)[1])
    dS = torch.randn(7, 7, generator=torch.Generator().manual_seed(2))
    a_txt, a_img = O.maxsim_closed_form_grads(ti, tt, dS)
    cnt = (O.maxsim_rowmax(ti, tt)[0] != 0).float().sum(dim=2) + 1e-8
    b_txt, b_img = O.maxsim_closed_form_grads_by_text(ti, tt, dS / cnt, ex["arg"])
    assert (a_txt - b_txt).abs().max().item() < 1e-6 and (a_img - b_img).abs().max().item() < 1e-6
