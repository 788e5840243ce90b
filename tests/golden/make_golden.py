"""Generate golden vectors by EXECUTING THE REFERENCE in the build container.

Run once (where /root/reference exists):  python tests/golden/make_golden.py
Outputs small .npz fixtures next to this file.  The GPU box has no /root/reference,
so tests only ever read the committed fixtures.

The reference's loss module is loaded by path (its package __init__ needs open_clip,
which is absent); the plain-mean scorer of model.py:626-628 cannot be imported for the
same reason, so for that one the three reference lines are evaluated through the same
torch calls (einsum string, torch.max(dim=3)[0], .mean(dim=-1)).
"""
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src/colxlip/loss.py"


def load_ref():
    spec = importlib.util.spec_from_file_location("ref_loss", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def synth(B, P, T, D, seed, pad_text=False, Bt=None):
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from oracle.colxlip_oracle import synth_inputs
    return synth_inputs(B, P, T, D, seed=seed, pad_text=pad_text, Bt=Bt)


def case_loss(ref, name, B, P, T, D, seed, alpha, scale, pad_text=False, adversarial=False):
    img, txt, tok_img, tok_txt = synth(B, P, T, D, seed, pad_text)
    if adversarial:
        tok_img = tok_img.clone()
        tok_txt = tok_txt.clone()
        tok_img[:, 1] = tok_img[:, 0]            # duplicated patches -> exact ties
        tok_txt[0] = 0.0                         # one all-zero text  -> count = 1e-8
        tok_txt[1] = -tok_img[1, :1].expand(T, D) * 0.5 + 0.0  # strongly negative vs image 1
    xs = [t.clone().requires_grad_() for t in (img, txt, tok_img, tok_txt)]
    ls = torch.tensor(scale, requires_grad=True)
    loss = ref.ColClipLoss(alpha=alpha)
    logits = loss.get_logits(*[x.detach() for x in xs], ls.detach())
    out = loss(*xs, ls, output_dict=True)
    out["total_loss"].backward()
    S = ref.compute_colbert_similarity(tok_img, tok_txt)
    sim = torch.einsum("mnd,kqd->mknq", tok_txt, tok_img)
    arg = torch.max(sim, dim=3)[1]
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        B=B, P=P, T=T, D=D, seed=seed, alpha=alpha, scale=scale,
        pad_text=pad_text, adversarial=adversarial,
        image_features=img.numpy(), text_features=txt.numpy(),
        token_image_features=tok_img.numpy(), token_text_features=tok_txt.numpy(),
        S=S.numpy(), argmax=arg.numpy().astype(np.int16),
        logits_per_image=logits["logits_per_image"].numpy(),
        logits_per_text_token=logits["logits_per_text_token"].numpy(),
        global_contrastive_loss=out["global_contrastive_loss"].item(),
        token_contrastive_loss=out["token_contrastive_loss"].item(),
        total_loss=out["total_loss"].item(),
        g_image_features=xs[0].grad.numpy(), g_text_features=xs[1].grad.numpy(),
        g_token_image_features=xs[2].grad.numpy(), g_token_text_features=xs[3].grad.numpy(),
        g_logit_scale=ls.grad.item())
    print(name, {k: round(v.item(), 6) for k, v in out.items()})


def case_mean(name, Bi, Bt, P, T, D, seed):
    _, _, tok_img, tok_txt = synth(Bi, P, T, D, seed, Bt=Bt)
    sim = torch.einsum("mnd,kqd->mknq", tok_txt, tok_img)       # model.py:626
    mx = torch.max(sim, dim=3)[0]                               # model.py:627
    S = mx.mean(dim=-1)                                         # model.py:628
    np.savez_compressed(os.path.join(HERE, name + ".npz"), Bi=Bi, Bt=Bt, P=P, T=T, D=D, seed=seed,
                        token_image_features=tok_img.numpy(), token_text_features=tok_txt.numpy(),
                        S=S.numpy())
    print(name, S.shape, float(S.mean()))


def case_bias(ref, name):
    img, txt, tok_img, tok_txt = synth(6, 5, 4, 16, 7)
    out = ref.ColClipLoss().get_logits(img, txt, tok_img, tok_txt, torch.tensor(10.0),
                                       logit_bias=torch.tensor(0.25))
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        image_features=img.numpy(), text_features=txt.numpy(),
                        token_image_features=tok_img.numpy(), token_text_features=tok_txt.numpy(),
                        logits_per_image=out["logits_per_image"].numpy(),
                        logits_per_text=out["logits_per_text"].numpy(),
                        logits_per_text_token=out["logits_per_text_token"].numpy())
    print(name, "bias case done")


def main():
    torch.set_num_threads(8)
    ref = load_ref()
    # small enough to commit (< 1 MB each); D kept at multiples of 64 for the tcgen05 path
    case_loss(ref, "loss_b8_p10_t7_d64", 8, 10, 7, 64, 11, 0.5, 1 / 0.07)
    case_loss(ref, "loss_b12_p49_t9_d64_pad", 12, 49, 9, 64, 12, 0.7, 20.0, pad_text=True)
    case_loss(ref, "loss_b10_p17_t5_d128_adv", 10, 17, 5, 128, 13, 0.3, 50.0, adversarial=True)
    case_loss(ref, "loss_b6_p200_t77_d64", 6, 200, 77, 64, 14, 0.5, 1 / 0.07, pad_text=True)
    case_mean("mean_bi7_bt9_p33_t12_d64", 7, 9, 33, 12, 64, 21)
    case_bias(ref, "logits_bias_b6")
    # headline scalar of SURVEY/BASELINE: cfg-1 shape, seed 1234 -> total 4.294426
    img, txt, tok_img, tok_txt = synth(64, 49, 77, 512, 1234)
    out = ref.ColClipLoss(alpha=0.5)(img, txt, tok_img, tok_txt, torch.tensor(1 / 0.07), output_dict=True)
    np.savez_compressed(os.path.join(HERE, "cfg1_scalars.npz"),
                        **{k: v.item() for k, v in out.items()})
    print("cfg1", {k: v.item() for k, v in out.items()})


if __name__ == "__main__":
    main()
