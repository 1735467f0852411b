"""Generate the golden vectors under tests/golden/ from the UNMODIFIED reference.

Run in the build container only (the reference does not travel to the GPU box):

    PYTHONDONTWRITEBYTECODE=1 PYTHONPATH=/root/reference python tests/golden/make_golden.py

For every model family on the hot path it builds the reference ``npf`` model, draws
seeded inputs, runs forward + loss (+ backward) on CPU in fp32 and stores inputs,
the reference ``state_dict`` and the reference outputs.  Small-width (r_dim=32)
variants keep the fixtures small; the notebook-width configs contribute their
parameter counts / state_dict layouts (``layouts.json``) and two fixtures with the
reference's own pretrained weights (``results/pretrained/RBF_Kernel``).
"""
import json
import math
import os
import sys
import warnings
from functools import partial

import torch

warnings.filterwarnings("ignore")
REF = os.environ.get("NPF_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

import npf  # noqa: E402  (the reference)
from npf import CNP, LNP, AttnCNP, AttnLNP, ConvCNP, ConvLNP, CNPFLoss, ELBOLossLNPF, NLLLossLNPF  # noqa: E402
from npf.architectures import CNN, MLP, ResConvBlock, SetConv, UnetCNN, UnsharedExpRBF, discard_ith_arg, merge_flat_input  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
torch.set_num_threads(4)


# ------------------------------------------------------------------ model builders
def build(name, R, **over):
    xenc = partial(MLP, n_hidden_layers=1, hidden_size=R)
    dec = merge_flat_input(partial(MLP, n_hidden_layers=4, hidden_size=R), is_sum_merge=True)
    xy = lambda h: merge_flat_input(partial(MLP, n_hidden_layers=2, hidden_size=h), is_sum_merge=True)  # noqa: E731
    if name == "CNP":            # jupyter/reproducibility/CNP.ipynb:176-196
        return CNP(x_dim=1, y_dim=1, XYEncoder=xy(2 * R), XEncoder=xenc, Decoder=dec, r_dim=R)
    if name == "AttnCNP":        # AttnCNP.ipynb:163-184
        return AttnCNP(x_dim=1, y_dim=1, XYEncoder=xy(R), is_self_attn=False, r_dim=R,
                       attention=over.get("attention", "transformer"), XEncoder=xenc, Decoder=dec)
    if name == "LNP":            # LNP.ipynb
        return LNP(x_dim=1, y_dim=1, XYEncoder=xy(2 * R), is_q_zCct=True, n_z_samples_train=over.get("nz", 3),
                   n_z_samples_test=over.get("nz", 3), XEncoder=xenc, Decoder=dec, r_dim=R,
                   encoded_path=over.get("encoded_path", "latent"))
    if name == "AttnLNP":        # AttnLNP.ipynb:160-180
        return AttnLNP(x_dim=1, y_dim=1, XYEncoder=xy(R), is_self_attn=False, is_q_zCct=True,
                       n_z_samples_train=over.get("nz", 3), n_z_samples_test=over.get("nz", 3), r_dim=R,
                       attention="transformer")
    if name == "ConvLNP":        # ConvLNP.ipynb:150-200 (1-D): 4 ResConvBlocks, BatchNorm1d, Linear decoder
        cnn = partial(CNN, ConvBlock=ResConvBlock, is_chan_last=True, n_conv_layers=2, n_blocks=4, Conv=torch.nn.Conv1d,
                      Normalization=torch.nn.BatchNorm1d, kernel_size=over.get("kernel_size", 19))
        return ConvLNP(x_dim=1, y_dim=1, Interpolator=SetConv, CNN=cnn, density_induced=over.get("density", 64),
                       is_global=over.get("is_global", True), is_q_zCct=over.get("is_q_zCct", False),
                       encoded_path=over.get("encoded_path", "latent"), n_z_samples_train=over.get("nz", 1),
                       n_z_samples_test=over.get("nz", 1), r_dim=R, Decoder=discard_ith_arg(torch.nn.Linear, i=0))
    if name in ("ConvCNP", "ConvCNP_unet", "ConvCNP_unshared"):
        cnn_kw = dict(ConvBlock=ResConvBlock, is_chan_last=True, n_conv_layers=2, Conv=torch.nn.Conv1d,
                      Normalization=torch.nn.BatchNorm1d, n_blocks=5, kernel_size=over.get("kernel_size", 19))
        if name == "ConvCNP_unet":   # SURVEY.md section 8, C3 definition
            cnn = partial(UnetCNN, Pool=torch.nn.AvgPool1d, upsample_mode="linear", max_nchannels=2 * R, **cnn_kw)
        else:                        # ConvCNP.ipynb:178-211
            cnn = partial(CNN, **cnn_kw)
        interp = partial(SetConv, RadialBasisFunc=UnsharedExpRBF) if name == "ConvCNP_unshared" else SetConv
        return ConvCNP(x_dim=1, y_dim=1, Interpolator=interp, CNN=cnn, density_induced=over.get("density", 64),
                       r_dim=R, Decoder=discard_ith_arg(partial(MLP, n_hidden_layers=4, hidden_size=R), i=0))
    raise KeyError(name)


def perturb(model, seed):
    """Make every tensor non-trivial: zero-initialised biases, LN/BN affine and BN running stats."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k, v in model.state_dict().items():
            if k.endswith("num_batches_tracked"):
                v.fill_(3)
            elif k.endswith("running_var"):
                v.copy_(0.5 + torch.rand(v.shape, generator=g))
            elif k.endswith("running_mean"):
                v.copy_(0.2 * torch.randn(v.shape, generator=g))
            elif k.endswith(".bias") or "norm" in k:
                v.add_(0.1 * torch.randn(v.shape, generator=g))


def chirp_inputs(B, Nc, Nt, seed):
    """Synthetic chirp tasks of SURVEY.md section 8(d); context = random subset of the target grid."""
    g = torch.Generator().manual_seed(seed)
    x = torch.linspace(-1, 1, Nt).view(1, Nt, 1).expand(B, Nt, 1).contiguous()
    tau = (x + 1) / 2
    f0 = 4 + 8 * torch.rand(B, 1, 1, generator=g)
    k = 20 + 60 * torch.rand(B, 1, 1, generator=g)
    ph = 2 * math.pi * torch.rand(B, 1, 1, generator=g)
    y = (0.1 + tau) ** 2 * torch.sin(2 * math.pi * (f0 * tau + 0.5 * k * tau ** 2) + ph)
    y = (y - y.mean(1, keepdim=True)) / y.std(1, keepdim=True)
    idx = torch.stack([torch.randperm(Nt, generator=g)[:Nc] for _ in range(B)]).unsqueeze(-1)
    return dict(X_cntxt=x.gather(1, idx), Y_cntxt=y.gather(1, idx), X_trgt=x, Y_trgt=y)


def dist_fields(p):
    return p.base_dist.loc.detach().clone(), p.base_dist.scale.detach().clone()


ONLY = [a for a in sys.argv[1:] if not a.startswith("-")]     # optional tag substrings: regenerate only those fixtures


def run_case(tag, name, R, B, Nc, Nt, seed, loss_cls, nz=3, state=None, **over):
    if ONLY and not any(f in tag for f in ONLY):
        return
    torch.manual_seed(seed)
    model = build(name, R, nz=nz, **over)
    if state is not None:
        model.load_state_dict(state, strict=True)
    else:
        perturb(model, seed + 1)
    inp = chirp_inputs(B, Nc, Nt, seed + 2)
    latent = name in ("LNP", "AttnLNP", "ConvLNP")
    fix = {"name": name, "r_dim": R, "kwargs": {k: v for k, v in over.items()}, "n_z": nz if latent else 1,
           "inputs": inp, "state_dict": {k: v.detach().clone() for k, v in model.state_dict().items()}}

    if latent:
        # G4: pin epsilon by patching rsample of the sampling distribution through a recorded draw
        n_lat = 1
        if name == "ConvLNP" and over.get("encoded_path", "latent") == "latent":
            n_lat = int(over.get("density", 64) * 3)     # one latent per induced point
        eps = torch.randn(nz, B, n_lat, R, generator=torch.Generator().manual_seed(seed + 3))
        fix["eps"] = eps
        orig = model.latent_path

        def latent_path(X_cntxt, R_, X_trgt, Y_trgt):
            q_zCc = model.infer_latent_dist(X_cntxt, R_)
            if model.is_q_zCct and Y_trgt is not None:
                q_zCct = model.infer_latent_dist(X_trgt, model.encode_globally(X_trgt, Y_trgt))
                sd_ = q_zCct
            else:
                q_zCct, sd_ = None, q_zCc
            z = sd_.base_dist.loc.unsqueeze(0) + sd_.base_dist.scale.unsqueeze(0) * eps   # == rsample with this eps
            return z, q_zCc, q_zCct

        model.latent_path = latent_path
        del orig

    # ---- train mode: forward + loss + backward
    model.train()
    crit = loss_cls()
    crit.train()
    out = model(**inp)
    loss = crit(out, inp["Y_trgt"])
    loss.backward()
    p, z, q_c, q_ct = out
    tr = {"loss": loss.detach().clone()}
    tr["loc"], tr["scale"] = dist_fields(p)
    tr["log_prob"] = p.log_prob(inp["Y_trgt"]).detach().clone()
    if latent:
        tr["z"] = z.detach().clone()
        tr["q_zCc_loc"], tr["q_zCc_scale"] = dist_fields(q_c)
        if q_ct is not None:
            tr["q_zCct_loc"], tr["q_zCct_scale"] = dist_fields(q_ct)
    tr["grads"] = {k: v.grad.detach().clone() for k, v in model.named_parameters() if v.grad is not None}
    tr["state_after"] = {k: v.detach().clone() for k, v in model.state_dict().items() if "running_" in k or "num_batches" in k}
    fix["train"] = tr

    # ---- eval mode: forward (no Y_trgt) + eval loss (always NLL / logsumexp, losses.py:63-69)
    model.load_state_dict(fix["state_dict"])
    model.eval()
    crit.eval()
    crit.reduction = None
    with torch.no_grad():
        out = model(inp["X_cntxt"], inp["Y_cntxt"], inp["X_trgt"])
        ev = {"loss_per_task": crit(out, inp["Y_trgt"]).clone()}
        ev["loc"], ev["scale"] = dist_fields(out[0])
    fix["eval"] = ev
    path = os.path.join(HERE, f"{tag}.pt")
    torch.save(fix, path)
    print(f"{tag:28s} loss={float(tr['loss']):+.5f}  params={sum(p.numel() for p in model.parameters()):,d}  "
          f"{os.path.getsize(path) / 1024:.0f} KiB")


def main():
    R = 32
    run_case("cnp_r32", "CNP", R, 3, 7, 12, 10, CNPFLoss)
    run_case("attncnp_r32", "AttnCNP", R, 3, 9, 20, 20, CNPFLoss)
    run_case("attncnp_multihead_r32", "AttnCNP", R, 2, 5, 11, 21, CNPFLoss, attention="multihead")
    run_case("attncnp_scaledot_r32", "AttnCNP", R, 2, 5, 11, 22, CNPFLoss, attention="scaledot")
    run_case("convcnp_r32", "ConvCNP", R, 3, 10, 24, 30, CNPFLoss, density=16, kernel_size=5)
    run_case("convcnp_unet_r32", "ConvCNP_unet", R, 2, 10, 24, 31, CNPFLoss, density=16, kernel_size=5)
    run_case("convcnp_unshared_r32", "ConvCNP_unshared", R, 2, 10, 24, 32, CNPFLoss, density=16, kernel_size=5)
    run_case("lnp_r32", "LNP", R, 3, 7, 12, 40, ELBOLossLNPF, nz=3)
    run_case("lnp_both_r32", "LNP", R, 2, 7, 12, 41, ELBOLossLNPF, nz=2, encoded_path="both")
    run_case("attnlnp_r32", "AttnLNP", R, 3, 9, 20, 50, ELBOLossLNPF, nz=3)
    run_case("attnlnp_nll_r32", "AttnLNP", R, 2, 6, 10, 51, NLLLossLNPF, nz=4)

    run_case("convlnp_nll_r32", "ConvLNP", R, 2, 10, 24, 70, NLLLossLNPF, nz=3, density=16, kernel_size=5,
             is_global=True, is_q_zCct=False)
    run_case("convlnp_both_r32", "ConvLNP", R, 2, 10, 24, 71, ELBOLossLNPF, nz=2, density=16, kernel_size=5,
             is_global=False, is_q_zCct=True, encoded_path="both")

    # reference's own trained weights at notebook width (fixtures for the headline model + BN statistics)
    pre = os.path.join(REF, "results/pretrained/RBF_Kernel")
    run_case("attncnp_pretrained", "AttnCNP", 128, 2, 16, 48, 60, CNPFLoss,
             state=torch.load(os.path.join(pre, "AttnCNP/run_0/params.pt"), map_location="cpu"))
    run_case("convcnp_pretrained", "ConvCNP", 128, 2, 16, 48, 61, CNPFLoss,
             state=torch.load(os.path.join(pre, "ConvCNP/run_0/params.pt"), map_location="cpu"))

    # notebook-width layouts and parameter counts (known answers: CNP.ipynb:164, AttnCNP.ipynb:151,
    # ConvCNP.ipynb:164, LNP.ipynb:149, AttnLNP.ipynb:148) + C3's UNet
    layouts = {}
    lpath = os.path.join(HERE, "layouts.json")
    if ONLY and os.path.exists(lpath):
        with open(lpath) as f:
            layouts = json.load(f)
    for nm in ["CNP", "AttnCNP", "ConvCNP", "ConvCNP_unet", "LNP", "AttnLNP", "ConvLNP"]:
        if ONLY and nm in layouts:
            continue
        m = build(nm, 128, nz=1)
        layouts[nm] = {"n_params": sum(p.numel() for p in m.parameters()),
                       "state_dict": {k: list(v.shape) for k, v in m.state_dict().items()}}
        print(nm, layouts[nm]["n_params"])
    with open(os.path.join(HERE, "layouts.json"), "w") as f:
        json.dump(layouts, f, indent=0, sort_keys=True)


if __name__ == "__main__":
    main()
