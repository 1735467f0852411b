"""Builders shared by the tests: the product models configured like the golden fixtures / notebook configs."""
import math
import warnings
from functools import partial

import torch

warnings.filterwarnings("ignore")


def build_product(name, R=128, nz=1, **over):
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.architectures import (CNN, MLP, ResConvBlock, SetConv, UnetCNN, UnsharedExpRBF,
                                                   discard_ith_arg, merge_flat_input)

    xenc = partial(MLP, n_hidden_layers=1, hidden_size=R)
    dec = merge_flat_input(partial(MLP, n_hidden_layers=4, hidden_size=R), is_sum_merge=True)
    xy = lambda h: merge_flat_input(partial(MLP, n_hidden_layers=2, hidden_size=h), is_sum_merge=True)  # noqa: E731
    if name == "CNP":
        return npf.CNP(x_dim=1, y_dim=1, XYEncoder=xy(2 * R), XEncoder=xenc, Decoder=dec, r_dim=R)
    if name == "AttnCNP":
        return npf.AttnCNP(x_dim=1, y_dim=1, XYEncoder=xy(R), is_self_attn=False, r_dim=R,
                           attention=over.get("attention", "transformer"), XEncoder=xenc, Decoder=dec)
    if name == "LNP":
        return npf.LNP(x_dim=1, y_dim=1, XYEncoder=xy(2 * R), is_q_zCct=True, n_z_samples_train=nz,
                       n_z_samples_test=nz, XEncoder=xenc, Decoder=dec, r_dim=R,
                       encoded_path=over.get("encoded_path", "latent"))
    if name == "AttnLNP":
        return npf.AttnLNP(x_dim=1, y_dim=1, XYEncoder=xy(R), is_self_attn=False, is_q_zCct=True,
                           n_z_samples_train=nz, n_z_samples_test=nz, r_dim=R, attention="transformer")
    if name == "ConvLNP":        # ConvLNP.ipynb:150-200 (1-D)
        cnn = partial(CNN, ConvBlock=ResConvBlock, is_chan_last=True, n_conv_layers=2, n_blocks=4, Conv=torch.nn.Conv1d,
                      Normalization=torch.nn.BatchNorm1d, kernel_size=over.get("kernel_size", 19))
        return npf.ConvLNP(x_dim=1, y_dim=1, Interpolator=SetConv, CNN=cnn, density_induced=over.get("density", 64),
                           is_global=over.get("is_global", True), is_q_zCct=over.get("is_q_zCct", False),
                           encoded_path=over.get("encoded_path", "latent"), n_z_samples_train=nz, n_z_samples_test=nz,
                           r_dim=R, Decoder=discard_ith_arg(torch.nn.Linear, i=0))
    if name in ("ConvCNP", "ConvCNP_unet", "ConvCNP_unshared"):
        kw = dict(ConvBlock=ResConvBlock, is_chan_last=True, n_conv_layers=2, Conv=torch.nn.Conv1d,
                  Normalization=torch.nn.BatchNorm1d, n_blocks=5, kernel_size=over.get("kernel_size", 19))
        if name == "ConvCNP_unet":
            cnn = partial(UnetCNN, Pool=torch.nn.AvgPool1d, upsample_mode="linear", max_nchannels=2 * R, **kw)
        else:
            cnn = partial(CNN, **kw)
        interp = partial(SetConv, RadialBasisFunc=UnsharedExpRBF) if name == "ConvCNP_unshared" else SetConv
        return npf.ConvCNP(x_dim=1, y_dim=1, Interpolator=interp, CNN=cnn, density_induced=over.get("density", 64),
                           r_dim=R, Decoder=discard_ith_arg(partial(MLP, n_hidden_layers=4, hidden_size=R), i=0))
    raise KeyError(name)


def perturb_(model, seed):
    """non-trivial biases, LN/BN affines and BN running statistics (same recipe as tests/golden/make_golden.py)"""
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


def chirp_batch(B, Nc, Nt, seed, device="cpu"):
    """Synthetic chirp tasks: targets on linspace(-1,1,Nt), context = random subset."""
    g = torch.Generator().manual_seed(seed)
    x = torch.linspace(-1, 1, Nt).view(1, Nt, 1).expand(B, Nt, 1).contiguous()
    tau = (x + 1) / 2
    f0 = 4 + 8 * torch.rand(B, 1, 1, generator=g)
    k = 20 + 60 * torch.rand(B, 1, 1, generator=g)
    ph = 2 * math.pi * torch.rand(B, 1, 1, generator=g)
    y = (0.1 + tau) ** 2 * torch.sin(2 * math.pi * (f0 * tau + 0.5 * k * tau ** 2) + ph)
    if Nt > 1:
        y = (y - y.mean(1, keepdim=True)) / y.std(1, keepdim=True)
    idx = torch.stack([torch.randperm(Nt, generator=g)[:Nc] for _ in range(B)]).unsqueeze(-1)
    out = dict(X_cntxt=x.gather(1, idx), Y_cntxt=y.gather(1, idx), X_trgt=x, Y_trgt=y)
    return {k_: v.to(device) for k_, v in out.items()}


def rel_err(a, b):
    """max |a-b| / max |b| : the normalised error all parity gates in tests/ use"""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
