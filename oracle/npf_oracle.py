"""CPU oracle for the npf forward / loss hot path.  TEST INFRASTRUCTURE ONLY.

This file is the checker, never the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  Nothing under
``npf_gwwaveform_b200/`` imports it.

It restates, as plain functional torch math on CPU tensors (fp32 or fp64), what
the reference ``npf`` modules compute on the 1-D hot path.  It is driven by a
reference-format ``state_dict`` (same key names as ``results/pretrained/**/
params.pt``) so weights can be shared verbatim with the reference and with the
CUDA implementation.  All ``file:line`` citations are relative to the
reference checkout (``/root/reference`` in the build container).

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the unmodified
reference in the build container, runs every model family below and stores
inputs / weights / outputs under ``tests/golden/*.pt``;
``tests/test_oracle_golden.py`` checks this oracle against those vectors.

Third-party arithmetic on the path is torch itself (softplus, softmax,
layer_norm, batch_norm, conv1d, interpolate, Normal.log_prob, kl_divergence;
unpinned in the reference's requirements.txt:2).  Their published formulas are
restated here with elementary ops so the oracle does not depend on
``torch.nn`` modules.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch

Tensor = torch.Tensor
SD = Dict[str, Tensor]

LOG_SQRT_2PI = 0.5 * math.log(2.0 * math.pi)


# --------------------------------------------------------------------------- #
# elementary pieces
# --------------------------------------------------------------------------- #
def softplus(x: Tensor) -> Tensor:
    """log(1+exp(x)) with torch's threshold-20 linearisation (F.softplus default)."""
    return torch.where(x > 20.0, x, torch.log1p(torch.exp(torch.clamp(x, max=20.0))))


def linear(sd: SD, prefix: str, x: Tensor) -> Tensor:
    """nn.Linear: x W^T + b (bias optional)."""
    y = x @ sd[prefix + ".weight"].to(x.dtype).t()
    b = sd.get(prefix + ".bias")
    if b is not None:
        y = y + b.to(x.dtype)
    return y


def mlp(sd: SD, prefix: str, x: Tensor) -> Tensor:
    """MLP.forward, ReLU activation, dropout=Identity, is_res=False.

    npf/architectures/mlp.py:95-109 : out(act(lin_n(...act(to_hidden(x))))).
    The number of hidden ``linears`` is read off the state_dict keys.
    """
    h = torch.relu(linear(sd, prefix + ".to_hidden", x))
    i = 0
    while f"{prefix}.linears.{i}.weight" in sd:
        h = torch.relu(linear(sd, f"{prefix}.linears.{i}", h))
        i += 1
    return linear(sd, prefix + ".out", h)


def merge_sum(sd: SD, prefix: str, x1: Tensor, x2: Tensor) -> Tensor:
    """MergeFlatInputs.forward with is_sum_merge=True.

    npf/architectures/encoders.py:175-183 : flat_module(relu(x1 + resizer(x2))).
    """
    return mlp(sd, prefix + ".flat_module", torch.relu(x1 + mlp(sd, prefix + ".resizer", x2)))


def layer_norm(sd: SD, prefix: str, x: Tensor, eps: float = 1e-5) -> Tensor:
    """nn.LayerNorm over the last dim (biased variance)."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * sd[prefix + ".weight"].to(x.dtype) + sd[prefix + ".bias"].to(x.dtype)


# --------------------------------------------------------------------------- #
# attention  (npf/architectures/attention.py)
# --------------------------------------------------------------------------- #
def multihead_attender(sd: SD, prefix: str, keys: Tensor, queries: Tensor, values: Tensor,
                       n_heads: int = 8) -> Tensor:
    """MultiheadAttender.forward, is_relative_pos=False.

    attention.py:457-505 : K = W_k k (no bias), Q = W_q q + b, V = W_v v (no bias);
    heads split on the feature dim (:507-527); DotAttender.score einsum / sqrt(head)
    (:204-220); softmax over keys, bmm with values (:129-164); optional
    post_processor Linear (:499-503).  value_head_size = kq_size // n_heads (:432).
    """
    K = linear(sd, prefix + ".key_transform", keys)
    Q = linear(sd, prefix + ".query_transform", queries)
    V = linear(sd, prefix + ".value_transform", values)
    B, Nk, kq = K.shape
    Nq = Q.shape[1]
    d = kq // n_heads
    Kh = K.view(B, Nk, n_heads, d).permute(0, 2, 1, 3)
    Qh = Q.view(B, Nq, n_heads, d).permute(0, 2, 1, 3)
    Vh = V.view(B, Nk, n_heads, d).permute(0, 2, 1, 3)
    logits = (Qh @ Kh.transpose(-1, -2)) / math.sqrt(d)
    attn = torch.softmax(logits, dim=-1)
    ctx = (attn @ Vh).permute(0, 2, 1, 3).reshape(B, Nq, n_heads * d)
    if prefix + ".post_processor.weight" in sd:
        ctx = linear(sd, prefix + ".post_processor", ctx)
    return ctx


def transformer_attender(sd: SD, prefix: str, keys: Tensor, queries: Tensor, values: Tensor,
                         n_heads: int = 8) -> Tensor:
    """TransformerAttender.forward (attention.py:569-588).

    Residual is added to the *raw* queries (:585); post_processor is None (:556).
    """
    ctx = multihead_attender(sd, prefix, keys, queries, values, n_heads)
    c = layer_norm(sd, prefix + ".layer_norm1", ctx + queries)
    return layer_norm(sd, prefix + ".layer_norm2", c + mlp(sd, prefix + ".mlp", c))


def dot_attender(keys: Tensor, queries: Tensor, values: Tensor) -> Tensor:
    """DotAttender(is_scale=True) when value_size == out_size (no resizer).

    attention.py:204-220,129-156.
    """
    logits = queries @ keys.transpose(-1, -2) / math.sqrt(queries.shape[-1])
    return torch.softmax(logits, dim=-1) @ values


# --------------------------------------------------------------------------- #
# SetConv  (npf/architectures/setcnn.py)
# --------------------------------------------------------------------------- #
def set_conv(sd: SD, prefix: str, keys: Tensor, queries: Tensor, values: Tensor, p: int = 2) -> Tensor:
    """SetConv.forward with ExpRBF (setcnn.py:234-268, 126-142).

    dist = |x_k - x_q| ; sigma = 1e-5 + softplus(theta) ; inp = -(dist/sigma)^p ;
    weight = softmax_k(inp) ; density = sum_k exp(inp) ;
    out = resizer(cat(sum_k weight*v_k, density)).
    """
    theta = sd[prefix + ".radial_basis_func.length_scale_param"].to(keys.dtype)
    if theta.numel() == 2:
        return set_conv_unshared(sd, prefix, keys, queries, values, p)
    sigma = 1e-5 + softplus(theta)
    diff = keys.unsqueeze(1) - queries.unsqueeze(2)                 # [B,Nq,Nk,1]
    dist = diff.abs().sum(-1, keepdim=True) if diff.shape[-1] == 1 else torch.norm(diff, p=p, dim=-1, keepdim=True)
    inp = -((dist / sigma) ** p)                                      # [B,Nq,Nk,1]
    weight = torch.softmax(inp, dim=-2)
    density = torch.exp(inp).sum(dim=-2)                              # [B,Nq,1]
    targets = (weight * values.unsqueeze(1)).sum(dim=2)               # [B,Nq,C]
    return linear(sd, prefix + ".resizer", torch.cat([targets, density], dim=-1))


def set_conv_unshared(sd: SD, prefix: str, keys: Tensor, queries: Tensor, values: Tensor, p: int = 2) -> Tensor:
    """SetConv with the fork's UnsharedExpRBF (setcnn.py:14-83).

    Two length scales: channel 0 weights, channel 1 density;
    weight = exp0 / (density + 1e-8) (explicit division, not softmax).
    """
    theta = sd[prefix + ".radial_basis_func.length_scale_param"].to(keys.dtype)
    sigma = 1e-5 + softplus(theta)                                    # [2]
    dist = (keys.unsqueeze(1) - queries.unsqueeze(2)).abs()           # [B,Nq,Nk,1]
    out = torch.exp(-((dist / sigma) ** p))                           # [B,Nq,Nk,2]
    density = out[..., 1:].sum(dim=-2)                                # [B,Nq,1]
    weight = out[..., 0:1] / (density.unsqueeze(2) + 1e-8)
    targets = (weight * values.unsqueeze(1)).sum(dim=2)
    return linear(sd, prefix + ".resizer", torch.cat([targets, density], dim=-1))


# --------------------------------------------------------------------------- #
# 1-D CNN / UNet  (npf/architectures/cnn.py) -- tensors are channel-LAST [B, L, C]
# --------------------------------------------------------------------------- #
def batch_norm(sd: SD, prefix: str, x: Tensor, training: bool, new_stats: Optional[dict],
               eps: float = 1e-5, momentum: float = 0.1) -> Tensor:
    """nn.BatchNorm1d on channel-last x[B,L,C].

    train: batch mean / biased var normalise; running stats updated with the
    unbiased var and momentum 0.1 (recorded into ``new_stats``).  eval: running stats.
    Absent keys (Normalization=nn.Identity) -> identity.
    """
    if prefix + ".weight" not in sd:
        return x
    w = sd[prefix + ".weight"].to(x.dtype)
    b = sd[prefix + ".bias"].to(x.dtype)
    if training:
        n = x.shape[0] * x.shape[1]
        mean = x.mean(dim=(0, 1))
        var = ((x - mean) ** 2).mean(dim=(0, 1))
        if new_stats is not None:
            # a module applied twice in one forward (ConvLNP encodes context and targets) updates its statistics twice
            rm = new_stats.get(prefix + ".running_mean", sd[prefix + ".running_mean"]).to(x.dtype)
            rv = new_stats.get(prefix + ".running_var", sd[prefix + ".running_var"]).to(x.dtype)
            nb = new_stats.get(prefix + ".num_batches_tracked", sd[prefix + ".num_batches_tracked"])
            new_stats[prefix + ".running_mean"] = ((1 - momentum) * rm + momentum * mean).detach()
            new_stats[prefix + ".running_var"] = ((1 - momentum) * rv + momentum * var * n / max(n - 1, 1)).detach()
            new_stats[prefix + ".num_batches_tracked"] = nb + 1
    else:
        mean = sd[prefix + ".running_mean"].to(x.dtype)
        var = sd[prefix + ".running_var"].to(x.dtype)
    return (x - mean) / torch.sqrt(var + eps) * w + b


def depthwise_conv(sd: SD, prefix: str, x: Tensor) -> Tensor:
    """Conv1d(C, C, k, padding=k//2, groups=C) on channel-last x[B,L,C] (zero padding)."""
    w = sd[prefix + ".weight"].to(x.dtype)              # [C,1,k]
    C, _, k = w.shape
    pad = k // 2
    B, L, _ = x.shape
    xp = torch.zeros(B, L + 2 * pad, C, dtype=x.dtype, device=x.device)
    xp[:, pad:pad + L] = x
    y = torch.zeros_like(x)
    for j in range(k):
        y = y + xp[:, j:j + L] * w[:, 0, j]
    b = sd.get(prefix + ".bias")
    if b is not None:
        y = y + b.to(x.dtype)
    return y


def pointwise_conv(sd: SD, prefix: str, x: Tensor) -> Tensor:
    """Conv1d(Cin, Cout, 1) on channel-last x == Linear over channels."""
    w = sd[prefix + ".weight"].to(x.dtype)[:, :, 0]     # [Cout,Cin]
    y = x @ w.t()
    b = sd.get(prefix + ".bias")
    if b is not None:
        y = y + b.to(x.dtype)
    return y


def res_conv_block(sd: SD, prefix: str, x: Tensor, training: bool, new_stats: Optional[dict]) -> Tensor:
    """ResConvBlock.forward (cnn.py:204-215), depth-separable convs (utils/helpers.py:395-398).

    n_conv_layers==2: out = pointwise1(depthwise1(relu(norm1(x)))) ;
    out = depthwise2(relu(norm2(out))) + x ; out = pointwise2(out).
    """
    out = x
    if prefix + ".conv1.depthwise.weight" in sd:
        out = torch.relu(batch_norm(sd, prefix + ".norm1", out, training, new_stats))
        out = pointwise_conv(sd, prefix + ".conv1.pointwise", depthwise_conv(sd, prefix + ".conv1.depthwise", out))
    out = torch.relu(batch_norm(sd, prefix + ".norm2", out, training, new_stats))
    out = depthwise_conv(sd, prefix + ".conv2_depthwise", out) + x
    return pointwise_conv(sd, prefix + ".conv2_pointwise", out)


def _n_blocks(sd: SD, prefix: str) -> int:
    n = 0
    while f"{prefix}.conv_blocks.{n}.conv2_pointwise.weight" in sd:
        n += 1
    return n


def cnn(sd: SD, prefix: str, x: Tensor, training: bool, new_stats: Optional[dict] = None) -> Tensor:
    """CNN.forward with is_chan_last=True (cnn.py:363-380): sequential blocks."""
    for i in range(_n_blocks(sd, prefix)):
        x = res_conv_block(sd, f"{prefix}.conv_blocks.{i}", x, training, new_stats)
    return x


def avg_pool2(x: Tensor) -> Tensor:
    """nn.AvgPool1d(2) on channel-last x[B,L,C] (floor(L/2) outputs)."""
    L2 = x.shape[1] // 2
    return 0.5 * (x[:, 0:2 * L2:2] + x[:, 1:2 * L2:2])


def upsample_linear2(x: Tensor) -> Tensor:
    """F.interpolate(mode='linear', scale_factor=2, align_corners=True) on channel-last x[B,L,C]."""
    L = x.shape[1]
    Lo = 2 * L
    if L == 1:
        return x.expand(-1, Lo, -1).clone()
    pos = torch.arange(Lo, dtype=x.dtype, device=x.device) * ((L - 1) / (Lo - 1))
    i0 = pos.floor().long().clamp(max=L - 1)
    i1 = (i0 + 1).clamp(max=L - 1)
    lam = (pos - i0.to(x.dtype)).view(1, Lo, 1)
    return x[:, i0] * (1 - lam) + x[:, i1] * lam


def unet_cnn(sd: SD, prefix: str, x: Tensor, training: bool, new_stats: Optional[dict] = None) -> Tensor:
    """UnetCNN.apply_convs (cnn.py:451-490) with Pool=AvgPool1d(2), linear upsampling,
    is_force_same_bottleneck=False; channel-last in/out."""
    n_blocks = _n_blocks(sd, prefix)
    n_down = n_blocks // 2
    residuals: List[Tensor] = []
    for i in range(n_down):
        x = res_conv_block(sd, f"{prefix}.conv_blocks.{i}", x, training, new_stats)
        residuals.append(x)
        x = avg_pool2(x)
    x = res_conv_block(sd, f"{prefix}.conv_blocks.{n_down}", x, training, new_stats)
    for i in range(n_down + 1, n_blocks):
        x = upsample_linear2(x)
        x = torch.cat((x, residuals[n_down - i]), dim=-1)
        x = res_conv_block(sd, f"{prefix}.conv_blocks.{i}", x, training, new_stats)
    return x


# --------------------------------------------------------------------------- #
# predictive distribution / losses
# --------------------------------------------------------------------------- #
def suffstat_to_loc_scale(suff: Tensor, y_dim: int = 1, heteroskedastic: bool = True) -> Tuple[Tensor, Tensor]:
    """NeuralProcessFamily.decode tail (base.py:347-362): split, scale = 0.01 + 0.99 softplus."""
    loc, s = suff.split(y_dim, dim=-1)
    scale = 0.01 + 0.99 * softplus(s)
    if not heteroskedastic:
        nz, B = scale.shape[:2]
        scale = scale.reshape(nz * B, -1, y_dim).mean(1, keepdim=True).view(nz, B, *([1] * (scale.dim() - 3)), y_dim).expand_as(loc)
    return loc, scale


def normal_log_prob(y: Tensor, loc: Tensor, scale: Tensor) -> Tensor:
    """torch.distributions.Normal.log_prob: -(y-mu)^2/(2 s^2) - log s - log sqrt(2 pi)."""
    return -((y - loc) ** 2) / (2 * scale ** 2) - torch.log(scale) - LOG_SQRT_2PI


def sum_log_prob(loc: Tensor, scale: Tensor, y: Tensor) -> Tensor:
    """losses.py:18-24 with Independent(Normal,1): sum over targets and y_dim -> [n_z, B]."""
    lp = normal_log_prob(y.unsqueeze(0), loc, scale)
    return lp.reshape(lp.shape[0], lp.shape[1], -1).sum(-1)


def kl_diag_normal(loc_q: Tensor, scale_q: Tensor, loc_p: Tensor, scale_p: Tensor) -> Tensor:
    """kl_divergence(Independent(Normal q), Independent(Normal p)), summed over the event dim.

    torch.distributions.kl._kl_normal_normal: 0.5*(var_ratio + t1 - 1 - log var_ratio).
    """
    var_ratio = (scale_q / scale_p) ** 2
    t1 = ((loc_q - loc_p) / scale_p) ** 2
    return (0.5 * (var_ratio + t1 - 1 - torch.log(var_ratio))).sum(-1)


def reduce_loss(loss: Tensor, reduction: Optional[str]) -> Tensor:
    """BaseLossNPF.forward tail (losses.py:71-81)."""
    if reduction is None:
        return loss
    if reduction == "mean":
        return loss.mean(0)
    if reduction == "sum":
        return loss.sum(0)
    raise ValueError(f"Unknown {reduction}")


def cnpf_loss(out: dict, y_trgt: Tensor, reduction: Optional[str] = "mean") -> Tensor:
    """CNPFLoss.get_loss (losses.py:112-123): -sum_t log p, n_z squeezed."""
    return reduce_loss(-sum_log_prob(out["loc"], out["scale"], y_trgt).squeeze(0), reduction)


def elbo_loss(out: dict, y_trgt: Tensor, reduction: Optional[str] = "mean") -> Tensor:
    """ELBOLossLNPF.get_loss (losses.py:126-150): -(mean_z sum_t log p - sum_lat KL(q_zCct || q_zCc))."""
    e = sum_log_prob(out["loc"], out["scale"], y_trgt).mean(0)
    kl = kl_diag_normal(out["q_zCct_loc"], out["q_zCct_scale"], out["q_zCc_loc"], out["q_zCc_scale"])
    kl = kl.reshape(kl.shape[0], -1).sum(-1)
    return reduce_loss(-(e - kl), reduction)


def nll_loss(out: dict, y_trgt: Tensor, reduction: Optional[str] = "mean", importance: bool = False) -> Tensor:
    """NLLLossLNPF.get_loss (losses.py:169-203): -(logsumexp_z sum_t log p [+ IS weights] - log n_z).

    This is also the eval-mode branch of every loss (losses.py:63-69) with q_zCct=None.
    """
    s = sum_log_prob(out["loc"], out["scale"], y_trgt)
    if importance:
        z = out["z"]

        def _slp(loc, scale):
            lp = normal_log_prob(z, loc.unsqueeze(0), scale.unsqueeze(0)).sum(-1)   # Independent over z_dim
            return lp.reshape(lp.shape[0], lp.shape[1], -1).sum(-1)

        s = s + _slp(out["q_zCc_loc"], out["q_zCc_scale"]) - _slp(out["q_zCct_loc"], out["q_zCct_scale"])
    n_z = s.shape[0]
    return reduce_loss(-(torch.logsumexp(s, 0) - math.log(n_z)), reduction)


# --------------------------------------------------------------------------- #
# model forwards  (npf/neuralproc/*.py)
# --------------------------------------------------------------------------- #
def _decode_merge(sd: SD, x_trgt_enc: Tensor, r_trgt: Tensor) -> Tuple[Tensor, Tensor]:
    """decode with Decoder = merge_flat_input(MLP, is_sum_merge=True) (base.py:327-367)."""
    return suffstat_to_loc_scale(merge_sum(sd, "decoder", x_trgt_enc, r_trgt))


def _latent_dist(sd: SD, r_lat: Tensor) -> Tuple[Tensor, Tensor]:
    """infer_latent_dist (base.py:516-547): latent_encoder MLP -> loc, 0.1 + 0.9 sigmoid."""
    suff = mlp(sd, "latent_encoder", r_lat)
    z_dim = suff.shape[-1] // 2
    loc, s = suff.split(z_dim, dim=-1)
    return loc, 0.1 + 0.9 * torch.sigmoid(s)


def _merge_r_z(sd: SD, r: Tensor, z: Tensor) -> Tensor:
    """merge_r_z (base.py:554-575): relu(Linear(cat(R expanded over n_z, z)))."""
    if r.shape != z.shape:
        r = r.unsqueeze(0).expand(*z.shape[:-1], r.shape[-1])
    return torch.relu(linear(sd, "r_z_merger", torch.cat((r, z), dim=-1)))


def _check_range(training: bool, x_c: Tensor, x_t: Tensor) -> None:
    """_validate_inputs (base.py:241-247): ValueError outside [-1,1] in training mode only."""
    if training and not (bool(((x_c >= -1) & (x_c <= 1)).all()) and bool(((x_t >= -1) & (x_t <= 1)).all())):
        raise ValueError("Features during training should be in [-1,1].")


def cnp_forward(sd: SD, x_c: Tensor, y_c: Tensor, x_t: Tensor, training: bool = False) -> dict:
    """CNP (np.py:86-110) through NeuralProcessFamily.forward (base.py:177-239)."""
    _check_range(training, x_c, x_t)
    xc, xt = mlp(sd, "x_encoder", x_c), mlp(sd, "x_encoder", x_t)
    B, Nc, _ = xc.shape
    r_dim = sd["decoder.resizer.to_hidden.weight"].shape[1]
    if Nc == 0:
        R = torch.zeros(B, 1, r_dim, dtype=xt.dtype, device=xt.device)
    else:
        R = merge_sum(sd, "xy_encoder", xc, y_c).mean(dim=1, keepdim=True)
    r_trgt = R.expand(B, xt.shape[1], r_dim).unsqueeze(0)
    loc, scale = _decode_merge(sd, xt, r_trgt)
    return {"loc": loc, "scale": scale}


def attncnp_forward(sd: SD, x_c: Tensor, y_c: Tensor, x_t: Tensor, training: bool = False,
                    attention: str = "transformer", n_heads: int = 8) -> dict:
    """AttnCNP (attnnp.py:105-131), is_self_attn=False."""
    _check_range(training, x_c, x_t)
    xc, xt = mlp(sd, "x_encoder", x_c), mlp(sd, "x_encoder", x_t)
    r_trgt = _attn_det_path(sd, xc, y_c, xt, attention, n_heads)[1]
    loc, scale = _decode_merge(sd, xt, r_trgt.unsqueeze(0))
    return {"loc": loc, "scale": scale}


def _attn_det_path(sd, xc, y_c, xt, attention, n_heads):
    B, Nc, _ = xc.shape
    r_dim = sd["decoder.resizer.to_hidden.weight"].shape[1]
    if Nc == 0:
        return (torch.zeros(B, 0, r_dim, dtype=xt.dtype, device=xt.device),
                torch.zeros(B, xt.shape[1], r_dim, dtype=xt.dtype, device=xt.device))
    R = merge_sum(sd, "xy_encoder", xc, y_c)
    if attention == "transformer":
        r_trgt = transformer_attender(sd, "attender", xc, xt, R, n_heads)
    elif attention == "multihead":
        r_trgt = multihead_attender(sd, "attender", xc, xt, R, n_heads)
    elif attention == "scaledot":
        r_trgt = dot_attender(xc, xt, R)
    else:
        raise ValueError(f"Unknown attention method {attention}")
    return R, r_trgt


def convcnp_forward(sd: SD, x_c: Tensor, y_c: Tensor, x_t: Tensor, training: bool = False,
                    density_induced: int = 64, unet: bool = False, x_induced: Optional[Tensor] = None,
                    new_stats: Optional[dict] = None) -> dict:
    """ConvCNP (convnp.py:104,128-168): grid linspace(-1.5,1.5,3*density); SetConv -> CNN -> SetConv;
    decoder = DiscardIthArg(MLP) (encoders.py:105-120)."""
    _check_range(training, x_c, x_t)
    B, Nc, _ = x_c.shape
    if x_induced is None:
        x_induced = torch.linspace(-1.5, 1.5, int(density_induced * 3), device=x_c.device)
    grid = x_induced.to(x_c.dtype).view(1, -1, 1).expand(B, -1, 1)
    r_dim = sd["cntxt_to_induced.resizer.weight"].shape[0]
    if Nc == 0:
        R = torch.zeros(B, grid.shape[1], r_dim, dtype=x_c.dtype, device=x_c.device)
    else:
        R = set_conv(sd, "cntxt_to_induced", x_c, grid, y_c)
    R = (unet_cnn if unet else cnn)(sd, "induced_to_induced", R, training, new_stats)
    r_trgt = set_conv(sd, "induced_to_trgt", grid, x_t, R).unsqueeze(0)
    loc, scale = suffstat_to_loc_scale(mlp(sd, "decoder.destination", r_trgt))
    return {"loc": loc, "scale": scale}


def _latent_path(sd, out, r_lat_c, r_lat_ct, n_z, eps):
    """latent_path (base.py:495-514): q_zCc, optional q_zCct, z = loc + scale*eps (rsample, G4)."""
    out["q_zCc_loc"], out["q_zCc_scale"] = _latent_dist(sd, r_lat_c)
    if r_lat_ct is not None:
        out["q_zCct_loc"], out["q_zCct_scale"] = _latent_dist(sd, r_lat_ct)
        loc, scale = out["q_zCct_loc"], out["q_zCct_scale"]
    else:
        loc, scale = out["q_zCc_loc"], out["q_zCc_scale"]
    if eps is None:
        eps = torch.randn(n_z, *loc.shape, dtype=loc.dtype)
    out["z"] = loc.unsqueeze(0) + scale.unsqueeze(0) * eps.to(loc.dtype)
    return out["z"]


def lnp_forward(sd: SD, x_c: Tensor, y_c: Tensor, x_t: Tensor, y_t: Optional[Tensor] = None,
                training: bool = False, n_z: int = 1, eps: Optional[Tensor] = None,
                encoded_path: str = "latent", is_q_zCct: bool = True) -> dict:
    """LNP (np.py:113-163 + base.py:374-575), z_dim == r_dim."""
    _check_range(training, x_c, x_t)
    xc, xt = mlp(sd, "x_encoder", x_c), mlp(sd, "x_encoder", x_t)
    B, Nc, _ = xc.shape
    r_dim = sd["decoder.resizer.to_hidden.weight"].shape[1]

    def enc(xe, y):
        if xe.shape[1] == 0:
            return torch.zeros(B, 1, r_dim, dtype=xt.dtype, device=xt.device)
        return merge_sum(sd, "xy_encoder", xe, y).mean(dim=1, keepdim=True)

    R = enc(xc, y_c)
    out: dict = {}
    r_ct = enc(xt, y_t) if (is_q_zCct and y_t is not None) else None
    z = _latent_path(sd, out, R, r_ct, n_z, eps)                        # [n_z,B,1,r]
    r_trgt = _merge_r_z(sd, R, z) if encoded_path == "both" else z
    r_trgt = r_trgt.expand(z.shape[0], B, xt.shape[1], r_dim)
    out["loc"], out["scale"] = _decode_merge(sd, xt, r_trgt)
    return out


def attnlnp_forward(sd: SD, x_c: Tensor, y_c: Tensor, x_t: Tensor, y_t: Optional[Tensor] = None,
                    training: bool = False, n_z: int = 1, eps: Optional[Tensor] = None,
                    attention: str = "transformer", n_heads: int = 8, is_q_zCct: bool = True) -> dict:
    """AttnLNP (attnnp.py:134-202), encoded_path='both'."""
    _check_range(training, x_c, x_t)
    xc, xt = mlp(sd, "x_encoder", x_c), mlp(sd, "x_encoder", x_t)
    B = xc.shape[0]
    r_dim = sd["decoder.resizer.to_hidden.weight"].shape[1]
    R, r_det = _attn_det_path(sd, xc, y_c, xt, attention, n_heads)

    def pool(Rm):   # rep_to_lat_input (attnnp.py:172-181)
        if Rm.shape[1] == 0:
            Rm = torch.zeros(B, 1, r_dim, dtype=xt.dtype, device=xt.device)
        return Rm.mean(dim=1, keepdim=True)

    out: dict = {}
    r_ct = None
    if is_q_zCct and y_t is not None:
        r_ct = pool(merge_sum(sd, "xy_encoder", xt, y_t))
    z = _latent_path(sd, out, pool(R), r_ct, n_z, eps)                  # [n_z,B,1,r]
    zt = z.expand(z.shape[0], B, xt.shape[1], r_dim)
    r_trgt = _merge_r_z(sd, r_det, zt)
    out["loc"], out["scale"] = _decode_merge(sd, xt, r_trgt)
    return out


def convlnp_forward(sd: SD, x_c: Tensor, y_c: Tensor, x_t: Tensor, y_t: Optional[Tensor] = None,
                    training: bool = False, n_z: int = 1, eps: Optional[Tensor] = None,
                    density_induced: int = 64, encoded_path: str = "latent", is_global: bool = False,
                    is_q_zCct: bool = True, new_stats: Optional[dict] = None) -> dict:
    """ConvLNP (convnp.py:184-335 + base.py:374-575): latent per induced point, CNN after sampling, optional
    global latent (neuralproc/helpers.py:21-34), SetConv on the n_z*B batch, decoder = DiscardIthArg(nn.Linear)."""
    _check_range(training, x_c, x_t)
    B, Nc, _ = x_c.shape
    grid1 = torch.linspace(-1.5, 1.5, int(density_induced * 3), device=x_c.device).to(x_c.dtype)
    grid = grid1.view(1, -1, 1).expand(B, -1, 1)
    U = grid.shape[1]
    r_dim = sd["cntxt_to_induced.resizer.weight"].shape[0]

    def encode(x, y):        # ConvCNP.encode_globally (convnp.py:128-154)
        if x.shape[1] == 0:
            R0 = torch.zeros(B, U, r_dim, dtype=x_t.dtype, device=x_t.device)
        else:
            R0 = set_conv(sd, "cntxt_to_induced", x, grid, y)
        return cnn(sd, "induced_to_induced", R0, training, new_stats)

    def lat_in(R):           # rep_to_lat_input (convnp.py:253-263)
        return R if encoded_path == "latent" else R.reshape(B, -1, r_dim).mean(dim=1, keepdim=True)

    R = encode(x_c, y_c)
    out: dict = {}
    r_ct = lat_in(encode(x_t, y_t)) if (is_q_zCct and y_t is not None) else None
    z = _latent_path(sd, out, lat_in(R), r_ct, n_z, eps)             # [n_z, B, U or 1, z]
    nz = z.shape[0]
    gridz = grid1.view(1, -1, 1).expand(nz * B, -1, 1)
    xt_rep = x_t.unsqueeze(0).expand(nz, *x_t.shape).reshape(nz * B, x_t.shape[1], 1)
    if encoded_path == "latent":
        h = z.reshape(nz * B, U, -1)
        if "reshaper_z.weight" in sd:
            h = linear(sd, "reshaper_z", h)
        h = cnn(sd, "induced_to_induced_post_sampling", h, training, new_stats)
        if is_global:
            half = h.shape[-1] // 2
            local, glob = h.split(half, dim=-1)
            h = torch.cat([local, glob.mean(dim=1, keepdim=True).expand(-1, U, -1)], dim=-1)
    else:
        h = _merge_r_z(sd, R, z.expand(nz, B, U, z.shape[-1])).reshape(nz * B, U, r_dim)
        h = cnn(sd, "induced_to_induced_post_sampling", h, training, new_stats)
    r_trgt = set_conv(sd, "induced_to_trgt", gridz, xt_rep, h).view(nz, B, x_t.shape[1], r_dim)
    out["loc"], out["scale"] = suffstat_to_loc_scale(linear(sd, "decoder.destination", r_trgt))
    return out


FORWARDS = {
    "CNP": cnp_forward,
    "AttnCNP": attncnp_forward,
    "ConvCNP": convcnp_forward,
    "LNP": lnp_forward,
    "AttnLNP": attnlnp_forward,
    "ConvLNP": convlnp_forward,
}


# --------------------------------------------------------------------------- #
# context / target split  (npf/utils/datasplit.py)
# --------------------------------------------------------------------------- #
def select_points(X: Tensor, y: Tensor, indcs: Tensor) -> Tuple[Tensor, Tensor]:
    """CntxtTrgtGetter.select (datasplit.py:246-255): gather rows of X and y by per-task indices."""
    B = X.shape[0]
    ix = indcs.unsqueeze(-1).expand(B, -1, X.shape[-1])
    iy = indcs.unsqueeze(-1).expand(B, -1, y.shape[-1])
    return torch.gather(X, 1, ix).contiguous(), torch.gather(y, 1, iy).contiguous()


def add_cntxts_to_trgts(num_points: int, target_indcs: Tensor, context_indcs: Tensor) -> Tensor:
    """CntxtTrgtGetter.add_cntxts_to_trgts (datasplit.py:229-237)."""
    return torch.cat([target_indcs, context_indcs], dim=-1)[:, :num_points]


# --------------------------------------------------------------------------- #
# autograd helper: loss + parameter gradients through the oracle
# --------------------------------------------------------------------------- #
def loss_and_grads(forward, loss_fn, sd: SD, inputs: dict, y_trgt: Tensor, dtype=torch.float32, **kw):
    """Run ``forward`` with leaf copies of every floating-point tensor of ``sd`` requiring grad;
    return (out, loss, {name: grad})."""
    leaf = {}
    for k, v in sd.items():
        if v.is_floating_point() and "running_" not in k:
            leaf[k] = v.detach().to(dtype).clone().requires_grad_(True)
        else:
            leaf[k] = v.detach().clone()
    out = forward(leaf, **{k: (v.to(dtype) if torch.is_tensor(v) and v.is_floating_point() else v)
                           for k, v in inputs.items()}, **kw)
    loss = loss_fn(out, y_trgt.to(dtype))
    loss.backward()
    grads = {k: v.grad for k, v in leaf.items() if v.requires_grad and v.grad is not None}
    return out, loss.detach(), grads
