"""K1: the fused MLP-chain kernel (npfk_mlp_chain_fwd, csrc/tc_mlp_chain.cu) and the decoder tail
(128 -> 2 head + predictive-distribution transform + Gaussian log-likelihood + sum over targets in the chain's epilogue,
npfk_nll_head_bwd for its backward).

References: (i) the same chain run layer by layer through npfk_linear_fwd (`set_chain_fusion(False)`): the fused kernel
issues the same tcgen05 k-steps in the same order, so the 128-wide activations must agree BIT FOR BIT; (ii) float64
torch for the head, the log-likelihood and every gradient.  Mirrors npf/architectures/mlp.py:95-109,
npf/architectures/encoders.py:175-183, npf/neuralproc/base.py:347-353,116 and npf/losses.py:18-24.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fast_mode():
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200 import kernels as Kn

    npf.set_mode("fast")
    Kn.set_chain_fusion(True)
    yield
    Kn.set_chain_fusion(True)
    npf.set_mode("fp32")
    torch.cuda.synchronize()
    from npf_gwwaveform_b200._lib import LIB

    assert LIB.load().npfk_tc_abort_code() == 0, "a tcgen05 kernel gave up on an mbarrier wait"


def _layers(n, seed, n_out=128, dev="cuda"):
    g = torch.Generator().manual_seed(seed)
    out = []
    for i in range(n):
        N = n_out if i == n - 1 else 128
        w = (torch.randn(N, 128, generator=g) / math.sqrt(128) * 1.4).to(dev).requires_grad_(True)
        b = (0.1 * torch.randn(N, generator=g)).to(dev).requires_grad_(True)
        out.append((w, b))
    return out


def _ref_chain(x, layers, res, out_relu, dtype=torch.float64):
    """float64 restatement with the operands the kernels see: weights rounded to tf32 (nearest), activations truncated"""
    def tf32_rn(t):
        i = t.float().contiguous().view(torch.int32)
        return ((i + 0x1000) & ~0x1FFF).view(torch.float32).to(dtype)

    h = x.cpu().to(dtype)
    for i, (w, b) in enumerate(layers):
        last = i == len(layers) - 1
        h = h @ tf32_rn(w.detach().cpu()).T + b.detach().cpu().to(dtype)
        if last and res is not None:
            h = h + res.cpu().to(dtype).repeat(h.shape[0] // res.shape[0], 1)
        if not last or out_relu:
            h = torch.relu(h)
    return h


@pytest.mark.parametrize("M,L,with_res,out_relu", [
    (128, 2, False, False), (256, 2, True, True), (1000, 3, False, True), (128 * 7 + 5, 4, True, True),
    (128 * 301, 6, False, False), (4096, 8, True, False), (129, 2, False, True),
])
def test_chain_matches_layerwise_bitwise(M, L, with_res, out_relu):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(M + L)
    x = torch.randn(M, 128, generator=g).cuda()
    res = torch.randn(M, 128, generator=g).cuda() if with_res else None
    layers = _layers(L, 3 * L)
    for w, _ in layers:
        Kn.tf32_weight(w)        # rounded copies are cached: keep their launches out of the counts below
    for grad in (True, False):
        outs = []
        for fused in (True, False):
            Kn.set_chain_fusion(fused)
            n0 = Kn.launch_count()
            with torch.set_grad_enabled(grad):
                y = Kn.mlp_chain(x.clone().requires_grad_(grad), layers, res=res, out_relu=out_relu)
            outs.append((y.detach().clone(), Kn.launch_count() - n0))
        assert outs[0][1] == 1 and outs[1][1] == L, (outs[0][1], outs[1][1])      # one launch vs one per layer
        assert torch.equal(outs[0][0], outs[1][0]), float((outs[0][0] - outs[1][0]).abs().max())
    ref = _ref_chain(x.cpu(), layers, None if res is None else res.cpu(), out_relu)
    err = float((outs[0][0].cpu().double() - ref.cpu()).abs().max() / ref.abs().max())
    assert err < 2e-3, err      # activations are truncated to tf32 by the MMA: 2^-11 per operand


def test_chain_broadcast_residual_and_gradients():
    """res rows tiled over z-samples (LNP decoders), gradients of every input against layer-by-layer launches"""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(11)
    Z, R = 3, 256
    x = torch.randn(Z * R, 128, generator=g).cuda()
    res = torch.randn(R, 128, generator=g).cuda()
    dy = torch.randn(Z * R, 128, generator=g).cuda()
    got = []
    for fused in (True, False):
        Kn.set_chain_fusion(fused)
        layers = _layers(3, 5)
        xs, rs = x.clone().requires_grad_(True), res.clone().requires_grad_(True)
        y = Kn.mlp_chain(xs, layers, res=rs, out_relu=True)
        y.backward(dy)
        got.append([y.detach(), xs.grad, rs.grad] + [t.grad for wb in layers for t in wb])
    for a, b in zip(*got):
        assert torch.equal(a, b) or float((a - b).abs().max() / b.abs().max()) < 1e-6     # backward kernels are the same


def _tail_reference(x, layers, y, T):
    h = x.double()
    ws = [(w.detach().double().cpu().requires_grad_(True), b.detach().double().cpu().requires_grad_(True)) for w, b in layers]
    hh = h.cpu().requires_grad_(True)
    z = hh
    for i, (w, b) in enumerate(ws):
        z = z @ w.T + b
        if i < len(ws) - 1:
            z = torch.relu(z)
    loc, raw = z[:, 0], z[:, 1]
    scale = 0.01 + 0.99 * torch.nn.functional.softplus(raw)
    yy = y.double().cpu().reshape(-1)
    yy = yy.repeat(z.shape[0] // yy.numel())
    lp = -((yy - loc) ** 2) / (2 * scale ** 2) - scale.log() - 0.5 * math.log(2 * math.pi)
    return z, lp.view(-1, T).sum(1), hh, ws


@pytest.mark.parametrize("Z,B,T,L", [(1, 2, 128, 2), (1, 3, 200, 5), (4, 2, 96, 3), (2, 1, 4096, 5)])
def test_decoder_tail_head_and_loglik(Z, B, T, L):
    """flat MLP 128 -> [128] x (L-1) -> 2 with the head and sum_t log N(y; loc, scale) in the chain's epilogue"""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(Z * 100 + T)
    M = Z * B * T
    x = (torch.randn(M, 128, generator=g).relu()).cuda()
    y = torch.randn(B, T, 1, generator=g).cuda()
    layers = _layers(L, 17, n_out=2)
    for w, _ in layers:
        Kn.tf32_weight(w)
    xs = x.clone().requires_grad_(True)
    n0 = Kn.launch_count()
    suff, lp = Kn.mlp_chain(xs, layers, x_is_relu=True, nll=(y, T))
    assert Kn.launch_count() - n0 == 1
    assert lp is not None and lp.shape == (Z * B,)
    wsum = torch.randn(Z * B, generator=g).cuda()
    ((lp * wsum).sum() + 0.3 * suff.sum()).backward()
    fused = [xs.grad.clone()] + [t.grad.clone() for wb in layers for t in wb]
    rel = lambda a, b: float((a.detach().cpu().double() - b.detach().cpu().double()).abs().max()  # noqa: E731
                             / b.detach().abs().max())
    # forward against float64 (tf32 operands: 2^-11 per product term)
    ref_suff, ref_lp, _, _ = _tail_reference(x, layers, y, T)
    assert rel(suff, ref_suff) < 3e-3
    assert rel(lp, ref_lp) < 2e-3
    # inference: same numbers without the stashes
    with torch.no_grad():
        suff2, lp2 = Kn.mlp_chain(x, layers, x_is_relu=True, nll=(y, T))
    assert torch.equal(suff2, suff.detach()) and rel(lp2, lp) < 1e-5
    # the unfused tail (separate GEMM launches, skinny head, loc/scale + log-likelihood kernels and their backward kernels;
    # gated against float64 in tests/test_gpu_fast.py and test_gpu_kernels.py) gives the same outputs and the same
    # gradients: same hidden activations bit for bit, hence the same ReLU masks
    Kn.set_chain_fusion(False)
    for wb in layers:
        for t in wb:
            t.grad = None
    xs3 = x.clone().requires_grad_(True)
    suff3, lp3 = Kn.mlp_chain(xs3, layers, x_is_relu=True, nll=(y, T))
    assert lp3 is None
    lp3 = Kn.gauss_sum_log_prob(suff3.view(Z, B, T, 2), y).reshape(-1)
    assert rel(suff3, suff) < 1e-5 and rel(lp3, lp) < 1e-5
    ((lp3 * wsum).sum() + 0.3 * suff3.sum()).backward()
    unfused = [xs3.grad] + [t.grad for wb in layers for t in wb]
    for a, b in zip(fused, unfused):
        assert rel(a, b) < 2e-4, (a.shape, rel(a, b))
