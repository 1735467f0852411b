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


# ------------------------------------------------------------------------------------------ K1b: two layers per backward launch
@pytest.mark.parametrize("M,L,x_is_relu,with_res", [
    (128, 2, True, False), (300, 2, False, True), (128 * 148 + 77, 3, True, False), (1000, 4, False, True),
    (4096, 5, True, False), (256, 8, False, False),
])
def test_chain_backward_pairs_match_layerwise(M, L, x_is_relu, with_res):
    """npfk_linear_bwd_fused2 (K1b, csrc/tc_chain_bwd.cu) against one npfk_linear_bwd_fused launch per layer: both issue the
    same tcgen05 k-steps on the same operands (the gradient between the two layers is rounded to tf32 in either path), so
    dx agrees bit for bit; the weight / bias gradients are sums over CTAs in a different order.  Autograd of
    mlp.py:95-109."""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(7 * M + L)
    x = torch.randn(M, 128, generator=g)
    x = (torch.relu(x) if x_is_relu else x).cuda()
    res = torch.randn(M, 128, generator=g).cuda() if with_res else None
    dy = torch.randn(M, 128, generator=g).cuda()
    got, launches = [], []
    for paired in (True, False):
        Kn.set_chain_bwd_fusion(paired)
        try:
            layers = _layers(L, 5 + L)
            for w, _ in layers:
                Kn.tf32_weight(w)
            xs = x.clone().requires_grad_(True)
            rs = res.clone().requires_grad_(True) if with_res else None
            y = Kn.mlp_chain(xs, layers, res=rs, out_relu=with_res, x_is_relu=x_is_relu)
            n0 = Kn.launch_count()
            y.backward(dy)
            launches.append(Kn.launch_count() - n0)
        finally:
            Kn.set_chain_bwd_fusion(True)
        got.append([xs.grad] + [t.grad for wb in layers for t in wb] + ([rs.grad] if with_res else []))
    # L layers: floor(L / 2) pair launches + (L % 2) single ones, against L (+ the ReLU backward of the output)
    assert launches[1] - launches[0] == L // 2, launches
    assert torch.equal(got[0][0], got[1][0]), float((got[0][0] - got[1][0]).abs().max())
    for a, b in zip(got[0][1:], got[1][1:]):
        assert float((a - b).abs().max() / b.abs().max()) < 1e-5


@pytest.mark.parametrize("M,masked0,with_db,accumulate", [
    (128, True, True, False), (300, False, True, False), (128 * 148 * 2 + 77, True, False, False), (4096, True, True, True),
])
def test_linear_bwd_fused2_c_abi_against_float64(M, masked0, with_db, accumulate):
    """the entry point itself, raw pointers: every output against the float64 definition in include/npfk.h"""
    import ctypes

    from npf_gwwaveform_b200 import kernels as Kn
    from npf_gwwaveform_b200._lib import LIB

    def ptr(t):
        return ctypes.c_void_p(t.data_ptr()) if t is not None else None

    g = torch.Generator().manual_seed(M)
    G = torch.randn(M, 128, generator=g).cuda()
    H1 = torch.relu(torch.randn(M, 128, generator=g)).cuda()
    H0 = torch.relu(torch.randn(M, 128, generator=g)).cuda()
    W1 = Kn.tf32_weight(torch.randn(128, 128, generator=g).cuda() / 11)
    W0 = Kn.tf32_weight(torch.randn(128, 128, generator=g).cuda() / 11)
    dx0 = torch.randn(M, 128, generator=g).cuda() if accumulate else None
    dx = dx0.clone() if accumulate else torch.empty(M, 128, device="cuda")
    dw1, dw0 = torch.zeros(128, 128, device="cuda"), torch.zeros(128, 128, device="cuda")
    db1 = torch.zeros(128, device="cuda") if with_db else None
    db0 = torch.zeros(128, device="cuda") if with_db else None
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    LIB.call("npfk_linear_bwd_fused2", ptr(G), 128, ptr(H1), 128, ptr(W1), 128, ptr(H0), 128, ptr(W0), 128,
             int(masked0) | (2 if accumulate else 0), ptr(dx), 128, ptr(dw1), 128, ptr(db1), ptr(dw0), 128, ptr(db0), M, st)
    torch.cuda.synchronize()
    Gd, H1d, H0d, W1d, W0d = (t.double().cpu() for t in (G, H1, H0, W1, W0))
    gm = (Gd @ W1d) * (H1d > 0)
    ref_dx = gm @ W0d
    if masked0:
        ref_dx = ref_dx * (H0d > 0)
    if accumulate:
        ref_dx = ref_dx + dx0.double().cpu()
    checks = [(dx, ref_dx), (dw1, Gd.T @ H1d), (dw0, gm.T @ H0d)]
    if with_db:
        checks += [(db1, Gd.sum(0)), (db0, gm.sum(0))]
    for got, ref in checks:
        err = float((got.double().cpu() - ref).abs().max() / ref.abs().max())
        assert err < 3e-3, err      # tf32 operands (2^-11 each), fp32 accumulate
