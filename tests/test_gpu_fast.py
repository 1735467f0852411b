"""NPFK_MODE_FAST (tcgen05 tf32 tensor-core kernels) against float64 references.

Tolerances: tf32 operands carry a 10-bit mantissa (unit round-off 2^-11 = 4.9e-4), accumulation is fp32 in
TMEM, so a K-term dot product is good to about 1e-3 of the output scale; the gates below are normalised max
errors  max|a-b| / max|b|  <= 2e-3 for single kernels (BASELINE.json north_star: <=1e-3 relative on the
per-target log-likelihood in bf16/TF32 mode is checked on the models in test_gpu_models_fast).
"""
import math

import pytest
import torch

from util_models import rel_err

pytestmark = pytest.mark.gpu
TOL = 2e-3


@pytest.fixture(autouse=True)
def fast_mode():
    from npf_gwwaveform_b200 import kernels as Kn

    Kn.set_mode("fast")
    yield
    Kn.set_mode("fp32")
    torch.cuda.synchronize()
    from npf_gwwaveform_b200._lib import LIB

    assert LIB.load().npfk_tc_abort_code() == 0, "a tcgen05 kernel gave up on an mbarrier wait"


@pytest.mark.parametrize("M,N,K,relu,res_mode", [
    (128, 128, 128, False, "none"), (300, 128, 128, True, "none"), (4096, 128, 32, True, "full"),
    (1000, 256, 128, False, "full"), (3 * 512, 128, 128, True, "tiled"), (777, 64, 256, True, "none"),
    (5000, 512, 256, False, "none"), (2049, 16, 128, False, "none"), (640, 128, 48, True, "none"),
    (100, 128, 128, True, "none"),      # M < 128: stays on the FFMA kernel
    (70000, 128, 128, True, "none"),    # more tiles than CTAs: persistent loop + TMEM double buffering
])
def test_linear_fast(M, N, K, relu, res_mode):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(M + N + K)
    x, w, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / math.sqrt(K), torch.randn(N, generator=g)
    res = None
    if res_mode == "full":
        res = torch.randn(M, N, generator=g)
    elif res_mode == "tiled":
        res = torch.randn(M // 3, N, generator=g)
    dy = torch.randn(M, N, generator=g)

    xs, ws, bs = (t.double().requires_grad_(True) for t in (x, w, b))
    rs = res.double().requires_grad_(True) if res is not None else None
    y_ref = xs @ ws.t() + bs
    if rs is not None:
        y_ref = y_ref + (rs if res_mode == "full" else rs.repeat(3, 1))
    xd, wd, bd = (t.cuda().requires_grad_(True) for t in (x, w, b))
    rd = res.cuda().requires_grad_(True) if res is not None else None
    y = Kn.linear(xd, wd, bd, res=rd, relu=relu)
    y.backward(dy.cuda())
    if relu:
        # pre-activations within tf32 noise of zero may land on the other side of the ReLU: the reference
        # gradient is taken with the kernel's own activation pattern
        y_ref = y_ref * (y.detach().cpu() > 0)
    y_ref.backward(dy.double())
    assert rel_err(y, y_ref) < TOL, ("y", rel_err(y, y_ref))
    for n, a, r in (("dx", xd.grad, xs.grad), ("dw", wd.grad, ws.grad), ("db", bd.grad, bs.grad)):
        assert rel_err(a, r) < TOL, (n, rel_err(a, r))
    if rd is not None:
        assert rel_err(rd.grad, rs.grad) < TOL


@pytest.mark.parametrize("M,x_is_relu", [(128, False), (333, True), (20000, True), (148 * 128 * 2 + 77, False)])
def test_mlp_chain_fast_fused_backward(M, x_is_relu):
    """3-layer 128-wide ReLU chain: every layer's backward runs npfk_linear_bwd_fused (dgrad with the ReLU mask
    taken from the saved input + wgrad + bias gradient in one tcgen05 kernel); float64 reference uses the
    kernel's own activation pattern."""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(M)
    x = torch.randn(M, 128, generator=g)
    if x_is_relu:
        x = torch.relu(x)
    ws = [torch.randn(128, 128, generator=g) / math.sqrt(128) for _ in range(3)]
    bs = [torch.randn(128, generator=g) * 0.1 for _ in range(3)]
    dy = torch.randn(M, 128, generator=g)
    xd = x.cuda().requires_grad_(True)
    wd = [w.cuda().requires_grad_(True) for w in ws]
    bd = [b.cuda().requires_grad_(True) for b in bs]
    y = Kn.mlp_chain(xd, list(zip(wd, bd)), x_is_relu=x_is_relu)
    y.backward(dy.cuda())

    # activation pattern of the kernels (pre-activations within tf32 noise of zero may land on either side of the
    # ReLU, and a flipped unit changes the gradient by O(1)): the same layers through Kn.linear expose h1, h2
    with torch.no_grad():
        h1 = Kn.linear(xd, wd[0], bd[0], relu=True)
        h2 = Kn.linear(h1, wd[1], bd[1], relu=True)
    masks = [(h1 > 0).double().cpu(), (h2 > 0).double().cpu()]
    xs = x.double().requires_grad_(True)
    wss = [w.double().requires_grad_(True) for w in ws]
    bss = [b.double().requires_grad_(True) for b in bs]
    h = xs
    for i in range(3):
        h = h @ wss[i].t() + bss[i]
        if i < 2:
            h = h * masks[i]
    h.backward(dy.double())
    assert rel_err(y, h) < TOL
    gx = xs.grad * (x > 0).double() if x_is_relu else xs.grad
    assert rel_err(xd.grad, gx) < 2 * TOL, ("dx", rel_err(xd.grad, gx))
    for i in range(3):
        assert rel_err(wd[i].grad, wss[i].grad) < 2 * TOL, (f"dw{i}", rel_err(wd[i].grad, wss[i].grad))
        assert rel_err(bd[i].grad, bss[i].grad) < 2 * TOL, (f"db{i}", rel_err(bd[i].grad, bss[i].grad))


def test_linear_fast_strided():
    """operands that are column slices of wider tensors (split r_z_merger, SetConv output with a density column)"""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(5)
    W = torch.randn(128, 256, generator=g) / 16
    X = torch.randn(4, 300, 160, generator=g)
    Wd, Xd = W.cuda().requires_grad_(True), X.cuda().requires_grad_(True)
    y = Kn.linear(Xd[..., 32:160], Wd[:, 128:])
    y.square().sum().backward()
    Ws, Xs = W.double().requires_grad_(True), X.double().requires_grad_(True)
    y_ref = Xs[..., 32:160] @ Ws[:, 128:].t()
    y_ref.square().sum().backward()
    assert rel_err(y, y_ref) < TOL
    assert rel_err(Wd.grad, Ws.grad) < TOL
    assert rel_err(Xd.grad, Xs.grad) < TOL


@pytest.mark.parametrize("B,Nq,Nk,H,D", [
    (2, 300, 77, 8, 16), (1, 4096, 512, 8, 16), (3, 129, 600, 8, 16), (2, 257, 64, 4, 32), (2, 128, 1, 8, 16),
    (5, 1, 130, 2, 16), (1, 1000, 1100, 8, 16),
])
def test_mha_fast(B, Nq, Nk, H, D):
    """tcgen05 flash attention (resident and streamed K/V, ragged tails, one key) against float64 softmax attention"""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(B * 1000 + Nq + Nk)
    E = H * D
    q, k, v = (torch.randn(B, n, E, generator=g) for n in (Nq, Nk, Nk))
    do = torch.randn(B, Nq, E, generator=g)
    scale = 1.0 / math.sqrt(D)

    def ref(q, k, v):
        qh, kh, vh = (t.view(B, -1, H, D).transpose(1, 2) for t in (q, k, v))
        a = torch.softmax(qh @ kh.transpose(-1, -2) * scale, -1)
        return (a @ vh).transpose(1, 2).reshape(B, -1, E)

    qs, ks, vs = (t.double().requires_grad_(True) for t in (q, k, v))
    o_ref = ref(qs, ks, vs)
    o_ref.backward(do.double())
    qd, kd, vd = (t.cuda().requires_grad_(True) for t in (q, k, v))
    o = Kn.mha(qd, kd, vd, H, scale)
    o.backward(do.cuda())
    assert rel_err(o, o_ref) < 3e-3, ("out", rel_err(o, o_ref))
    for n, a, r in (("dq", qd.grad, qs.grad), ("dk", kd.grad, ks.grad), ("dv", vd.grad, vs.grad)):
        if Nk == 1 and n != "dv":      # a single key: softmax is constant, dq = dk = 0 up to rounding
            assert float(a.abs().max()) < 2e-2      # tf32 rounding of dO.V against the fp32 dO.O
            continue
        assert rel_err(a, r) < 5e-3, (n, rel_err(a, r))


@pytest.mark.parametrize("B,Nq,Nk,mode", [(2, 100, 37, "fast"), (3, 256, 128, "fast"), (2, 200, 64, "fast"),
                                         (5, 1192, 300, "fast"), (2, 100, 37, "fp32")])
def test_transformer_block_gradient_side_channels(B, Nq, Nk, mode):
    """TransformerAttender with its three GradStash side channels (LN2 -> LN1 dout2, LN1 -> query transform dX
    reduce-add, LN1 -> attention backward delta) against the same module with plain autograd fan-out sums."""
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.architectures import attenders

    g = torch.Generator().manual_seed(B * 1000 + Nq)
    keys, queries, values = (torch.randn(B, n, 128, generator=g) for n in (Nk, Nq, Nk))
    dy = torch.randn(B, Nq, 128, generator=g)
    torch.manual_seed(1)
    mod = attenders.TransformerAttender(128, 128, 128, n_heads=8).cuda()
    npf.set_mode(mode)
    old = attenders._USE_STASH
    try:
        res = []
        for use in (False, True):
            attenders._USE_STASH = use
            ins = [t.detach().cuda().requires_grad_(True) for t in (keys, queries, values)]
            for p_ in mod.parameters():
                p_.grad = None
            out = mod(*ins)
            out.backward(dy.cuda())
            res.append((out.detach(), [t.grad for t in ins] + [p_.grad.clone() for p_ in mod.parameters()]))
        assert torch.equal(res[0][0], res[1][0])
        names = ["dkeys", "dqueries", "dvalues"] + [n for n, _ in mod.named_parameters()]
        for n, a, b in zip(names, res[1][1], res[0][1]):
            err = float((a - b).abs().max() / (b.abs().max() + 1e-12))
            assert err < (5e-4 if mode == "fast" else 2e-5), (n, err)     # fast: tf32 operands re-truncate dS
    finally:
        attenders._USE_STASH = old
        npf.set_mode("fp32")
