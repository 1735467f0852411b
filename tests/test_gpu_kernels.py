"""Per-kernel parity: every npfk entry point (through the C-ABI via npf_gwwaveform_b200.kernels) against the CPU
oracle's restatement of the same reference op, forward and backward, on ragged / edge shapes.

Tolerance (fp32 mode): normalised max error  max|a-b| / max|b|  <= 1e-4  (BASELINE.json north_star); gradients
are reported against the same bound relaxed to 5e-4 (atomics change the summation order).
"""
import math

import pytest
import torch

from oracle import npf_oracle as O
from util_models import rel_err

pytestmark = pytest.mark.gpu
TOL, GTOL = 1e-4, 5e-4


def dev(t, grad=False):
    return t.detach().cuda().requires_grad_(grad)


def cpu(t, grad=False):
    return t.detach().double().requires_grad_(grad)


def check_grads(names, ours, theirs, tol=GTOL):
    for n, a, b in zip(names, ours, theirs):
        assert a is not None, n
        assert rel_err(a, b) < tol, (n, rel_err(a, b))


@pytest.mark.parametrize("M,N,K,relu,res_mode", [
    (300, 128, 128, True, "none"), (257, 128, 1, True, "none"), (129, 2, 128, False, "none"),
    (500, 128, 32, True, "full"), (3 * 70, 64, 129, False, "tiled"), (1, 256, 128, True, "none"),
    (1000, 128, 256, False, "full"), (64, 32, 2, True, "none"),
])
def test_linear(M, N, K, relu, res_mode):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(M + N + K)
    x, w, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / math.sqrt(K), torch.randn(N, generator=g)
    res = None
    if res_mode == "full":
        res = torch.randn(M, N, generator=g)
    elif res_mode == "tiled":
        res = torch.randn(M // 3, N, generator=g)
    dy = torch.randn(M, N, generator=g)

    xs, ws, bs = cpu(x, True), cpu(w, True), cpu(b, True)
    rs = cpu(res, True) if res is not None else None
    y_ref = xs @ ws.t() + bs
    if rs is not None:
        y_ref = y_ref + (rs if res_mode == "full" else rs.repeat(3, 1))
    if relu:
        y_ref = torch.relu(y_ref)
    y_ref.backward(dy.double())

    xd, wd, bd = dev(x, True), dev(w, True), dev(b, True)
    rd = dev(res, True) if res is not None else None
    y = Kn.linear(xd, wd, bd, res=rd, relu=relu)
    y.backward(dy.cuda())
    assert rel_err(y, y_ref) < TOL
    check_grads(["dx", "dw", "db"], [xd.grad, wd.grad, bd.grad], [xs.grad, ws.grad, bs.grad])
    if rd is not None:
        check_grads(["dres"], [rd.grad], [rs.grad])


def test_linear_strided_operands():
    """column slices of a wider weight (split r_z_merger, base.py:570-575) and of a wider input"""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(5)
    W = torch.randn(64, 96, generator=g)
    X = torch.randn(4, 50, 80, generator=g)
    Wd, Xd = dev(W, True), dev(X, True)
    y = Kn.linear(Xd[..., :32], Wd[:, 64:])
    y.sum().backward()
    Ws, Xs = cpu(W, True), cpu(X, True)
    y_ref = Xs[..., :32] @ Ws[:, 64:].t()
    y_ref.sum().backward()
    assert rel_err(y, y_ref) < TOL
    check_grads(["dW", "dX"], [Wd.grad, Xd.grad], [Ws.grad, Xs.grad])


@pytest.mark.parametrize("Z,B,T,C", [(1, 3, 17, 32), (4, 2, 300, 128), (2, 1, 1, 8), (3, 2, 70, 6), (5, 2, 129, 160)])
def test_bcast_add_relu(Z, B, T, C):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(Z * 100 + T)
    a, r, dy = torch.randn(B, T, C, generator=g), torch.randn(Z, B, C, generator=g), torch.randn(Z, B, T, C, generator=g)
    as_, rs = cpu(a, True), cpu(r, True)
    ref = torch.relu(as_.unsqueeze(0) + rs.unsqueeze(2))
    ref.backward(dy.double())
    ad, rd = dev(a, True), dev(r, True)
    out = Kn.bcast_add_relu(ad, rd)
    out.backward(dy.cuda())
    assert rel_err(out, ref) < 1e-6
    check_grads(["da", "dr"], [ad.grad, rd.grad], [as_.grad, rs.grad])


@pytest.mark.parametrize("B,N,C", [(3, 7, 32), (2, 513, 128), (1, 1, 5)])
def test_set_mean(B, N, C):
    from npf_gwwaveform_b200 import kernels as Kn

    x = torch.randn(B, N, C, generator=torch.Generator().manual_seed(N))
    xs = cpu(x, True)
    ref = xs.mean(1, keepdim=True)
    ref.square().sum().backward()
    xd = dev(x, True)
    out = Kn.set_mean(xd)
    out.square().sum().backward()
    assert rel_err(out, ref) < 1e-5
    check_grads(["dx"], [xd.grad], [xs.grad])


@pytest.mark.parametrize("B,Nq,Nk,H,D", [
    (2, 37, 19, 8, 4), (2, 130, 65, 8, 16), (1, 5, 300, 1, 32), (2, 33, 7, 1, 128), (1, 129, 1, 2, 64), (3, 9, 9, 4, 8),
])
def test_mha(B, Nq, Nk, H, D):
    """heads read in place from the feature dim; compared with the reference's permute/bmm/softmax formulation"""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(Nq * 7 + Nk)
    E = H * D
    q, k, v = (torch.randn(B, n, E, generator=g) for n in (Nq, Nk, Nk))
    dy = torch.randn(B, Nq, E, generator=g)
    qs, ks, vs = cpu(q, True), cpu(k, True), cpu(v, True)
    Qh, Kh, Vh = (t.view(B, -1, H, D).permute(0, 2, 1, 3) for t in (qs, ks, vs))
    ref = (torch.softmax(Qh @ Kh.transpose(-1, -2) / math.sqrt(D), -1) @ Vh).permute(0, 2, 1, 3).reshape(B, Nq, E)
    ref.backward(dy.double())
    qd, kd, vd = dev(q, True), dev(k, True), dev(v, True)
    out = Kn.mha(qd, kd, vd, H, 1.0 / math.sqrt(D))
    out.backward(dy.cuda())
    assert rel_err(out, ref) < TOL
    check_grads(["dq", "dk", "dv"], [qd.grad, kd.grad, vd.grad], [qs.grad, ks.grad, vs.grad])


@pytest.mark.parametrize("rows,C,with_b", [(33, 128, True), (7, 32, True), (1000, 256, False), (5, 1000, True),
                                           (20001, 128, False), (1, 128, True)])
def test_add_layernorm(rows, C, with_b):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(rows + C)
    a, b = torch.randn(rows, C, generator=g), torch.randn(rows, C, generator=g)
    gam, bet, dy = 1 + 0.1 * torch.randn(C, generator=g), torch.randn(C, generator=g), torch.randn(rows, C, generator=g)
    sd = {"ln.weight": cpu(gam, True), "ln.bias": cpu(bet, True)}
    as_, bs = cpu(a, True), cpu(b, True)
    ref = O.layer_norm(sd, "ln", as_ + bs if with_b else as_)
    ref.backward(dy.double())
    ad, bd, gd, btd = dev(a, True), dev(b, True), dev(gam, True), dev(bet, True)
    out = Kn.add_layernorm(ad, bd if with_b else None, gd, btd)
    out.backward(dy.cuda())
    assert rel_err(out, ref) < TOL
    check_grads(["da", "dgamma", "dbeta"], [ad.grad, gd.grad, btd.grad], [as_.grad, sd["ln.weight"].grad, sd["ln.bias"].grad])
    if with_b:
        check_grads(["db"], [bd.grad], [bs.grad])


def test_add_layernorm_parked_fanout_gradient():
    """c = LN1(x + y); out = LN2(c + f(c)) with the GradStash side channel (LN2 parks its gradient for c, LN1 adds it
    to f's while reading) against the same block through plain autograd fan-out."""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(11)
    rows, C = 300, 128
    x, y = torch.randn(rows, C, generator=g), torch.randn(rows, C, generator=g)
    W = torch.randn(C, C, generator=g) / math.sqrt(C)
    g1, b1, g2, b2 = (torch.randn(C, generator=g) * 0.3 + 1 for _ in range(4))
    dy = torch.randn(rows, C, generator=g)

    def block(use_stash):
        ts = [dev(t, True) for t in (x, y, W, g1, b1, g2, b2)]
        xd, yd, Wd, g1d, b1d, g2d, b2d = ts
        st = Kn.GradStash() if use_stash else None
        c = Kn.add_layernorm(xd, yd, g1d, b1d, extra=st)
        out = Kn.add_layernorm(c, Kn.linear(c, Wd, relu=True), g2d, b2d, park_a=st)
        out.backward(dy.cuda())
        assert st is None or st.t is None          # consumed by LN1's backward
        return out, [t.grad for t in ts]

    o_ref, g_ref = block(False)
    o_st, g_st = block(True)
    assert torch.equal(o_ref, o_st)
    check_grads(["dx", "dy", "dW", "dg1", "db1", "dg2", "db2"], g_st, [t.double().cpu() for t in g_ref], tol=1e-5)


@pytest.mark.parametrize("mode,rows", [("fp32", 300), ("fast", 300), ("fast", 148 * 128 + 50)])
def test_linear_dx_onto_parked_gradient(mode, rows):
    """q = Linear(x); c = LN(f(q) + x): LN parks its gradient for x and the Linear's backward adds its dx onto it
    (TMA reduce-add in the fused tensor-core kernel in fast mode, a plain add otherwise) - against autograd's fan-out."""
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(5)
    C = 128
    x = torch.randn(rows, C, generator=g)
    Wq, bq = torch.randn(C, C, generator=g) / math.sqrt(C), torch.randn(C, generator=g) * 0.1
    g1, b1 = torch.randn(C, generator=g) * 0.3 + 1, torch.randn(C, generator=g) * 0.1
    dy = torch.randn(rows, C, generator=g)
    npf.set_mode(mode)
    try:
        def block(use_stash):
            xd, Wd, bd, g1d, b1d = (dev(t, True) for t in (x, Wq, bq, g1, b1))
            st = Kn.GradStash() if use_stash else None
            q = Kn.linear(xd, Wd, bd, dx_stash=st)
            out = Kn.add_layernorm(torch.tanh(q), xd, g1d, b1d, park_b=st)
            out.backward(dy.cuda())
            assert st is None or st.t is None
            return out, [xd.grad, Wd.grad, bd.grad, g1d.grad, b1d.grad]

        o_ref, g_ref = block(False)
        o_st, g_st = block(True)
        assert torch.equal(o_ref, o_st)
        check_grads(["dx", "dW", "db", "dg1", "db1"], g_st, [t.double().cpu() for t in g_ref],
                    tol=1e-5 if mode == "fp32" else 1e-4)
    finally:
        npf.set_mode("fp32")


def _setconv_ref(xk, xq, v, theta, W, bias, unshared):
    sd = {"sc.radial_basis_func.length_scale_param": theta, "sc.resizer.weight": W, "sc.resizer.bias": bias}
    return O.set_conv(sd, "sc", xk, xq, v)


@pytest.mark.parametrize("B,Nk,Nq,C,unshared,shared_grid", [
    (2, 13, 48, 1, False, False),     # context -> induced grid (queries share one grid)
    (2, 48, 70, 32, False, True),     # induced grid -> targets (keys share one grid)
    (1, 192, 130, 128, False, True),
    (2, 5, 9, 160, False, False),     # C > 128: two channel passes
    (2, 40, 33, 16, True, True),      # fork's UnsharedExpRBF
    (1, 1, 3, 2, False, False),
])
def test_setconv(B, Nk, Nq, C, unshared, shared_grid):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(Nk * 3 + Nq)
    if shared_grid:
        xk = torch.linspace(-1.5, 1.5, Nk).view(1, Nk, 1).expand(B, Nk, 1)
    else:
        xk = torch.rand(B, Nk, 1, generator=g) * 2 - 1
    xq = torch.rand(B, Nq, 1, generator=g) * 2 - 1
    v = torch.randn(B, Nk, C, generator=g)
    # length scale of the order of the key spacing, as at initialisation (sigma 0.012 vs grid step 0.016)
    sig = 3.0 / max(Nk, 8)
    theta = torch.full((2 if unshared else 1,), math.log(math.expm1(sig)))
    if unshared:
        theta[1] += 0.3
    W, bias = torch.randn(24, C + 1, generator=g) / math.sqrt(C + 1), torch.randn(24, generator=g)
    dy = torch.randn(B, Nq, 24, generator=g)

    ts, vs, Ws, bs = cpu(theta, True), cpu(v, True), cpu(W, True), cpu(bias, True)
    ref = _setconv_ref(xk.double(), xq.double(), vs, ts, Ws, bs, unshared)
    ref.backward(dy.double())

    td, vd, Wd, bd = dev(theta, True), dev(v, True), dev(W, True), dev(bias, True)
    t = Kn.setconv_rbf(xk.cuda(), xq.cuda(), vd, td, 2.0, unshared)
    out = Kn.linear(t, Wd, bd)
    out.backward(dy.cuda())
    assert rel_err(out, ref) < TOL
    check_grads(["dv", "dtheta", "dW"], [vd.grad, td.grad, Wd.grad], [vs.grad, ts.grad, Ws.grad])


@pytest.mark.parametrize("B,Nk,Nq,C,O_,unshared,keys", [
    (2, 48, 70, 32, 24, False, "grid"),       # induced grid -> targets
    (1, 192, 4096, 128, 128, False, "grid"),  # C3 shape of one task: most key chunks are skipped
    (2, 192, 130, 128, 128, False, "grid"),
    (2, 100, 300, 128, 128, False, "random"),  # resident-value kernel, unsorted keys: window = every key, in rounds
    (2, 64, 257, 128, 128, True, "grid"),      # resident-value kernel, UnsharedExpRBF
    (3, 192, 1111, 128, 128, False, "grid_sorted_q"),   # sorted targets: one round per warp (the ConvCNP case)
    (1, 256, 600, 128, 128, False, "clustered"),
    (2, 70, 200, 16, 160, False, "random"),   # unsorted keys, two channel passes of the output
    (2, 40, 90, 16, 24, True, "grid"),        # fork's UnsharedExpRBF
    (1, 1, 3, 2, 8, False, "random"),
    (2, 33, 65, 8, 16, False, "clustered"),   # queries far from every key (softmax max far below 0)
])
def test_setconv_interp_hoisted_resizer(B, Nk, Nq, C, O_, unshared, keys):
    """SetConv.forward with the resizer applied to the keys' values first (npfk_setconv_interp_*), against the
    oracle's unhoisted restatement of setcnn.py:254-268."""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(Nk * 5 + Nq)
    if keys == "grid":
        xk = torch.linspace(-1.5, 1.5, Nk).view(1, Nk, 1).expand(B, Nk, 1)
    elif keys == "clustered":
        xk = torch.rand(B, Nk, 1, generator=g) * 0.2 - 0.9
    else:
        xk = torch.rand(B, Nk, 1, generator=g) * 2 - 1
    xq = torch.rand(B, Nq, 1, generator=g) * 2 - 1
    if keys == "grid_sorted_q":
        xk = torch.linspace(-1.5, 1.5, Nk).view(1, Nk, 1).expand(B, Nk, 1)
        xq = torch.linspace(-1, 1, Nq).view(1, Nq, 1).expand(B, Nq, 1).contiguous()
    v = torch.randn(B, Nk, C, generator=g)
    sig = 3.0 / max(Nk, 8)
    theta = torch.full((2 if unshared else 1,), math.log(math.expm1(sig)))
    if unshared:
        theta[1] += 0.3
    W, bias = torch.randn(O_, C + 1, generator=g) / math.sqrt(C + 1), torch.randn(O_, generator=g)
    dy = torch.randn(B, Nq, O_, generator=g)

    ts, vs, Ws, bs = cpu(theta, True), cpu(v, True), cpu(W, True), cpu(bias, True)
    ref = _setconv_ref(xk.double(), xq.double(), vs, ts, Ws, bs, unshared)
    ref.backward(dy.double())

    td, vd, Wd, bd = dev(theta, True), dev(v, True), dev(W, True), dev(bias, True)
    vp = Kn.linear(vd, Wd[:, :C])
    out = Kn.setconv_interp(xk.cuda(), xq.cuda(), vp, td, Wd[:, C], bd, 2.0, unshared)
    out.backward(dy.cuda())
    assert rel_err(out, ref) < TOL
    check_grads(["dv", "dW", "db"], [vd.grad, Wd.grad, bd.grad], [vs.grad, Ws.grad, bs.grad])
    check_grads(["dtheta"], [td.grad], [ts.grad], tol=2e-3)


@pytest.mark.parametrize("B,Nk,Nq,C,unshared", [
    (2, 512, 200, 1, False),    # C3's context -> grid shape
    (3, 37, 70, 1, False), (2, 50, 9, 3, False), (2, 33, 40, 4, False), (2, 45, 33, 2, True), (1, 1, 3, 1, False),
])
def test_setconv_data_values(B, Nk, Nq, C, unshared):
    """context -> grid: the values are data, so only the length scale (and the resizer) get gradients; few-channel
    inputs take the warp-per-query kernels (forward for any C <= 4, backward when dv is not requested)."""
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(Nk * 7 + Nq)
    xk = torch.rand(B, Nk, 1, generator=g) * 2 - 1
    xq = torch.linspace(-1.1, 1.1, Nq).view(1, Nq, 1).expand(B, Nq, 1)
    v = torch.randn(B, Nk, C, generator=g)
    sig = 3.0 / max(Nk, 8)
    theta = torch.full((2 if unshared else 1,), math.log(math.expm1(sig)))
    if unshared:
        theta[1] += 0.3
    W, bias = torch.randn(24, C + 1, generator=g) / math.sqrt(C + 1), torch.randn(24, generator=g)
    dy = torch.randn(B, Nq, 24, generator=g)

    ts, Ws, bs = cpu(theta, True), cpu(W, True), cpu(bias, True)
    ref = _setconv_ref(xk.double(), xq.double(), v.double(), ts, Ws, bs, unshared)
    ref.backward(dy.double())

    td, Wd, bd = dev(theta, True), dev(W, True), dev(bias, True)
    t = Kn.setconv_rbf(xk.cuda(), xq.cuda(), v.cuda(), td, 2.0, unshared)
    out = Kn.linear(t, Wd, bd)
    out.backward(dy.cuda())
    assert rel_err(out, ref) < TOL
    check_grads(["dtheta", "dW"], [td.grad, Wd.grad], [ts.grad, Ws.grad])


def test_setconv_far_queries_are_defined():
    """G3: softmax weights stay defined when every exp underflows; density -> 0 (setcnn.py:134-140)"""
    from npf_gwwaveform_b200 import kernels as Kn

    xk = torch.tensor([[-0.9, -0.8, -0.7]]).view(1, 3, 1)
    xq = torch.tensor([[1.0, 0.95]]).view(1, 2, 1)
    v = torch.tensor([[1.0, 2.0, 3.0]]).view(1, 3, 1)
    theta = torch.tensor([math.log(math.expm1(0.012))])
    out = Kn.setconv_rbf(xk.cuda(), xq.cuda(), v.cuda(), theta.cuda(), 2.0, False).cpu()
    assert torch.isfinite(out).all()
    torch.testing.assert_close(out[0, :, 0], torch.tensor([3.0, 3.0]))   # nearest key wins
    assert float(out[0, :, 1].abs().max()) == 0.0


@pytest.mark.parametrize("B,L,C,ks,norm,training,with_res", [
    (3, 48, 32, 5, True, True, True), (2, 192, 128, 19, True, False, True), (2, 24, 64, 11, False, True, False),
    (1, 7, 40, 3, True, True, False), (2, 100, 256, 19, True, True, True),
])
def test_norm_relu_dwconv(B, L, C, ks, norm, training, with_res):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(L + C)
    x = torch.randn(B, L, C, generator=g) + 0.3
    w, bias = torch.randn(C, 1, ks, generator=g) / ks, torch.randn(C, generator=g)
    gam, bet = 1 + 0.2 * torch.randn(C, generator=g), 0.2 * torch.randn(C, generator=g)
    rm, rv = 0.2 * torch.randn(C, generator=g), 0.5 + torch.rand(C, generator=g)
    res = torch.randn(B, L, C, generator=g) if with_res else None
    dy = torch.randn(B, L, C, generator=g)

    sd = {"c.weight": cpu(w, True), "c.bias": cpu(bias, True)}
    if norm:
        sd.update({"n.weight": cpu(gam, True), "n.bias": cpu(bet, True), "n.running_mean": rm.double(),
                   "n.running_var": rv.double(), "n.num_batches_tracked": torch.tensor(0)})
    xs = cpu(x, True)
    rs = cpu(res, True) if with_res else None
    new_stats = {}
    h = torch.relu(O.batch_norm(sd, "n", xs, training, new_stats))
    ref = O.depthwise_conv(sd, "c", h)
    if with_res:
        ref = ref + rs
    ref.backward(dy.double())

    xd, wd, bd = dev(x, True), dev(w, True), dev(bias, True)
    gd, btd = (dev(gam, True), dev(bet, True)) if norm else (None, None)
    rmd, rvd = (rm.cuda(), rv.cuda()) if norm else (None, None)
    rd = dev(res, True) if with_res else None
    nbt = torch.tensor(3, device="cuda") if norm and training else None
    y = Kn.norm_act_dwconv(xd, gd, btd, wd, bd, rd, rmd, rvd, training, 1e-5, 0.1, True, nbt)
    y.backward(dy.cuda())
    assert nbt is None or int(nbt) == 4        # BatchNorm1d.num_batches_tracked += 1, on the device
    assert rel_err(y, ref) < TOL
    check_grads(["dx", "dw", "dbias"], [xd.grad, wd.grad, bd.grad], [xs.grad, sd["c.weight"].grad, sd["c.bias"].grad])
    if norm:
        check_grads(["dgamma", "dbeta"], [gd.grad, btd.grad], [sd["n.weight"].grad, sd["n.bias"].grad])
        if training:   # running statistics updated with the unbiased variance, momentum 0.1
            assert rel_err(rmd, new_stats["n.running_mean"]) < 1e-5
            assert rel_err(rvd, new_stats["n.running_var"]) < 1e-5
    if with_res:
        check_grads(["dres"], [rd.grad], [rs.grad])


@pytest.mark.parametrize("B,L,C1,C2", [(2, 24, 32, 32), (1, 48, 256, 256), (3, 5, 8, 0), (2, 1, 4, 4)])
def test_pool_upsample_concat(B, L, C1, C2):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(L)
    x = torch.randn(B, L, C1, generator=g)
    skip = torch.randn(B, 2 * L, C2, generator=g) if C2 else None
    xs = cpu(x, True)
    ss = cpu(skip, True) if C2 else None
    up = O.upsample_linear2(xs)
    ref = torch.cat((up, ss), -1) if C2 else up
    dy = torch.randn(ref.shape, generator=g)
    ref.backward(dy.double())
    xd = dev(x, True)
    sdv = dev(skip, True) if C2 else None
    out = Kn.upsample2_cat(xd, sdv)
    out.backward(dy.cuda())
    assert rel_err(out, ref) < 1e-5
    check_grads(["dx"], [xd.grad], [xs.grad])
    if C2:
        check_grads(["dskip"], [sdv.grad], [ss.grad])
    # pooling (odd lengths drop the last sample like nn.AvgPool1d)
    for Lp in (L, 2 * L + 1):
        z = torch.randn(B, Lp, C1, generator=g)
        if Lp < 2:
            continue
        zs, zd = cpu(z, True), dev(z, True)
        pr = O.avg_pool2(zs)
        pr.square().sum().backward()
        po = Kn.avgpool2(zd)
        po.square().sum().backward()
        assert rel_err(po, pr) < 1e-6
        check_grads(["dz"], [zd.grad], [zs.grad])


@pytest.mark.parametrize("Z,B,T,ydim", [(1, 3, 50, 1), (4, 2, 3000, 1), (2, 2, 17, 3)])
def test_gauss_nll_and_loc_scale(Z, B, T, ydim):
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(T)
    suff = torch.randn(Z, B, T, 2 * ydim, generator=g) * 2
    suff[0, 0, 0, ydim:] = 30.0       # softplus threshold branch
    suff[0, 0, 1 % T, ydim:] = -30.0  # scale floor 0.01
    y = torch.randn(B, T, ydim, generator=g)
    ss = cpu(suff, True)
    loc, scale = O.suffstat_to_loc_scale(ss, ydim)
    ref = O.sum_log_prob(loc, scale, y.double())
    wts = torch.randn(Z, B, generator=g).double()
    (ref * wts).sum().backward()
    sdv = dev(suff, True)
    out = Kn.gauss_sum_log_prob(sdv, y.cuda())
    (out * wts.float().cuda()).sum().backward()
    assert rel_err(out, ref) < TOL
    check_grads(["dsuff"], [sdv.grad], [ss.grad])
    # explicit loc / scale tensors (what the returned distribution exposes)
    s2 = dev(suff, True)
    l, s = Kn.loc_scale(s2)
    assert rel_err(l, loc) < 1e-6 and rel_err(s, scale) < 1e-5
    (l.sum() + (s * s).sum()).backward()
    s3 = cpu(suff, True)
    l3, sc3 = O.suffstat_to_loc_scale(s3, ydim)
    (l3.sum() + (sc3 * sc3).sum()).backward()
    check_grads(["dsuff(loc_scale)"], [s2.grad], [s3.grad])


def test_latent_pieces():
    from npf_gwwaveform_b200 import kernels as Kn

    g = torch.Generator().manual_seed(11)
    B, zd, S = 5, 48, 3
    suff = torch.randn(B, 1, 2 * zd, generator=g)
    suff2 = torch.randn(B, 1, 2 * zd, generator=g)
    eps = torch.randn(S, B, 1, zd, generator=g)
    a, b = cpu(suff, True), cpu(suff2, True)
    la, sa = a.split(zd, -1)
    lb, sb = b.split(zd, -1)
    sa, sb = 0.1 + 0.9 * torch.sigmoid(sa), 0.1 + 0.9 * torch.sigmoid(sb)
    z_ref = la.unsqueeze(0) + sa.unsqueeze(0) * eps.double()
    kl_ref = O.kl_diag_normal(la, sa, lb, sb)
    (z_ref.square().sum() + (kl_ref * torch.arange(B).view(B, 1).double()).sum()).backward()
    ad, bd = dev(suff, True), dev(suff2, True)
    lad, sad = Kn.latent_stats(ad)
    lbd, sbd = Kn.latent_stats(bd)
    z = Kn.rsample(lad, sad, eps.cuda())
    kl = Kn.kl_diag_normal(lad, sad, lbd, sbd)
    (z.square().sum() + (kl * torch.arange(B).view(B, 1).float().cuda()).sum()).backward()
    assert rel_err(z, z_ref) < 1e-5 and rel_err(kl, kl_ref) < TOL
    check_grads(["dsuff_q", "dsuff_p"], [ad.grad, bd.grad], [a.grad, b.grad])


def test_range_flag_and_adam():
    from npf_gwwaveform_b200 import kernels as Kn

    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    Kn.range_flag(torch.linspace(-1, 1, 1000).cuda(), flag)
    assert int(flag.item()) == 0
    x = torch.zeros(5000)
    x[4321] = 1.0001
    Kn.range_flag(x.cuda(), flag)
    assert int(flag.item()) == 1
    flag.zero_()
    x[4321] = float("nan")
    Kn.range_flag(x.cuda(), flag)
    assert int(flag.item()) == 1
    # Adam against torch.optim.Adam on CPU
    g = torch.Generator().manual_seed(3)
    p0 = torch.randn(1000, generator=g)
    pc = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pc], lr=1e-3)
    pd, m, v = p0.clone().cuda(), torch.zeros(1000).cuda(), torch.zeros(1000).cuda()
    for step in range(1, 4):
        grad = torch.randn(1000, generator=g)
        pc.grad = grad.clone()
        opt.step()
        Kn.adam_step(pd, grad.cuda(), m, v, 1e-3, 0.9, 0.999, 1e-8, step)
    assert rel_err(pd, pc) < 1e-6


def test_launch_counter_counts_our_kernels():
    from npf_gwwaveform_b200 import kernels as Kn

    n0 = Kn.launch_count()
    Kn.linear(torch.randn(10, 8).cuda(), torch.randn(4, 8).cuda())
    assert Kn.launch_count() == n0 + 1
