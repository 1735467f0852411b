"""Workload for `ncu --set full`: two launches of each hot kernel at the bench's own shapes (C2: 256 tasks,
512 context / 4096 target points, r_dim 128; C3 SetConv: 32 tasks, 192 grid points), fast (tf32) mode."""
import math
import os
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402

npf.set_mode("fast")
B, Nq, Nk, H, D = 256, 4096, 512, 8, 16
E = H * D
q, k, v, do = (torch.randn(B, n, E, device="cuda") for n in (Nq, Nk, Nk, Nq))
for t in (q, k, v):
    t.requires_grad_(True)
for _ in range(2):
    o = Kn.mha(q, k, v, H, 1 / math.sqrt(D))
    o.backward(do)
    q.grad = k.grad = v.grad = None
del o
# MLP chain on the target rows: fused forward chain (K1), backward as one pair launch (K1b: layers 3 + 2) and one
# single-layer fused launch (layer 1: dgrad + mask + wgrad)
x = q.detach().reshape(B * Nq, E).requires_grad_(True)
w1 = (torch.randn(128, 128, device="cuda") * 0.05).requires_grad_(True)
w2 = (torch.randn(128, 128, device="cuda") * 0.05).requires_grad_(True)
w3 = (torch.randn(128, 128, device="cuda") * 0.05).requires_grad_(True)
b1 = torch.zeros(128, device="cuda", requires_grad=True)
dy = do.reshape(B * Nq, E)
for _ in range(2):
    y = Kn.mlp_chain(x, [(w1, b1), (w2, b1), (w3, b1)])
    y.backward(dy)
    x.grad = None
del y
# separate dgrad / wgrad kernels (shapes the fused kernel does not take still use them)
xd = x.detach()
for _ in range(2):
    y = Kn.mlp_chain(xd, [(w1, b1)])          # wgrad only
    y.backward(dy)
dx = torch.empty_like(xd)
from npf_gwwaveform_b200._lib import LIB  # noqa: E402

st = torch.cuda.current_stream().cuda_stream
wt = Kn.tf32_weight(w1)
for _ in range(2):
    LIB.call("npfk_linear_bwd_dx", dy.data_ptr(), 128, wt.data_ptr(), 128, xd.data_ptr(), 128, dx.data_ptr(), 128,
             B * Nq, 128, 128, 1, st)
# residual + LayerNorm
g_, b_ = torch.ones(128, device="cuda", requires_grad=True), torch.zeros(128, device="cuda", requires_grad=True)
for _ in range(2):
    o = Kn.add_layernorm(x, dy, g_, b_)
    o.backward(dy)
    x.grad = None
del o
torch.cuda.synchronize()
# SetConv induced -> target (C3 shape per task: 192 grid points, 4096 targets, 128 channels) and the Gaussian NLL
Bs = 32
grid = torch.linspace(-1.5, 1.5, 192, device="cuda")
xt = torch.linspace(-1, 1, Nq, device="cuda")
vals = torch.randn(Bs, 192, 128, device="cuda", requires_grad=True)
theta = torch.tensor([-4.4], device="cuda", requires_grad=True)
wd = torch.randn(128, device="cuda", requires_grad=True)
bs = torch.zeros(128, device="cuda", requires_grad=True)
for _ in range(2):
    o = Kn.setconv_interp(grid, xt, vals, theta, wd, bs)
    o.backward(torch.randn_like(o))
# K1: fused MLP chains at the C2 target-row count -- the decoder's flat module with the head + log-likelihood tail
# (4 layers + 128 -> 2), training (every activation written) and inference (nothing but [M, 2] leaves the chip), the
# resizer (2 layers + residual), and the tail's backward
def _layers(n, n_out=128):
    return [((torch.randn(n_out if i == n - 1 else 128, 128, device="cuda") * 0.1).requires_grad_(True),
             torch.zeros(n_out if i == n - 1 else 128, device="cuda", requires_grad=True)) for i in range(n)]


M = B * Nq
xr = torch.randn(M, 128, device="cuda").relu()
rs = torch.randn(M, 128, device="cuda")
yc = torch.randn(B, Nq, 1, device="cuda")
flat, resz = _layers(5, 2), _layers(2)
for _ in range(2):
    sf, lp = Kn.mlp_chain(xr.clone().requires_grad_(True), flat, x_is_relu=True, nll=(yc, Nq))
    lp.sum().backward()
    with torch.no_grad():
        Kn.mlp_chain(xr, flat, x_is_relu=True, nll=(yc, Nq))
    Kn.mlp_chain(xr.clone().requires_grad_(True), resz, res=rs, out_relu=True)
del sf, lp
torch.cuda.synchronize()
suff = torch.randn(1, 256, Nq, 2, device="cuda", requires_grad=True)
yt = torch.randn(256, Nq, 1, device="cuda")
for _ in range(2):
    lp = Kn.gauss_sum_log_prob(suff, yt)
    lp.sum().backward()
torch.cuda.synchronize()
# the 1 -> 128 first layer of the x encoder on the target rows: streaming expand forward, one-wave reduce (wgrad)
x1 = torch.rand(M, 1, device="cuda") * 2 - 1
w1s = torch.randn(128, 1, device="cuda", requires_grad=True)
b1s = torch.zeros(128, device="cuda", requires_grad=True)
for _ in range(2):
    y1 = Kn.linear(x1, w1s, b1s, relu=False)
    y1.backward(dy)
torch.cuda.synchronize()
