"""Workload for `ncu --set full`: a few launches of the hot kernels at a reduced batch (64 tasks, 512/4096)."""
import math, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import torch
import npf_gwwaveform_b200 as npf
from npf_gwwaveform_b200 import kernels as Kn
npf.set_mode("fast")
B, Nq, Nk, H, D = 64, 4096, 512, 8, 16
E = H * D
q, k, v, do = (torch.randn(B, n, E, device="cuda") for n in (Nq, Nk, Nk, Nq))
for t in (q, k, v):
    t.requires_grad_(True)
for _ in range(2):
    o = Kn.mha(q, k, v, H, 1 / math.sqrt(D))
    o.backward(do)
x = torch.randn(B * Nq, 128, device="cuda", requires_grad=True)
w1 = (torch.randn(128, 128, device="cuda") * 0.05).requires_grad_(True)
w2 = (torch.randn(128, 128, device="cuda") * 0.05).requires_grad_(True)
b1 = torch.zeros(128, device="cuda", requires_grad=True)
for _ in range(2):
    y = Kn.mlp_chain(x, [(w1, b1), (w2, b1)])
    y.sum().backward()
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
suff = torch.randn(1, 256, Nq, 2, device="cuda", requires_grad=True)
yt = torch.randn(256, Nq, 1, device="cuda")
for _ in range(2):
    lp = Kn.gauss_sum_log_prob(suff, yt)
    lp.sum().backward()
torch.cuda.synchronize()
