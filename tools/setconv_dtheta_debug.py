"""Diagnose the SetConv length-scale gradient against float64 autograd (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from npf_gwwaveform_b200 import kernels as Kn


def ref(xk, xq, v, theta):
    sigma = 1e-5 + torch.nn.functional.softplus(theta)
    inp = -((xk[:, None, :] - xq[:, :, None]).abs() / sigma) ** 2      # [B,Nq,Nk]
    w = torch.softmax(inp, -1)
    dens = torch.exp(inp).sum(-1, keepdim=True)
    return torch.cat([w @ v, dens], -1), inp, w


for (B, Nk, Nq, C, th) in [(2, 33, 192, 1, -4.4), (2, 192, 129, 128, -4.4), (4, 512, 192, 1, -4.4), (2, 33, 192, 1, -2.0)]:
    g = torch.Generator().manual_seed(1)
    xk = torch.rand(B, Nk, generator=g) * 2 - 1
    xq = torch.linspace(-1.5, 1.5, Nq).expand(B, Nq).contiguous()
    v = torch.randn(B, Nk, C, generator=g)
    do = torch.randn(B, Nq, C + 1, generator=g)
    theta = torch.tensor([th])
    t64 = theta.double().requires_grad_(True)
    o64, inp, w = ref(xk.double(), xq.double(), v.double(), t64)
    (o64 * do.double()).sum().backward()
    t32 = theta.clone().requires_grad_(True)
    o32, _, _ = ref(xk, xq, v, t32)
    (o32 * do).sum().backward()
    td = theta.cuda().requires_grad_(True)
    od = Kn.setconv_rbf(xk.cuda().unsqueeze(-1), xq.cuda().unsqueeze(-1), v.cuda(), td)
    (od * do.cuda()).sum().backward()
    # per-element terms in float64 to see how much cancels
    dT, dD = do[..., :C].double(), do[..., C:].double()
    T = o64[..., :C].detach()
    dp = dT @ v.double().transpose(1, 2)
    dot = (dT * T).sum(-1, keepdim=True)
    ds = w.detach() * (dp - dot)
    sig = float(1e-5 + torch.nn.functional.softplus(theta.double()))
    terms = (ds * inp.detach() + dD * torch.exp(inp.detach()) * inp.detach()) * (-2 / sig)
    print(f"B{B} Nk{Nk} Nq{Nq} C{C} th{th}: ref64 {float(t64.grad):+.6e} torch32 {float(t32.grad):+.6e} "
          f"ours {float(td.grad):+.6e} rel(ours) {abs(float(td.grad) - float(t64.grad)) / abs(float(t64.grad)):.2e} "
          f"rel(torch32) {abs(float(t32.grad) - float(t64.grad)) / abs(float(t64.grad)):.2e} "
          f"sum|terms| {float(terms.abs().sum()):.3e} out_err {float((od.cpu() - o64.detach()).abs().max()):.2e}")
