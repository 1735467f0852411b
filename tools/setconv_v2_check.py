"""Developer probe: resident-value SetConv kernels (NPFK_SETCONV_V2=1) against float64 at a ConvCNP shape."""
import math, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from npf_gwwaveform_b200 import kernels as Kn
B, Nk, Nq, C = 2, 192, 200, 128
g = torch.Generator().manual_seed(1)
xk = torch.linspace(-1.5, 1.5, Nk).view(1, Nk, 1).expand(B, Nk, 1).contiguous()
xq = torch.linspace(-1, 1, Nq).view(1, Nq, 1).expand(B, Nq, 1).contiguous()
v = torch.randn(B, Nk, C, generator=g); wd = torch.randn(C, generator=g); bs = torch.randn(C, generator=g)
theta = torch.tensor([math.log(math.expm1(0.01204))])
dy = torch.randn(B, Nq, C, generator=g)
def ref():
    vs, ts, ws = v.double().requires_grad_(True), theta.double().requires_grad_(True), wd.double().requires_grad_(True)
    sig = 1e-5 + torch.nn.functional.softplus(ts)
    inp = -(((xq.double() - xk.double().transpose(1, 2)).abs() / sig) ** 2)          # [B,Nq,Nk]
    w = torch.softmax(inp, -1); dens = torch.exp(inp).sum(-1, keepdim=True)
    out = w @ vs + dens * ws + bs.double()
    out.backward(dy.double())
    return out.detach(), vs.grad, ts.grad, ws.grad
ro, rdv, rdt, rdw = ref()
vd, td, wdd, bd = v.cuda().requires_grad_(True), theta.cuda().requires_grad_(True), wd.cuda().requires_grad_(True), bs.cuda().requires_grad_(True)
o = Kn.setconv_interp(xk.cuda(), xq.cuda(), vd, td, wdd, bd, 2.0, False)
o.backward(dy.cuda())
rel = lambda a, b: float((a.detach().cpu().double() - b).abs().max() / b.abs().max())
print(f"V2={os.environ.get('NPFK_SETCONV_V2','1')}: out {rel(o, ro):.2e} dv {rel(vd.grad, rdv):.2e} dtheta {rel(td.grad, rdt):.2e} ({float(td.grad):.6e} vs {float(rdt):.6e}) dwd {rel(wdd.grad, rdw):.2e}")
