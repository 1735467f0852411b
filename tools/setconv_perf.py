"""Developer probe: SetConv grid->target kernels at the C3 shape (B tasks x 192 grid points -> 4096 targets x 128)."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import torch
from npf_gwwaveform_b200 import kernels as Kn
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
Nq = 4096
grid = torch.linspace(-1.5, 1.5, 192, device="cuda"); xt = torch.linspace(-1, 1, Nq, device="cuda")
vals = torch.randn(B, 192, 128, device="cuda", requires_grad=True)
theta = torch.tensor([-4.4], device="cuda", requires_grad=True)
wd = torch.randn(128, device="cuda", requires_grad=True); bs = torch.zeros(128, device="cuda", requires_grad=True)
do = torch.randn(B, Nq, 128, device="cuda")
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
with torch.no_grad():
    f = t(lambda: Kn.setconv_interp(grid, xt, vals, theta, wd, bs))
o = Kn.setconv_interp(grid, xt, vals, theta, wd, bs)
fb = t(lambda: torch.autograd.grad(o, (vals, theta, wd, bs), do, retain_graph=True))
mb = B * Nq * 128 * 4 / 1e6
print(f"B={B}: fwd {f:.1f} us ({mb / f * 1e-3:.2f} TB/s of {mb:.0f} MB written)  bwd {fb:.1f} us ({mb / fb * 1e-3:.2f} TB/s read)")
