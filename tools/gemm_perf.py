"""Developer probe: per-call time of the 128->128 Linear kernels on 1.05 M rows (C2 target rows), fast mode."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import torch
import npf_gwwaveform_b200 as npf
from npf_gwwaveform_b200 import kernels as Kn
from npf_gwwaveform_b200._lib import LIB
npf.set_mode("fast")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 256 * 4096
g = torch.randn(M, 128, device="cuda"); h = torch.relu(torch.randn(M, 128, device="cuda"))
w = Kn.tf32_weight(torch.randn(128, 128, device="cuda") / 11)
dx = torch.empty(M, 128, device="cuda"); dw = torch.zeros(128, 128, device="cuda"); db = torch.zeros(128, device="cuda")
p = lambda t: t.data_ptr()
st = torch.cuda.current_stream().cuda_stream
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
u = M * 512
f_dx = lambda: LIB.call("npfk_linear_bwd_dx", p(g), 128, p(w), 128, p(h), 128, p(dx), 128, M, 128, 128, 1, st)
f_dw = lambda: LIB.call("npfk_linear_bwd_dw", p(g), 128, p(h), 128, p(dw), 128, p(db), M, 128, 128, 1, st)
f_fu = lambda: LIB.call("npfk_linear_bwd_fused", p(g), 128, p(h), 128, p(w), 128, 1, p(dx), 128, p(dw), 128, p(db), M, 128, 128, 1, st)
for name, fn, units in (("dgrad(masked)", f_dx, 3), ("wgrad", f_dw, 2), ("fused", f_fu, 3)):
    us = t(fn)
    print(f"{name:14s} {us:8.1f} us  {units * u / us / 1e3:7.1f} GB/s algorithmic ({units} x {u / 1e6:.0f} MB)")
print("abort code", LIB.load().npfk_tc_abort_code())
