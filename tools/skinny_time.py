"""timing of the 1 -> 128 first layer (skinny expand) and its wgrad (skinny reduce) at the C2 target rows"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import npf_gwwaveform_b200 as npf
from npf_gwwaveform_b200 import kernels as Kn
npf.set_mode("fast")
M = 1048576
x = torch.rand(M, 1, device="cuda") * 2 - 1
w = torch.randn(128, 1, device="cuda", requires_grad=True)
b = torch.randn(128, device="cuda", requires_grad=True)
y = Kn.linear(x, w, b, relu=True)
ref = torch.relu(x.double() @ w.detach().double().T + b.detach().double())
print("fwd max err", float((y.double() - ref).abs().max()))
dy = torch.randn_like(y)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
with torch.no_grad():
    print(f"expand fwd: {t(lambda: Kn.linear(x, w, b, relu=True)):.1f} us")
def fb():
    yy = Kn.linear(x, w, b, relu=True)
    yy.backward(dy)
print(f"fwd+bwd: {t(fb):.1f} us")
import ctypes
from npf_gwwaveform_b200._lib import LIB
p = lambda t_: ctypes.c_void_p(t_.data_ptr()) if t_ is not None else None
dw, db = torch.zeros(128, 1, device="cuda"), torch.zeros(128, device="cuda")
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
print(f"reduce (wgrad of the 1 -> 128 layer): {t(lambda: LIB.call('npfk_linear_bwd_dw', p(dy), 128, p(x), 1, p(dw), 1, p(db), M, 128, 1, 1, st)):.1f} us")
