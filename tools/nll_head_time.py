"""timing of npfk_nll_head_bwd at the C2 decoder shape (1,048,576 rows, 4096 targets per task)"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from npf_gwwaveform_b200._lib import LIB
M, T = 1048576, 4096
p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
h = torch.relu(torch.randn(M, 128, device="cuda"))
hw = torch.randn(2, 128, device="cuda") * 0.1
suff = torch.randn(M, 2, device="cuda")
y = torch.randn(M, device="cuda")
dlp = torch.randn(M // T, device="cuda")
g = torch.empty(M, 128, device="cuda")
dw, db = torch.zeros(2, 128, device="cuda"), torch.zeros(2, device="cuda")
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def run():
    LIB.call("npfk_nll_head_bwd", p(h), 128, p(hw), p(suff), p(y), M, T, p(dlp), None, 1, p(g), p(dw), p(db), M, st)
for _ in range(3): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): run()
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 20 * 1e3
print(f"nll_head_bwd: {us:.1f} us = {2 * M * 512 / us / 1e6:.2f} TB/s")
