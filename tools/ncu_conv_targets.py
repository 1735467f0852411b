"""Workload for `ncu --set full` on the ConvCNP (C3) CNN kernels: BatchNorm statistics, depthwise conv forward /
backward and the pointwise GEMMs at the UNet's shapes (32 tasks, 192 grid points, 128..512 channels, kernel size 19)."""
import os
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402

npf.set_mode("fast")
B, ks = 32, 19
for L, C in ((192, 128), (192, 512), (96, 256)):
    x = torch.randn(B, L, C, device="cuda", requires_grad=True)
    res = torch.randn(B, L, C, device="cuda")
    gamma = torch.ones(C, device="cuda", requires_grad=True)
    beta = torch.zeros(C, device="cuda", requires_grad=True)
    w = (torch.randn(C, 1, ks, device="cuda") / ks).requires_grad_(True)
    bias = torch.zeros(C, device="cuda", requires_grad=True)
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    wp = (torch.randn(C, C, device="cuda") * 0.05).requires_grad_(True)
    for _ in range(2):
        y = Kn.norm_act_dwconv(x, gamma, beta, w, bias, res, rm, rv, True, 1e-5, 0.1)
        z = Kn.linear(y, wp, bias)
        z.backward(torch.randn_like(z))
torch.cuda.synchronize()
