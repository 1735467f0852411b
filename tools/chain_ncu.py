"""Workload for ncu: a few launches of the fused MLP chain (6 layers, training and inference) at the C2 row count."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402

npf.set_mode("fast")
M = 256 * 4096
x = torch.randn(M, 128, device="cuda").relu()
ls = [((torch.randn(128, 128, device="cuda") * 0.1).requires_grad_(True), torch.zeros(128, device="cuda", requires_grad=True))
      for _ in range(6)]
for train in (False, True, False, True):
    with torch.set_grad_enabled(train):
        Kn.mlp_chain(x.clone().requires_grad_(train), ls, out_relu=True)
torch.cuda.synchronize()
