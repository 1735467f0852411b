"""Developer probe for ncu: a few launches of npfk_mha_fwd / npfk_mha_bwd at the C2 shape (fast mode)."""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402

npf.set_mode("fast")
B, Nq, Nk, H, D = 256, 4096, 512, 8, 16
E = H * D
q, k, v, do = (torch.randn(B, n, E, device="cuda") for n in (Nq, Nk, Nk, Nq))
for t in (q, k, v):
    t.requires_grad_(True)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    o = Kn.mha(q, k, v, H, 1 / math.sqrt(D))
    o.backward(do)
    q.grad = k.grad = v.grad = None
torch.cuda.synchronize()
