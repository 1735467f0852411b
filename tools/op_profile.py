"""Developer probe: which torch ops (not npfk kernels) launch device kernels in one training step, with the Python
source line that issued them.  Usage: op_profile.py [model] [B Nc Nt]"""
import os
import sys
from collections import Counter

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402
from npf_gwwaveform_b200.neuralproc import set_lazy_validation  # noqa: E402
from npf_gwwaveform_b200.training import FlatAdamTrainer  # noqa: E402
from util_models import build_product, chirp_batch  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "ConvCNP"
B, Nc, Nt = (int(a) for a in sys.argv[2:5]) if len(sys.argv) > 4 else (32, 512, 4096)
set_lazy_validation(True)
npf.set_mode("fast")
torch.manual_seed(0)
m = build_product(name, 128).cuda().train()
inp = {k: v.cuda() for k, v in chirp_batch(B, Nc, Nt, 0).items()}
crit = npf.CNPFLoss().train()
tr = FlatAdamTrainer(m, crit, lr=1e-3)
for _ in range(3):
    tr.step(inp)
torch.cuda.synchronize()
Kn.set_direct_grad(os.environ.get("NPFK_DIRECT_GRAD", "0") == "1")
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True, record_shapes=True) as prof:
    tr.step(inp)
    torch.cuda.synchronize()
cnt = Counter()
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CPU and ev.name.startswith("aten::") and ev.cuda_time_total > 0 \
            and not any(c.cuda_time_total > 0 and c.name.startswith("aten::") for c in ev.cpu_children):
        src = next((s for s in ev.stack if "npf_gwwaveform_b200/" in s or "tests/" in s), ev.stack[0] if ev.stack else "?")
        cnt[(ev.name, str(ev.input_shapes)[:60], src.split("/root/repo/")[-1][:90])] += 1
for (n, sh, src), c in cnt.most_common(60):
    print(f"{c:4d}  {n:28s} {sh:60s} {src}")
