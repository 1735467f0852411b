"""Developer probe: which aten::add / add_ calls with large operands remain in one AttnCNP training step."""
import os, sys
import torch
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import npf_gwwaveform_b200 as npf
from npf_gwwaveform_b200 import kernels as Kn
from util_models import build_product, chirp_batch
npf.set_mode("fast")
torch.manual_seed(0)
m = build_product("AttnCNP", 128).cuda().train()
crit = npf.CNPFLoss().train()
b = {k: v.cuda() for k, v in chirp_batch(32, 512, 4096, 1).items()}
for _ in range(2):
    m.zero_grad(); crit(m(**b), b["Y_trgt"]).backward()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU], record_shapes=True, with_stack=True) as prof:
    m.zero_grad(); crit(m(**b), b["Y_trgt"]).backward()
for e in prof.events():
    if e.name in ("aten::add", "aten::add_") and e.input_shapes and e.input_shapes[0] and len(e.input_shapes[0]) >= 2:
        n = 1
        for d in e.input_shapes[0]: n *= d
        if n >= 1_000_000:
            st = [f for f in (e.stack or []) if "npf_gwwaveform" in f or "autograd" in f][:3]
            print(e.name, e.input_shapes[:2], st)
