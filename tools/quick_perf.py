"""Developer probe: time forward+loss+backward of a model at a given shape with CUDA events (not the bench)."""
import os
import sys
import time

import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200.neuralproc import set_lazy_validation  # noqa: E402
from util_models import build_product, chirp_batch  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "AttnCNP"
B, Nc, Nt = (int(a) for a in sys.argv[2:5]) if len(sys.argv) > 4 else (32, 512, 4096)
nz = int(sys.argv[5]) if len(sys.argv) > 5 else 1
set_lazy_validation(True)
npf.set_mode(os.environ.get("NPFK_MODE", "fast"))
torch.manual_seed(0)
m = build_product(name, 128, nz=nz).cuda().train()
inp = {k: v.cuda() for k, v in chirp_batch(B, Nc, Nt, 0).items()}
crit = (npf.NLLLossLNPF if name == "ConvLNP" else npf.ELBOLossLNPF if "LNP" in name else npf.CNPFLoss)().train()


def step():
    for p in m.parameters():
        p.grad = None
    loss = crit(m(**inp), inp["Y_trgt"])
    loss.backward()
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
n0 = npf.launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.time()
e0.record()
K = 5
for _ in range(K):
    loss = step()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
print(f"{name} B={B} Nc={Nc} Nt={Nt} nz={nz}: {ms:.2f} ms/step  {B * Nt / ms * 1e3 / 1e6:.2f} M target-pts/s  "
      f"launches/step={(npf.launch_count() - n0) / K:.0f} wall={(time.time() - t0) / K * 1e3:.1f} ms loss={float(loss):.4f} "
      f"mem={torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")
