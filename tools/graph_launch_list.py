"""Workload for `ncu --profile-from-start off`: only REPLAYS of the captured training step are inside the profiled range
(cudaProfilerStart/Stop), so the launch list shows exactly which kernels one optimisation step consists of."""
import os
import sys

import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200.neuralproc import set_lazy_validation  # noqa: E402
from npf_gwwaveform_b200.training import FlatAdamTrainer  # noqa: E402
from util_models import build_product, chirp_batch  # noqa: E402

name, B = (sys.argv[1], int(sys.argv[2])) if len(sys.argv) > 2 else ("AttnCNP", 256)
npf.set_mode("fast")
set_lazy_validation(True)
torch.manual_seed(0)
model = build_product(name, 128).cuda().train()
tr = FlatAdamTrainer(model, npf.CNPFLoss().train(), lr=1e-3)
batch = {k: v.cuda() for k, v in chirp_batch(B, 512, 4096, 1234).items()}
for _ in range(3):
    tr.step(batch)
tr.capture(batch)
for _ in range(2):
    tr.step_graphed(batch)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
for _ in range(2):
    tr.step_graphed(batch)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
