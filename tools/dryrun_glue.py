"""Developer aid (no GPU needed): run the Python glue of every model with the kernel launches stubbed out.

Kernel calls still go through ctypes (argument counts and types are checked against include/npfk.h) but their
CUDA launch errors are ignored, so outputs are garbage; only shapes / autograd wiring are exercised.
Never used by the package, the tests or the bench.
"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import _lib, kernels  # noqa: E402
from util_models import build_product, chirp_batch  # noqa: E402

kernels.require_cuda = lambda *a: None
kernels._st = lambda: None
_orig = _lib.LIB.call
calls = {}


def call(name, *args):
    calls[name] = calls.get(name, 0) + 1
    dll = _lib.LIB.load()
    getattr(dll, name)(*args)          # ctypes validates the argument list; the CUDA error is ignored


_lib.LIB.call = call
kernels.LIB.call = call
import npf_gwwaveform_b200.neuralproc.family as fam  # noqa: E402

# the range flag is garbage in a dry run
fam.NeuralProcessFamily._validate_inputs = lambda self, *a: None

for name, loss, nz in [("CNP", npf.CNPFLoss, 1), ("AttnCNP", npf.CNPFLoss, 1), ("ConvCNP", npf.CNPFLoss, 1),
                       ("ConvCNP_unet", npf.CNPFLoss, 1), ("ConvCNP_unshared", npf.CNPFLoss, 1), ("LNP", npf.ELBOLossLNPF, 3),
                       ("AttnLNP", npf.ELBOLossLNPF, 3), ("AttnLNP", npf.NLLLossLNPF, 2)]:
    for over in ([{}] if name != "LNP" else [{}, {"encoded_path": "both"}]):
        m = build_product(name, 32, nz=nz, density=16, kernel_size=5, **over)
        inp = chirp_batch(2, 5, 12, 0)
        for mode in ("train", "eval"):
            m.train(mode == "train")
            crit = loss()
            crit.train(mode == "train")
            if mode == "train":
                out = m(**inp)
            else:
                out = m(inp["X_cntxt"], inp["Y_cntxt"], inp["X_trgt"])
            p = out[0]
            assert tuple(p.base_dist.loc.shape) == (nz, 2, 12, 1), p.base_dist.loc.shape
            l = crit(out, inp["Y_trgt"])
            assert l.dim() == 0
            if mode == "train":
                l.backward()
                missing = [k for k, v in m.named_parameters() if v.grad is None]
                assert not missing, (name, missing)
                bad = [k for k, v in m.named_parameters() if v.grad.shape != v.shape]
                assert not bad, bad
        # empty context
        m.eval()
        out = m(torch.zeros(2, 0, 1), torch.zeros(2, 0, 1), inp["X_trgt"])
        print("ok", name, over, loss.__name__)
print(sorted(calls.items()))
