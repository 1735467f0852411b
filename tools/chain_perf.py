"""Developer probe: npfk_mlp_chain_fwd alone at the C2 target-row count (CUDA events, fast mode).
training = every layer's activation written (what backward needs); inference = only the result leaves the chip."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402
from npf_gwwaveform_b200._lib import LIB  # noqa: E402

npf.set_mode("fast")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 256 * 4096
x = torch.randn(M, 128, device="cuda").relu()
res = torch.randn(M, 128, device="cuda")
y = torch.randn(M // 4096, 4096, 1, device="cuda")


def layers(n, n_out=128):
    out = []
    for i in range(n):
        N = n_out if i == n - 1 else 128
        out.append(((torch.randn(N, 128, device="cuda") * 0.1).requires_grad_(True), torch.zeros(N, device="cuda", requires_grad=True)))
    return out


def timeit(fn, iters=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


u = M * 512 / 1e9   # GB per activation tensor
for L, head, use_res in [(2, False, False), (2, False, True), (4, False, False), (6, False, False), (5, True, False)]:
    ls = layers(L, 2 if head else 128)
    for w, _ in ls:
        Kn.tf32_weight(w)
    nl = (y, 4096) if head else None
    Lsq = L - 1 if head else L
    for train in (True, False):
        xx = x.clone().requires_grad_(train)

        def run():
            with torch.set_grad_enabled(train):
                return Kn.mlp_chain(xx, ls, res=res if use_res else None, out_relu=not head, nll=nl)

        ms = timeit(run)
        n_out = Lsq if train else (0 if head else 1)
        gb = u * (1 + (1 if use_res else 0) + n_out)
        tf = 2.0 * M * 128 * 128 * Lsq / 1e12
        print(f"chain L={Lsq}{'+head+nll' if head else ''}{' +res' if use_res else ''} {'training ' if train else 'inference'}: "
              f"{ms * 1e3:8.1f} us  {gb / (ms * 1e-3):7.0f} GB/s (algorithmic)  {tf / (ms * 1e-3):6.1f} TF/s  abort={LIB.load().npfk_tc_abort_code()}",
              flush=True)
    Kn.set_chain_fusion(False)
    xx = x.clone().requires_grad_(True)
    ms = timeit(lambda: Kn.mlp_chain(xx, ls, res=res if use_res else None, out_relu=not head, nll=nl))
    print(f"   layer by layer (training): {ms * 1e3:8.1f} us")
    Kn.set_chain_fusion(True)
