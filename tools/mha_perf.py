"""Developer probe: time npfk_mha_fwd / npfk_mha_bwd alone (CUDA events) and check them against float64."""
import math
import os
import sys

import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402
from npf_gwwaveform_b200._lib import LIB  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "fast"
npf.set_mode(mode)


def check(B, Nq, Nk, H, D):
    g = torch.Generator().manual_seed(Nq + Nk)
    E = H * D
    q, k, v = (torch.randn(B, n, E, generator=g) for n in (Nq, Nk, Nk))
    do = torch.randn(B, Nq, E, generator=g)
    scale = 1 / math.sqrt(D)
    qs, ks, vs = (t.double().requires_grad_(True) for t in (q, k, v))
    qh, kh, vh = (t.view(B, -1, H, D).transpose(1, 2) for t in (qs, ks, vs))
    o = (torch.softmax(qh @ kh.transpose(-1, -2) * scale, -1) @ vh).transpose(1, 2).reshape(B, -1, E)
    o.backward(do.double())
    qd, kd, vd = (t.cuda().requires_grad_(True) for t in (q, k, v))
    od = Kn.mha(qd, kd, vd, H, scale)
    od.backward(do.cuda())
    torch.cuda.synchronize()
    rel = lambda a, r: float((a.detach().cpu().double() - r).norm() / r.norm())
    print(f"check B{B} Nq{Nq} Nk{Nk} H{H} D{D}: out {rel(od, o):.2e} dq {rel(qd.grad, qs.grad):.2e} "
          f"dk {rel(kd.grad, ks.grad):.2e} dv {rel(vd.grad, vs.grad):.2e} abort={LIB.load().npfk_tc_abort_code()}", flush=True)


def bench(B, Nq, Nk, H, D, iters=5):
    E = H * D
    q, k, v, do = (torch.randn(B, n, E, device="cuda") for n in (Nq, Nk, Nk, Nq))
    q.requires_grad_(True); k.requires_grad_(True); v.requires_grad_(True)
    scale = 1 / math.sqrt(D)
    for _ in range(2):
        o = Kn.mha(q, k, v, H, scale)
        o.backward(do)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    tf = tb = 0.0
    for _ in range(iters):
        q.grad = k.grad = v.grad = None
        ev[0].record()
        o = Kn.mha(q, k, v, H, scale)
        ev[1].record()
        o.backward(do)
        ev[2].record()
        torch.cuda.synchronize()
        tf += ev[0].elapsed_time(ev[1]); tb += ev[1].elapsed_time(ev[2])
    tf /= iters; tb /= iters
    fl = 4.0 * B * H * Nq * Nk * D
    print(f"bench B{B} Nq{Nq} Nk{Nk} H{H} D{D} [{mode}]: fwd {tf:.3f} ms ({fl / tf / 1e9:.1f} TF/s)  bwd {tb:.3f} ms "
          f"({2.5 * fl / tb / 1e9:.1f} TF/s algorithmic) abort={LIB.load().npfk_tc_abort_code()}", flush=True)


for shp in [(1, 64, 128, 2, 16), (2, 300, 77, 8, 16), (1, 4096, 512, 8, 16), (3, 129, 600, 8, 16), (2, 257, 64, 4, 32),
            (2, 128, 1, 8, 16), (5, 1, 130, 2, 16)]:
    check(*shp)
bench(256, 4096, 512, 8, 16)
