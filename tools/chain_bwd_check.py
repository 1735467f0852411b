"""K1b check: npfk_linear_bwd_fused2 (two layers, one kernel) against two npfk_linear_bwd_fused launches and against
float64 torch, plus timing of both at the benchmark shape.  Run on the GPU box:  python tools/chain_bwd_check.py"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from npf_gwwaveform_b200._lib import LIB  # noqa: E402


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def round_tf32(t):
    out = t.clone()
    LIB.call("npfk_round_tf32", _p(t), _p(out), t.numel(), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return out


def run(M, masked0, with_db=True, seed=0, time_it=False, accumulate=False):
    g = torch.Generator(device="cuda").manual_seed(seed)
    G = torch.randn(M, 128, device="cuda", generator=g)
    H1 = torch.relu(torch.randn(M, 128, device="cuda", generator=g))
    H0 = torch.relu(torch.randn(M, 128, device="cuda", generator=g))
    W1 = round_tf32(torch.randn(128, 128, device="cuda", generator=g) / 11)
    W0 = round_tf32(torch.randn(128, 128, device="cuda", generator=g) / 11)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dx_init = torch.randn(M, 128, device="cuda", generator=g) if accumulate else None

    def layerwise():
        gm = torch.empty(M, 128, device="cuda")
        dx = dx_init.clone() if accumulate else torch.empty(M, 128, device="cuda")
        dw1, dw0 = torch.zeros(128, 128, device="cuda"), torch.zeros(128, 128, device="cuda")
        db1 = torch.zeros(128, device="cuda") if with_db else None
        db0 = torch.zeros(128, device="cuda") if with_db else None
        LIB.call("npfk_linear_bwd_fused", _p(G), 128, _p(H1), 128, _p(W1), 128, 1, _p(gm), 128, _p(dw1), 128, _p(db1),
                 M, 128, 128, 1, st)
        LIB.call("npfk_linear_bwd_fused", _p(gm), 128, _p(H0), 128, _p(W0), 128, int(masked0) | (2 if accumulate else 0),
                 _p(dx), 128, _p(dw0), 128, _p(db0), M, 128, 128, 1, st)
        return dx, dw1, db1, dw0, db0

    def fused2():
        dx = dx_init.clone() if accumulate else torch.empty(M, 128, device="cuda")
        dw1, dw0 = torch.zeros(128, 128, device="cuda"), torch.zeros(128, 128, device="cuda")
        db1 = torch.zeros(128, device="cuda") if with_db else None
        db0 = torch.zeros(128, device="cuda") if with_db else None
        LIB.call("npfk_linear_bwd_fused2", _p(G), 128, _p(H1), 128, _p(W1), 128, _p(H0), 128, _p(W0), 128,
                 int(masked0) | (2 if accumulate else 0), _p(dx), 128, _p(dw1), 128, _p(db1), _p(dw0), 128, _p(db0), M, st)
        return dx, dw1, db1, dw0, db0

    a = layerwise()
    b = fused2()
    torch.cuda.synchronize()
    abort = int(LIB.load().npfk_tc_abort_code())
    names = ["dx", "dw1", "db1", "dw0", "db0"]
    worst = 0.0
    line = []
    for n, x, y in zip(names, a, b):
        if x is None:
            continue
        err = float((x - y).abs().max() / (x.abs().max() + 1e-30))
        worst = max(worst, err)
        line.append(f"{n} {err:.2e}")
    # float64 reference of the whole pair
    Gd, H1d, H0d, W1d, W0d = (t.double() for t in (G, H1, H0, W1, W0))
    gm = (Gd @ W1d) * (H1d > 0)
    dxr = gm @ W0d
    if masked0:
        dxr = dxr * (H0d > 0)
    if accumulate:
        dxr = dxr + dx_init.double()
    e64 = float((b[0].double() - dxr).abs().max() / dxr.abs().max())
    e64w = float((b[3].double() - gm.t() @ H0d).abs().max() / (gm.t() @ H0d).abs().max())
    print(f"M={M} masked0={masked0} db={with_db} acc={accumulate}: vs layer-wise {'  '.join(line)} | vs fp64 dx {e64:.2e} dw0 {e64w:.2e}"
          f" | abort={abort}", flush=True)
    ok = worst < 2e-3 and e64 < 3e-3 and e64w < 3e-3 and abort == 0
    if time_it:
        for fn, nm in ((layerwise, "layer-wise (2 launches)"), (fused2, "fused2")):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                fn()
            e1.record()
            torch.cuda.synchronize()
            print(f"    {nm}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us (incl. zero fills)", flush=True)
    return ok


if __name__ == "__main__":
    ok = True
    for M, m0, db in ((128, True, True), (256, False, True), (300, True, False), (128 * 148 * 2 + 77, True, True),
                      (131072, True, True)):
        ok &= run(M, m0, db)
    ok &= run(4096, True, True, accumulate=True)
    ok &= run(1048576, True, True, time_it=True)
    print("OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)
