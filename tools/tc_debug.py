"""Direct C-ABI calls of the tcgen05 GEMMs with the abort code printed after each (GPU box debugging aid)."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from npf_gwwaveform_b200._lib import LIB  # noqa: E402

P = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())  # noqa: E731
ST = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)  # noqa: E731


def abort():
    torch.cuda.synchronize()
    return LIB.load().npfk_tc_abort_code()


def err(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())


def main():
    torch.manual_seed(0)
    for (M, N, K) in [(128, 128, 128), (300, 128, 128), (4096, 256, 128), (1000, 64, 256)]:
        x = torch.randn(M, K, device="cuda")
        w = torch.randn(N, K, device="cuda") / K ** 0.5
        g = torch.randn(M, N, device="cuda")
        y = torch.full((M, N), 7.0, device="cuda")
        LIB.call("npfk_linear_fwd", P(x), K, P(w), K, None, None, 0, 0, P(y), N, M, N, K, 0, 1, ST())
        print(M, N, K, "fwd abort", abort(), "err", err(y, x @ w.t()))
        dx = torch.full((M, K), 7.0, device="cuda")
        LIB.call("npfk_linear_bwd_dx", P(g), N, P(w), K, None, 0, P(dx), K, M, N, K, 1, ST())
        print(M, N, K, "dx  abort", abort(), "err", err(dx, g @ w), dx[0, :4].tolist(), (g @ w)[0, :4].tolist())
        dw = torch.zeros(N, K, device="cuda")
        db = torch.zeros(N, device="cuda")
        LIB.call("npfk_linear_bwd_dw", P(g), N, P(x), K, P(dw), K, P(db), M, N, K, 1, ST())
        print(M, N, K, "dw  abort", abort(), "err", err(dw, g.t() @ x), "db err", err(db, g.sum(0)),
              dw[0, :4].tolist(), (g.t() @ x)[0, :4].tolist())


if __name__ == "__main__":
    main()


def via_autograd():
    import math

    from npf_gwwaveform_b200 import kernels as Kn

    Kn.set_mode("fast")
    for (M, N, K, relu) in [(300, 128, 128, True), (300, 128, 128, False), (4096, 128, 128, True)]:
        g = torch.Generator().manual_seed(M + N + K)
        x, w, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / math.sqrt(K), torch.randn(N, generator=g)
        dy = torch.randn(M, N, generator=g)
        xd, wd, bd = (t.cuda().requires_grad_(True) for t in (x, w, b))
        y = Kn.linear(xd, wd, bd, relu=relu)
        y.backward(dy.cuda())
        yr = xd.detach().double() @ wd.detach().double().t() + bd.detach().double()
        gg = dy.cuda().double() * ((yr > 0) if relu else 1.0)
        dxr = gg @ wd.detach().double()
        d = (xd.grad.double() - dxr).abs()
        i = int(d.argmax())
        print(M, N, K, relu, "abort", abort(), "dx err", float(d.max() / dxr.abs().max()), "at", divmod(i, K),
              "bad rows", (d.max(1).values > 1e-2).nonzero().flatten().tolist()[:20],
              "mask mismatches", int(((y > 0) != (yr > 0)).sum()))


if __name__ == "__main__":
    via_autograd()
