"""Developer probe: does a chain of fused Linear backward kernels run faster when the rows are walked in L2-sized chunks
(dX of layer l is re-read as G of layer l-1 while still in the 126 MB L2) than layer by layer over all rows?"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import npf_gwwaveform_b200 as npf  # noqa: E402
from npf_gwwaveform_b200 import kernels as Kn  # noqa: E402
from npf_gwwaveform_b200._lib import LIB  # noqa: E402

npf.set_mode("fast")
M, L = 256 * 4096, 4
dev = "cuda"
H = [torch.randn(M, 128, device=dev).relu() for _ in range(L)]          # saved inputs of the layers
W = [Kn.tf32_weight((torch.randn(128, 128, device=dev) * 0.1)) for _ in range(L)]
dW = [torch.zeros(128, 128, device=dev) for _ in range(L)]
dB = [torch.zeros(128, device=dev) for _ in range(L)]
G0 = torch.randn(M, 128, device=dev)
st = torch.cuda.current_stream().cuda_stream


def bwd(g, h, w, dx, dw, db, rows):
    LIB.call("npfk_linear_bwd_fused", g.data_ptr(), 128, h.data_ptr(), 128, w.data_ptr(), 128, 1, dx.data_ptr(), 128,
             dw.data_ptr(), 128, db.data_ptr(), rows, 128, 128, 1, torch.cuda.current_stream().cuda_stream)


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()          # replayed as a CUDA graph: the host cost of 64+ launches is not the question
    with torch.cuda.graph(graph):
        fn()
    fn = graph.replay
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


full = [torch.empty(M, 128, device=dev) for _ in range(2)]


def layerwise():
    g = G0
    for l in range(L - 1, -1, -1):
        dx = full[l & 1]
        bwd(g, H[l], W[l], dx, dW[l], dB[l], M)
        g = dx


print(f"layer by layer, {L} layers x {M} rows: {timeit(layerwise) * 1e3:.0f} us")
for nchunk in (4, 8, 16, 32, 64):
    R = M // nchunk
    pp = [torch.empty(R, 128, device=dev) for _ in range(2)]
    out = torch.empty(M, 128, device=dev)

    def chunked():
        for c in range(nchunk):
            g = G0[c * R:(c + 1) * R]
            for l in range(L - 1, -1, -1):
                dx = out[c * R:(c + 1) * R] if l == 0 else pp[l & 1]
                bwd(g, H[l][c * R:(c + 1) * R], W[l], dx, dW[l], dB[l], R)
                g = dx

    print(f"chunked x{nchunk} ({R} rows = {R * 512 / 1e6:.0f} MB per tensor): {timeit(chunked) * 1e3:.0f} us  abort={LIB.load().npfk_tc_abort_code()}")
