"""npfk_bn_stats_static: the batch-statistics kernel with its scratch in a static __device__ array of the library
(include/npfk.h), for callers that cannot allocate.  Checked through the C-ABI (ctypes, raw pointers) against
npfk_bn_stats with caller scratch, on consecutive calls that share the array.  Collected last on purpose
(``zz``)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-6      # the two entry points run the same kernel; only the order of the double atomics differs


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _both(lib, x, momentum, running):
    """(stats, running_mean, running_var, num_batches) from npfk_bn_stats and from npfk_bn_stats_static."""
    rows, C = x.shape
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = []
    for static in (False, True):
        stats = torch.full((2 * C,), float("nan"), device=x.device)
        rm = torch.linspace(-1, 1, C, device=x.device) if running else None
        rv = torch.linspace(0.5, 2, C, device=x.device) if running else None
        nb = torch.tensor([7], dtype=torch.int64, device=x.device) if running else None
        if static:
            lib.call("npfk_bn_stats_static", _p(x), rows, C, _p(stats), _p(rm), _p(rv), momentum, _p(nb), st)
        else:
            scratch = torch.empty(2 * C + 1, dtype=torch.float64, device=x.device)
            lib.call("npfk_bn_stats", _p(x), rows, C, _p(scratch), _p(stats), _p(rm), _p(rv), momentum, _p(nb), st)
        outs.append((stats, rm, rv, nb))
    torch.cuda.synchronize()
    return outs


def _check(outs, x):
    (s0, rm0, rv0, nb0), (s1, rm1, rv1, nb1) = outs
    C = x.shape[1]
    assert torch.isfinite(s1).all()
    torch.testing.assert_close(s1, s0, rtol=TOL, atol=TOL)
    # and against the definition, so a pair of equally wrong answers cannot pass
    xd = x.double()
    torch.testing.assert_close(s1[:C].double(), xd.mean(0), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(s1[C:].double(), xd.var(0, unbiased=False), rtol=1e-5, atol=1e-6)
    if rm0 is not None:
        torch.testing.assert_close(rm1, rm0, rtol=TOL, atol=TOL)
        torch.testing.assert_close(rv1, rv0, rtol=TOL, atol=TOL)
        assert int(nb1) == int(nb0) == 8


def test_bn_stats_static_matches_bn_stats_on_consecutive_calls():
    from npf_gwwaveform_b200._lib import LIB

    g = torch.Generator(device="cuda").manual_seed(11)
    # consecutive calls share the array: wide then narrow then wide again, so stale sums / a stale ticket would show
    shapes = [(4096, 128, True), (777, 50, False), (131, 1023, True), (4096, 128, True), (64, 4096, False)]
    for rows, C, running in shapes:
        x = torch.randn(rows, C, device="cuda", generator=g) * 3 + 1.5
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        stats = torch.empty(2 * C, device="cuda")
        held = torch.cuda.memory_allocated()
        LIB.call("npfk_bn_stats_static", _p(x), rows, C, _p(stats), None, None, 0.0, None, st)
        assert torch.cuda.memory_allocated() == held      # nothing allocated through torch by the call
        _check(_both(LIB, x, 0.1, running), x)
    # back-to-back on one stream without a host sync in between: stream order serialises the shared array
    xs = [torch.randn(2048, 256, device="cuda", generator=g) + i for i in range(3)]
    outs = [torch.empty(512, device="cuda") for _ in xs]
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for x, o in zip(xs, outs):
        LIB.call("npfk_bn_stats_static", _p(x), 2048, 256, _p(o), None, None, 0.0, None, st)
    torch.cuda.synchronize()
    for x, o in zip(xs, outs):
        torch.testing.assert_close(o[:256].double(), x.double().mean(0), rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(o[256:].double(), x.double().var(0, unbiased=False), rtol=1e-5, atol=1e-6)


def test_bn_stats_static_rejects_sizes_beyond_the_array():
    from npf_gwwaveform_b200._lib import LIB, NpfkError

    C = 4097
    x = torch.randn(8, C, device="cuda")
    stats = torch.empty(2 * C, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    with pytest.raises(NpfkError, match="4096"):
        LIB.call("npfk_bn_stats_static", _p(x), 8, C, _p(stats), None, None, 0.0, None, st)
    # the refusal leaves the library usable
    C = 4096
    x = torch.randn(8, C, device="cuda")
    stats = torch.empty(2 * C, device="cuda")
    LIB.call("npfk_bn_stats_static", _p(x), 8, C, _p(stats), None, None, 0.0, None, st)
    torch.cuda.synchronize()
    torch.testing.assert_close(stats[:C].double(), x.double().mean(0), rtol=1e-5, atol=1e-6)
