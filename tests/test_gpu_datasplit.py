"""Device-side context / target split (npf_gwwaveform_b200/utils/datasplit.py) against the oracle's restatement of
CntxtTrgtGetter.select and against the properties GetRandomIndcs guarantees in the reference (datasplit.py:60-255):
n distinct in-range indices per task, independent across tasks, uniform over positions, deterministic per seed."""
import random

import numpy as np
import pytest
import torch

from oracle import npf_oracle as O

pytestmark = pytest.mark.gpu


def _xy(B, N, xd=1, yd=1, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(B, N, xd, generator=g) * 2 - 1, torch.randn(B, N, yd, generator=g)


@pytest.mark.parametrize("B,N,a,b", [(5, 100, 0.1, 0.5), (64, 4096, 512, 512), (3, 7, 7, 7), (2, 50, 0, 0), (1, 1, 1, 1)])
def test_random_subset_properties(B, N, a, b):
    from npf_gwwaveform_b200.utils.datasplit import GetRandomIndcs

    random.seed(1)
    np.random.seed(1)
    getter = GetRandomIndcs(a=a, b=b, seed=123)
    idx = getter(B, N, device="cuda").cpu()
    n = idx.shape[1]
    lo_n = a if a >= 1 or a == 0 else int(a * N)
    hi_n = b if b >= 1 or b == 0 else int(b * N)
    assert lo_n <= n <= hi_n
    assert idx.dtype == torch.int64 and idx.shape[0] == B
    if n:
        assert int(idx.min()) >= 0 and int(idx.max()) < N
        for r in idx:
            assert len(set(r.tolist())) == n                      # without replacement
    if B > 1 and 1 < n < N:
        assert not torch.equal(idx[0], idx[1])                    # independent per task
    if n == N and n:
        assert torch.equal(idx.sort(dim=1).values, torch.arange(N).expand(B, N))   # a full permutation
    # same seed and call count -> same draw; next call -> a different one
    random.seed(1)
    np.random.seed(1)
    again = GetRandomIndcs(a=a, b=b, seed=123)(B, N, device="cuda").cpu()
    assert torch.equal(idx, again)


def test_random_subset_is_uniform():
    """each position is selected with probability n/N (512 of 4096 over 2048 tasks: 256 +- 5 sigma per position)"""
    from npf_gwwaveform_b200.utils.datasplit import GetRandomIndcs

    B, N, n = 2048, 4096, 512
    idx = GetRandomIndcs(a=n, b=n, seed=7)(B, N, device="cuda")
    counts = torch.bincount(idx.reshape(-1), minlength=N).float().cpu()
    p = n / N
    sigma = (B * p * (1 - p)) ** 0.5
    assert float((counts - B * p).abs().max()) < 5.5 * sigma
    # and the ORDER is uniform too: the first selected index has mean (N-1)/2
    first = idx[:, 0].float().mean().item()
    assert abs(first - (N - 1) / 2) < 5 * (N / 12 ** 0.5) / B ** 0.5


def test_batch_share_and_range():
    from npf_gwwaveform_b200.utils.datasplit import GetRandomIndcs

    idx = GetRandomIndcs(a=10, b=10, is_batch_share=True, seed=3)(6, 100, device="cuda").cpu()
    assert all(torch.equal(idx[0], r) for r in idx) and len(set(idx[0].tolist())) == 10
    idx = GetRandomIndcs(a=5, b=5, range_indcs=(40, 60), seed=3)(4, 100, device="cuda").cpu()
    assert int(idx.min()) >= 40 and int(idx.max()) < 60


@pytest.mark.parametrize("B,N,xd,yd,n", [(4, 64, 1, 1, 20), (3, 33, 2, 3, 33), (2, 10, 1, 1, 0)])
def test_select_matches_reference_gather(B, N, xd, yd, n):
    from npf_gwwaveform_b200.utils.datasplit import CntxtTrgtGetter

    X, y = _xy(B, N, xd, yd)
    g = torch.Generator().manual_seed(9)
    idx = torch.stack([torch.randperm(N, generator=g)[:n] for _ in range(B)]) if n else torch.zeros(B, 0, dtype=torch.int64)
    xr, yr = O.select_points(X, y, idx)
    xs, ys = CntxtTrgtGetter().select(X.cuda(), y.cuda(), idx.cuda())
    assert torch.equal(xs.cpu(), xr) and torch.equal(ys.cpu(), yr)        # a gather is exact


def test_getter_end_to_end_and_collate():
    from npf_gwwaveform_b200.utils.datasplit import (CntxtTrgtGetter, GetRandomIndcs, GetRangeIndcs, cntxt_trgt_collate,
                                                     get_all_indcs)

    B, N = 6, 128
    X, y = _xy(B, N)
    Xd, yd_ = X.cuda(), y.cuda()
    getter = CntxtTrgtGetter(contexts_getter=GetRandomIndcs(a=0.1, b=0.5, seed=11), targets_getter=get_all_indcs)
    ci, _, ti, _ = getter(Xd, yd_, is_return_indcs=True)
    Xc, Yc, Xt, Yt = getter(Xd, yd_, context_indcs=ci)
    xr, yr = O.select_points(X, y, ci.cpu())
    assert torch.equal(Xc.cpu(), xr) and torch.equal(Yc.cpu(), yr)
    assert torch.equal(Xt.cpu(), X) and torch.equal(Yt.cpu(), y)
    assert torch.equal(ti.cpu(), torch.arange(N).expand(B, N))
    # targets = a range + the context points appended (is_add_cntxts_to_trgts, datasplit.py:229-237)
    getter2 = CntxtTrgtGetter(contexts_getter=GetRandomIndcs(a=8, b=8, seed=5), targets_getter=GetRangeIndcs((0, 100)),
                              is_add_cntxts_to_trgts=True)
    ci, _, ti, _ = getter2(Xd, yd_, is_return_indcs=True)
    assert torch.equal(ti.cpu(), O.add_cntxts_to_trgts(N, torch.arange(100).expand(B, 100), ci.cpu()))
    Xc, Yc, Xt, Yt = getter2(Xd, yd_, context_indcs=ci, target_indcs=None)
    assert Xt.shape == (B, 108, 1)
    # out-of-range index -> IndexError (torch.gather raises on the host in the reference)
    with pytest.raises(IndexError):
        getter.select(Xd, yd_, torch.full((B, 2), N, device="cuda"))
    # collate: list of (x_i, y_i) samples -> (inputs dict on the device, Y_trgt)
    collate = cntxt_trgt_collate(getter)
    inputs, targets = collate([(X[i], y[i]) for i in range(B)])
    assert set(inputs) == {"X_cntxt", "Y_cntxt", "X_trgt", "Y_trgt"} and targets is inputs["Y_trgt"]
    assert inputs["X_cntxt"].is_cuda and inputs["X_cntxt"].shape[1] == inputs["Y_cntxt"].shape[1]


def test_split_feeds_the_model():
    """one training step of AttnCNP on a batch split on the device"""
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.utils.datasplit import CntxtTrgtGetter, GetRandomIndcs
    from util_models import build_product, chirp_batch

    inp = chirp_batch(4, 8, 96, 3)
    Xc, Yc, Xt, Yt = CntxtTrgtGetter(contexts_getter=GetRandomIndcs(a=16, b=32, seed=2))(inp["X_trgt"].cuda(), inp["Y_trgt"].cuda())
    m = build_product("AttnCNP", 128).cuda().train()
    loss = npf.CNPFLoss()(m(Xc, Yc, Xt, Yt), Yt)
    loss.backward()
    assert torch.isfinite(loss) and 16 <= Xc.shape[1] <= 32
