"""Context / target split of a batch that already lives in HBM.

Mirror of the reference's index selectors and ``CntxtTrgtGetter`` (npf/utils/datasplit.py:29-255) and of the collate
function that calls it (utils/data/dataloader.py:6-37).  Same constructors, call signatures and return values; the
differences are where the work happens:

* the reference draws the index matrix with numpy on the host for every batch (``indep_shuffle_``,
  npf/utils/helpers.py:82-96) and gathers on the CPU inside a DataLoader worker; here ``GetRandomIndcs`` draws the subset
  on the device (``npfk_random_subset``: partial Fisher-Yates per task, counter-based RNG) and ``select`` gathers with
  ``npfk_gather_rows``, so a training step needs no host work and no H2D copy beyond the raw batch;
* the NUMBER of points is still drawn on the host with ``random`` / ``numpy`` exactly like the reference
  (datasplit.py:106-126): it fixes tensor shapes, so it has to be known to the host anyway;
* random streams differ from numpy's (like every device RNG); the distribution (uniform subsets, independent per task)
  is the same.  Seeds: ``GetRandomIndcs(seed=...)`` or the global ``torch`` seed at construction time.

CPU tensors raise ``NpfkError``: there is no host fallback for this path either.
"""
import random

import numpy as np
import torch

from .. import kernels as K
from .._lib import LIB

__all__ = ["get_all_indcs", "GetRangeIndcs", "GetIndcsMerger", "GetRandomIndcs", "CntxtTrgtGetter", "cntxt_trgt_collate"]


def _ratio_to_int(percentage, max_val):
    """npf/utils/helpers.py:99-108 ratio_to_int: a value in [1, max_val] is a count (1 is ONE point, as in the
    reference, which tests the count range first), a value in [0, 1) a fraction of ``max_val``."""
    if 1 <= percentage <= max_val:
        out = percentage
    elif 0 <= percentage < 1:
        out = percentage * max_val
    else:
        raise ValueError(f"percentage={percentage} outside of [0,{max_val}].")
    return int(out)


def get_all_indcs(batch_size, n_possible_points, device=None):
    """All indices, the same for every task (datasplit.py:30-34)."""
    return torch.arange(n_possible_points, device=device).expand(batch_size, n_possible_points)


class GetRangeIndcs:
    """All indices of a range (datasplit.py:37-45)."""

    def __init__(self, arange):
        self.arange = arange

    def __call__(self, batch_size, n_possible_points, device=None):
        indcs = torch.arange(*self.arange, device=device)
        return indcs.expand(batch_size, len(indcs))


class GetIndcsMerger:
    """Concatenates the indices of several indexers (datasplit.py:48-57)."""

    def __init__(self, indexers):
        self.indexers = indexers

    def __call__(self, batch_size, n_possible_points, device=None):
        return torch.cat([ix(batch_size, n_possible_points, device=device) for ix in self.indexers], dim=1)


class GetRandomIndcs:
    """Random subset of indices per task (datasplit.py:60-144), drawn on the device.

    Parameters as in the reference: ``a``, ``b`` (min / max number of indices, fractions if < 1), ``is_batch_share``,
    ``range_indcs``, ``is_ensure_one``, ``is_beta_binomial``, ``proba_uniform``; plus ``seed`` for the device stream."""

    def __init__(self, a=0.1, b=0.5, is_batch_share=False, range_indcs=None, is_ensure_one=False,
                 is_beta_binomial=False, proba_uniform=0, seed=None):
        self.a, self.b = a, b
        self.is_batch_share = is_batch_share
        self.range_indcs = range_indcs
        self.is_ensure_one = is_ensure_one
        self.is_beta_binomial = is_beta_binomial
        self.proba_uniform = proba_uniform
        self.seed = int(torch.initial_seed() if seed is None else seed) & 0x7FFFFFFFFFFFFFFF
        self.n_calls = 0

    def reseed(self, seed):
        """restart the device stream (eval_loglike does this so that every evaluation sees the same context sets,
        like ``set_seed`` in utils/evaluate.py:11)"""
        self.seed = int(seed) & 0x7FFFFFFFFFFFFFFF
        self.n_calls = 0

    def n_indcs(self, n_possible_points):
        """number of indices of this call: host-side draw, same rules as datasplit.py:106-126"""
        if np.random.uniform(size=1) < self.proba_uniform:
            n = random.randint(0, n_possible_points)
        elif self.is_beta_binomial:
            from scipy.stats import betabinom

            n = int(betabinom(n_possible_points, self.a, self.b).rvs())
        else:
            n = random.randint(_ratio_to_int(self.a, n_possible_points), _ratio_to_int(self.b, n_possible_points))
        if self.is_ensure_one and n < 1:
            n = 1
        return n

    def __call__(self, batch_size, n_possible_points, device=None):
        device = torch.device("cuda" if device is None else device)
        if device.type != "cuda":
            raise K.NpfkError("GetRandomIndcs draws its indices on the GPU: pass device='cuda' (no host path)")
        lo = 0
        if self.range_indcs is not None:
            lo = int(self.range_indcs[0])
            n_possible_points = int(self.range_indcs[1] - self.range_indcs[0])
        n = self.n_indcs(n_possible_points)
        indcs = torch.empty(batch_size, n, dtype=torch.int64, device=device)
        self.n_calls += 1
        seed = (self.seed + 0x9E3779B97F4A7C15 * self.n_calls) & 0x7FFFFFFFFFFFFFFF
        with torch.cuda.device(device):
            LIB.call("npfk_random_subset", indcs.data_ptr(), batch_size, n_possible_points, n, lo, seed,
                     int(self.is_batch_share), torch.cuda.current_stream().cuda_stream)
        return indcs


class CntxtTrgtGetter:
    """Splits ``X[B,N,x_dim]``, ``y[B,N,y_dim]`` into context and target points by indices (datasplit.py:147-255)."""

    def __init__(self, contexts_getter=None, targets_getter=get_all_indcs, is_add_cntxts_to_trgts=False):
        self.contexts_getter = GetRandomIndcs() if contexts_getter is None else contexts_getter
        self.targets_getter = targets_getter
        self.is_add_cntxts_to_trgts = is_add_cntxts_to_trgts

    def __call__(self, X, y=None, context_indcs=None, target_indcs=None, is_return_indcs=False):
        K.require_cuda(X, y)
        batch_size, num_points = self.getter_inputs(X)
        # indices handed in by the caller are checked synchronously (IndexError here, as torch.gather in the reference);
        # indices drawn by the getters on the device are checked without a host synchronisation (raised one call late)
        check_c = "sync" if context_indcs is not None else "async"
        check_t = "sync" if target_indcs is not None else "async"
        if context_indcs is None:
            context_indcs = self.contexts_getter(batch_size, num_points, device=X.device)
        all_targets = target_indcs is None and self.targets_getter is get_all_indcs and not self.is_add_cntxts_to_trgts
        if target_indcs is None:
            target_indcs = self.targets_getter(batch_size, num_points, device=X.device)
        if self.is_add_cntxts_to_trgts:
            target_indcs = self.add_cntxts_to_trgts(num_points, target_indcs, context_indcs)
        X_pre_cntxt = self.preprocess_context(X)
        if is_return_indcs:
            return context_indcs, X_pre_cntxt, target_indcs, X
        X_cntxt, Y_cntxt = self.select(X_pre_cntxt, y, context_indcs, check=check_c)
        if all_targets:        # every point is a target: the gather would be an identity copy of two tensors
            X_trgt, Y_trgt = X, y
        else:
            X_trgt, Y_trgt = self.select(X, y, target_indcs, check=check_t)
        return X_cntxt, Y_cntxt, X_trgt, Y_trgt

    def preprocess_context(self, X):
        return X

    def add_cntxts_to_trgts(self, num_points, target_indcs, context_indcs):
        target_indcs = torch.cat([target_indcs.to(context_indcs.device), context_indcs], dim=-1)
        return target_indcs[:, :num_points]

    def getter_inputs(self, X):
        batch_size, num_points, _ = X.shape
        return batch_size, num_points

    def select(self, X, y, indcs, check="sync"):
        """(X[b, indcs[b]], y[b, indcs[b]]) as fresh contiguous tensors; an index outside [0, N) raises IndexError
        (``check``: see ``_gather``; out-of-range rows are written as zeros, never read out of bounds)."""
        return _gather(X, indcs, check), _gather(y, indcs, check)


_pending_bad = []      # (event, pinned host flag, N) of gathers whose out-of-range flag has not been looked at yet


def _poll_bad(block: bool = False):
    """raise IndexError for an earlier gather whose index flag has arrived (never waits unless ``block``)"""
    while _pending_bad and (block or _pending_bad[0][0].query()):
        ev, host, N = _pending_bad.pop(0)
        if block:
            ev.synchronize()
        if int(host.item()):
            _pending_bad.clear()
            raise IndexError(f"select: index out of range for {N} points (detected on an earlier gather)")


def _gather(src, indcs, check: str = "async"):
    """``check``: "async" (default) copies the kernel's out-of-range flag to pinned host memory without waiting and
    raises from a LATER call once it has arrived (no host synchronisation on the training path; the kernel clamps, so
    nothing is read out of bounds); "sync" waits and raises here, as the reference's torch.gather would."""
    K.require_cuda(src)
    B, N, D = src.shape
    src = src.contiguous().float()
    indcs = indcs.to(device=src.device, dtype=torch.int64)
    if indcs.stride(0) == 0 or not indcs.is_contiguous():      # expanded (batch-shared) indices
        indcs = indcs.expand(B, indcs.shape[-1]).contiguous()
    n = indcs.shape[1]
    dst = torch.empty(B, n, D, device=src.device, dtype=torch.float32)
    bad = torch.zeros(1, dtype=torch.int32, device=src.device)
    LIB.call("npfk_gather_rows", src.data_ptr(), indcs.data_ptr(), dst.data_ptr(), B, N, n, D, bad.data_ptr(),
             torch.cuda.current_stream().cuda_stream)
    if not torch.cuda.is_current_stream_capturing() and n > 0:
        if check == "sync":
            if int(bad.item()):
                raise IndexError(f"select: index out of range for {N} points")
        else:
            _poll_bad()
            host = torch.empty(1, dtype=torch.int32).pin_memory()
            host.copy_(bad, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            _pending_bad.append((ev, host, N))
            if len(_pending_bad) > 64:
                _poll_bad(block=True)
    return dst


def cntxt_trgt_collate(get_cntxt_trgt, is_duplicate_batch=False, device="cuda", **kwargs):
    """collate_fn (utils/data/dataloader.py:6-37): default-collates the samples, moves X, y to ``device`` and splits
    them there.  Returns ``(inputs dict, Y_trgt)`` like the reference."""

    def mycollate(batch):
        collated = torch.utils.data.dataloader.default_collate(batch)
        X = collated[0].to(device, non_blocking=True)
        y = collated[1].to(device, non_blocking=True)
        if is_duplicate_batch:
            X = torch.cat([X, X], dim=0)
            y = torch.cat([y, y], dim=0)
        X_cntxt, Y_cntxt, X_trgt, Y_trgt = get_cntxt_trgt(X, y, **kwargs)
        inputs = dict(X_cntxt=X_cntxt, Y_cntxt=Y_cntxt, X_trgt=X_trgt, Y_trgt=Y_trgt)
        return inputs, Y_trgt

    mycollate.get_cntxt_trgt = get_cntxt_trgt
    return mycollate
