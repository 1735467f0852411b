"""Prediction helpers: the reference's ``SamplePredictor`` (npf/utils/predict.py:8-24) and the sharded batched-inference
driver for BASELINE.json configs[4] (1M tasks of 16k-sample waveforms over 8 GPUs, no communication).

``predict_sharded`` is replicas only (inference row): rank r of W takes the tasks
``shard_range(n_tasks, r, W)``, walks them in device-resident chunks, and keeps what it computes ON THE DEVICE.  The
predictive mean/std of one chunk are 8 B per target (131 GB for the whole sweep), so they are handed to a ``reduce``
callback that turns them into a small per-task summary (default: mean predictive mean, mean predictive std, and the
log-likelihood per task when targets are supplied); only the summaries are concatenated.  The next chunk's inputs are
copied host -> device on a side stream from pinned memory while the current chunk computes (double buffering), so a host
data source is overlapped as a DataLoader with pinned memory would be.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Iterable, Iterator, Optional, Tuple

import torch

__all__ = ["SamplePredictor", "predict_sharded", "iter_chunks", "default_summary"]


class SamplePredictor:
    """Multi-sample prediction with a trained model; ``is_dist`` returns the predictive distribution instead of its
    mean (reference npf/utils/predict.py:8-24)."""

    def __init__(self, model, is_dist=False):
        self.model = model
        self.is_dist = is_dist

    def __call__(self, *args):
        p_y_pred, *_ = self.model(*args)
        if self.is_dist:
            return p_y_pred
        return p_y_pred.base_dist.loc.detach()


def default_summary(p_yCc, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
    """[B, 2] (or [B, 3] with targets): mean predictive mean, mean predictive std (over z-samples and targets) and the
    per-task log-likelihood log(1/n_z sum_z prod_t p(y_t | z)) -- what the reference's evaluation reads from a
    prediction (utils/visualize/visualize_1d.py:280-306 plots loc/scale; utils/evaluate.py:9-28 the log-likelihood)."""
    loc, scale = p_yCc.base_dist.loc, p_yCc.base_dist.scale              # [n_z, B, Nt, 1]
    cols = [loc.mean(dim=(0, 2, 3)), scale.mean(dim=(0, 2, 3))]
    y = batch.get("Y_trgt")
    if y is not None:
        from ..losses import sum_log_prob

        lp = sum_log_prob(p_yCc, y)                                        # [n_z, B]
        cols.append(torch.logsumexp(lp, 0) - math.log(lp.shape[0]))
    return torch.stack(cols, dim=1)


def iter_chunks(tasks: Dict[str, torch.Tensor], lo: int, hi: int, chunk: int) -> Iterator[Dict[str, torch.Tensor]]:
    """views of ``tasks`` (dict of [n_tasks, ...] tensors, host or device) over [lo, hi) in pieces of ``chunk`` tasks"""
    for s in range(lo, hi, chunk):
        e = min(hi, s + chunk)
        yield {k: v[s:e] for k, v in tasks.items()}


def predict_sharded(model, tasks, n_tasks: Optional[int] = None, chunk: int = 64, rank: Optional[int] = None,
                    world_size: Optional[int] = None, reduce: Callable = default_summary,
                    device: Optional[torch.device] = None, cuda_graph: bool = True) -> Tuple[torch.Tensor, Tuple[int, int]]:
    """Batched inference over this rank's shard of a task set.

    tasks: a dict of tensors ``X_cntxt[n,Nc,1], Y_cntxt[n,Nc,1], X_trgt[n,Nt,1]`` (optionally ``Y_trgt``), host or
           device; or a callable ``tasks(lo, hi) -> iterable of such dicts`` that yields the chunks of tasks [lo, hi)
           (a data source that synthesises / reads its shard itself; ``n_tasks`` is then required).
    ``cuda_graph``: the forward + ``reduce`` of a chunk is captured into a CUDA graph the second time a chunk shape is
    seen and replayed for every later chunk of that shape (the sweep is launch-bound otherwise: ~90 kernels for 4 ms of
    work); chunks of any other shape (the shard's tail) run eagerly.  Models that draw random numbers per call (latent
    paths) are not captured.
    Returns ``(summaries, (lo, hi))``: the ``reduce`` outputs of the shard's chunks concatenated along dim 0 (on the
    device), and the shard's task range.  Ranks do not exchange anything; gather the summaries yourself if you need them
    in one place (they are a few bytes per task).
    """
    from ..training import shard_range

    if rank is None or world_size is None:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            rank, world_size = dist.get_rank(), dist.get_world_size()
        else:
            rank, world_size = 0, 1
    if device is None:
        device = next(model.parameters()).device
    if device.type != "cuda":
        raise RuntimeError("predict_sharded runs on a CUDA device (the npf hot path has no CPU fallback)")
    if callable(tasks):
        if n_tasks is None:
            raise ValueError("n_tasks is required when `tasks` is a callable data source")
        lo, hi = shard_range(n_tasks, rank, world_size)
        chunks: Iterable[Dict[str, torch.Tensor]] = tasks(lo, hi)
    else:
        n = next(iter(tasks.values())).shape[0]
        n_tasks = n if n_tasks is None else min(n_tasks, n)
        lo, hi = shard_range(n_tasks, rank, world_size)
        chunks = iter_chunks(tasks, lo, hi, chunk)

    was_training = model.training
    model.eval()
    main = torch.cuda.current_stream(device)
    copy = torch.cuda.Stream(device)
    out = []

    def stage(batch):
        """host -> device on the copy stream (non-blocking for pinned tensors); device tensors pass through"""
        if all(v.is_cuda for v in batch.values()):
            return batch, None
        with torch.cuda.stream(copy):
            dev = {k: v.to(device, non_blocking=True) for k, v in batch.items()}
            ev = torch.cuda.Event()
            ev.record(copy)
        return dev, ev

    latent = getattr(model, "encoded_path", "deterministic") != "deterministic"
    # shapes -> (graph, static inputs, static output) once captured; None after the first eager pass.  Kept on the model
    # so that later calls (the next shard, the next sweep) replay instead of capturing again; the graph reads the weights
    # in place (and re-rounds them to tf32 on every replay), so it stays valid when they are updated
    from .. import kernels as _K

    cache = model.__dict__.setdefault("_npfk_predict_graphs", {})
    graphs = cache.setdefault((id(reduce), _K.get_mode()), {})

    def run(cur):
        # as the reference predicts (visualize_1d.py:280-306): no Y_trgt, so latent models sample z from q(z|context)
        p_yCc = model(cur["X_cntxt"], cur["Y_cntxt"], cur["X_trgt"])[0]
        return reduce(p_yCc, cur)

    def step(cur):
        key = tuple((k, tuple(v.shape)) for k, v in sorted(cur.items()))
        if not cuda_graph or latent:
            return run(cur)
        if key not in graphs:                       # first chunk of this shape: eager (also warms the kernels up)
            graphs[key] = None
            return run(cur)
        if graphs[key] is None:                     # second chunk: capture, then replay for this and all later ones
            static = {k: torch.empty_like(v) for k, v in cur.items()}
            torch.cuda.synchronize(device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                out_static = run(static)
            graphs[key] = (g, static, out_static)
        g, static, out_static = graphs[key]
        for k, v in cur.items():
            static[k].copy_(v, non_blocking=True)
        g.replay()
        return out_static.clone()

    try:
        it = iter(chunks)
        nxt = next(it, None)
        staged = stage(nxt) if nxt is not None else None
        with torch.no_grad():
            while staged is not None:
                cur, ev = staged
                nxt = next(it, None)
                staged = stage(nxt) if nxt is not None else None      # overlaps with the forward below
                if ev is not None:
                    main.wait_event(ev)
                    for v in cur.values():
                        v.record_stream(main)
                out.append(step(cur))
    finally:
        model.train(was_training)
    if not out:
        return torch.empty(0, device=device), (lo, hi)
    return torch.cat(out, dim=0), (lo, hi)
