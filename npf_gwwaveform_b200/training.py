"""Data-parallel training step over task batches (one process per GPU, torch.distributed / NCCL).

The reference trains through skorch on one device (utils/train.py:243-263: Adam, lr 1e-3) and has no distributed
code.  Tasks are independent, so the batch is sharded across ranks; the only exchange per step is the
gradient all-reduce.  Design:

* all parameters live in ONE flat fp32 buffer (views), all gradients in one flat buffer, laid out in the order
  in which backward finishes them (measured on a calibration step);
* the flat gradient is cut into a few buckets; a bucket's NCCL all-reduce is issued from the
  post-accumulate-grad hook of its last parameter, so it travels over NVLink while backward is still running;
* Adam runs as one npfk kernel over the flat buffers after the last bucket has landed;
* ``capture`` records the whole step - forward, loss, backward, the per-bucket all-reduces (forked onto NCCL's stream
  in the middle of backward) and Adam (step index on the device) - into ONE CUDA graph; ``step_graphed`` replays it.

With world_size == 1 the same code runs without any collective.  BatchNorm statistics stay per replica
(SURVEY.md 8e default).
"""
from __future__ import annotations

import os
from typing import Callable, List, Optional

import torch
import torch.distributed as dist

from . import kernels as K

__all__ = ["FlatAdamTrainer", "shard_range"]


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous, balanced [lo, hi) slice of ``n_items`` independent tasks for ``rank`` (no communication)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class FlatAdamTrainer:
    """forward -> loss -> backward (with overlapped gradient all-reduce) -> Adam, on flat buffers."""

    def __init__(self, model: torch.nn.Module, criterion: Callable, lr: float = 1e-3, betas=(0.9, 0.999),
                 eps: float = 1e-8, n_buckets: int = 2, process_group=None, adam_impl: Optional[Callable] = None,
                 overlap: bool = True):
        self.model, self.criterion = model, criterion
        self.lr, self.betas, self.eps = lr, betas, eps
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.n_buckets = max(1, n_buckets)
        self.overlap = overlap and self.world > 1
        self.adam_impl = adam_impl or K.adam_step
        self.step_count = 0
        self.params: List[torch.nn.Parameter] = [p for p in model.parameters() if p.requires_grad]
        self._order: Optional[List[int]] = None          # readiness order measured on the first step
        self._fired: List[int] = []
        self._pending = []
        self._hooks = []
        self._build_flat(list(range(len(self.params))))
        self._install_hooks()

    # ------------------------------------------------------------------ flat storage
    def _build_flat(self, order: List[int]) -> None:
        dev = self.params[0].device
        # every parameter starts on a 128-byte boundary: the tensor-core weight-gradient kernels need 16-byte aligned
        # outputs to take gradients straight into this buffer (kernels.set_direct_grad); the padding stays zero
        pad = lambda n: (n + 31) & ~31  # noqa: E731
        total = sum(pad(p.numel()) for p in self.params)
        flat_p = torch.zeros(total, device=dev, dtype=torch.float32)
        flat_g = torch.zeros(total, device=dev, dtype=torch.float32)
        old_m = getattr(self, "exp_avg", None)
        self.offsets = {}
        off = 0
        with torch.no_grad():
            for i in order:
                p = self.params[i]
                n = p.numel()
                flat_p[off:off + n].copy_(p.detach().reshape(-1))
                p.data = flat_p[off:off + n].view_as(p)
                p.grad = flat_g[off:off + n].view_as(p)
                self.offsets[i] = (off, n)
                off += pad(n)
        self.flat_param, self.flat_grad = flat_p, flat_g
        if old_m is None:
            self.exp_avg = torch.zeros_like(flat_p)
            self.exp_avg_sq = torch.zeros_like(flat_p)
        # bucket boundaries: equal element counts, cut at parameter boundaries, in layout (= readiness) order
        self.bucket_of = {}
        self.bucket_slices = []
        target = total / self.n_buckets
        start, b = 0, 0
        for i in order:
            o, n = self.offsets[i]
            self.bucket_of[i] = b
            if (o + n) >= target * (b + 1) and b < self.n_buckets - 1:
                self.bucket_slices.append((start, o + n))
                start, b = o + n, b + 1
        self.bucket_slices.append((start, total))
        self.bucket_last = {}
        for i in order:                                   # last parameter (in readiness order) of each bucket
            self.bucket_last[self.bucket_of[i]] = i
        self.layout_order = order

    def _install_hooks(self) -> None:
        for i, p in enumerate(self.params):
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))

    def _make_hook(self, i: int):
        def hook(_param):
            self._fired.append(i)
            if getattr(self, "_event_cb", None) is not None:      # rehearsal / capture: the trainer decides what to launch
                self._event_cb(i)
                return
            if (self.overlap and self._order is not None and not getattr(self, "_capturing", False)
                    and self.bucket_last.get(self.bucket_of[i]) == i):
                self._launch_bucket(self.bucket_of[i])
        return hook

    def _launch_bucket(self, b: int) -> None:
        lo, hi = self.bucket_slices[b]
        K.join_side_stream()      # weight gradients issued on the side stream (kernels._fork_side) are part of the bucket
        if hi > lo:
            self._pending.append(dist.all_reduce(self.flat_grad[lo:hi], op=dist.ReduceOp.SUM, group=self.group,
                                                 async_op=True))

    def _start_arena(self) -> None:
        """one zero fill for every small "+=" buffer of the backward pass (kernels.zeros)"""
        if getattr(self, "_zero_arena", None) is None:
            self._zero_arena = torch.empty(1 << 21, device=self.flat_grad.device, dtype=torch.float32)      # 8 MB
        self._zero_arena.zero_()
        K.set_zero_arena(self._zero_arena)

    # ------------------------------------------------------------------ one optimisation step
    def step(self, batch: dict) -> torch.Tensor:
        """``batch`` holds X_cntxt, Y_cntxt, X_trgt, Y_trgt of this rank's shard; returns the local loss (device)."""
        self.flat_grad.zero_()
        self._fired.clear()
        self._pending.clear()
        self._start_arena()
        try:
            out = self.model(**batch)
            loss = self.criterion(out, batch["Y_trgt"])
            loss.backward()
        finally:
            K.set_zero_arena(None)
        for p_idx, p in enumerate(self.params):          # autograd may have replaced a .grad view (it must not)
            o, n = self.offsets[p_idx]
            if p.grad is None or p.grad.data_ptr() != self.flat_grad.data_ptr() + 4 * o:
                if p.grad is not None:
                    self.flat_grad[o:o + n].copy_(p.grad.reshape(-1))
                p.grad = self.flat_grad[o:o + n].view_as(p)
        if self.world > 1:
            if self.overlap and self._order is not None:
                for w in self._pending:
                    w.wait()
            else:
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
        self.step_count += 1
        self.adam_impl(self.flat_param, self.flat_grad, self.exp_avg, self.exp_avg_sq, self.lr, self.betas[0],
                       self.betas[1], self.eps, self.step_count, 1.0 / self.world)
        K.tf32_cache_clear()      # a custom adam_impl changes the weights without bumping tensor versions either
        if getattr(self, "_step_dev", None) is not None:  # keep the captured graphs' step counter in sync
            self._step_dev.fill_(self.step_count)
        if self._order is None:                           # calibration: re-lay the buffers in readiness order
            seen, order = set(), []
            for i in self._fired:
                if i not in seen:
                    seen.add(i)
                    order.append(i)
            order += [i for i in range(len(self.params)) if i not in seen]
            self._order = order
            m, v = self._gather_state(self.exp_avg), self._gather_state(self.exp_avg_sq)
            self._build_flat(order)
            self._scatter_state(self.exp_avg, m)
            self._scatter_state(self.exp_avg_sq, v)
        return loss.detach()

    # ------------------------------------------------------------------ CUDA-graph step
    def _rehearse(self) -> None:
        """One eager forward + backward in direct-gradient mode on the static batch that only records, in order, every
        moment a parameter's gradient is (about to be) written: ``K._grad_sink`` calls of the "+=" weight-gradient
        kernels and the post-accumulate hooks of everything else.  ``_bucket_ready[b]`` = index of the last such event
        of bucket b's parameters: while capturing, bucket b's all-reduce is recorded as soon as a later event arrives
        (the kernel behind a sink call is launched before the next call), so it forks off the main stream in the
        middle of backward."""
        index = {id(p): i for i, p in enumerate(self.params)}
        seq: List[int] = []
        self._event_cb = lambda i: seq.append(i)
        K.set_sink_callback(lambda p: seq.append(index[id(p)]) if id(p) in index else None)
        K.set_direct_grad(os.environ.get("NPFK_DIRECT_GRAD", "1") != "0")
        try:
            self.flat_grad.zero_()
            out = self.model(**self._static)
            self.criterion(out, self._static["Y_trgt"]).backward()
        finally:
            K.set_direct_grad(False)
            K.set_sink_callback(None)
            self._event_cb = None
        self.flat_grad.zero_()
        self._n_events = len(seq)
        last = {}
        for n, i in enumerate(seq):
            last[i] = n
        self._bucket_ready = []
        for b in range(len(self.bucket_slices)):
            members = [i for i in range(len(self.params)) if self.bucket_of[i] == b]
            # a parameter that produced no event (no gradient this step) never blocks its bucket
            self._bucket_ready.append(max([last[i] for i in members if i in last], default=-1))

    def _capture_event(self, _i=None) -> None:
        """called at every gradient event while capturing: record the all-reduce of every bucket that is complete"""
        n = self._event_no
        self._event_no += 1
        while self._next_bucket < len(self.bucket_slices) and self._bucket_ready[self._next_bucket] < n:
            self._launch_bucket(self._next_bucket)
            self._next_bucket += 1

    def capture(self, batch: dict, optimizer_in_graph: bool = True, micro_batches: int = 1) -> None:
        """Record one WHOLE optimisation step into a CUDA graph: zero the flat gradient, forward, loss, backward, the
        gradient all-reduce (world > 1: one NCCL all-reduce per bucket, forked onto NCCL's stream as soon as the
        bucket's last weight-gradient kernel has been recorded, joined before Adam) and the Adam update (step index
        kept on the device, ``npfk_adam_step_dev``).  Static copies of ``batch``'s tensors are made; every later batch
        must have the same shapes.  Call after at least two eager ``step`` calls (buffers laid out in readiness order,
        kernels' one-time host initialisation done).  ``optimizer_in_graph=False`` keeps the round-1 behaviour (one
        blocking all-reduce and an eager Adam launch after every replay).

        ``micro_batches > 1`` (gradient accumulation, e.g. BASELINE.json configs[3]: a fixed global batch of 128 tasks at
        1/2/4/8 GPUs): three more graphs are recorded -- first (zero + forward/backward), middle (forward/backward into
        the same flat gradient) and last (forward/backward + all-reduce + Adam with the gradient scaled by
        1 / (world * micro_batches)) -- and ``step_graphed(batch, first=, last=)`` replays the matching one."""
        assert self._order is not None, "run eager steps before capture()"
        torch.cuda.synchronize()
        self._static = {k: v.clone() for k, v in batch.items()}
        self._opt_in_graph = bool(optimizer_in_graph) and self.adam_impl is K.adam_step
        self._micro = max(1, int(micro_batches))
        if getattr(self, "_step_dev", None) is None:      # ONE device-side step counter shared by every captured graph
            self._step_dev = torch.zeros(1, dtype=torch.int32, device=self.flat_param.device)
        self._step_dev.fill_(self.step_count)
        if self._opt_in_graph and self.world > 1:
            self._rehearse()
            torch.cuda.synchronize()
        self._losses = {}
        self._graphs = {(True, True): self._capture_variant(True, True)}
        if self._micro > 1:
            assert self._opt_in_graph, "gradient accumulation over micro-batches needs the optimiser inside the graph"
            for key in ((True, False), (False, False), (False, True)):
                self._graphs[key] = self._capture_variant(*key)
        self._graph = self._graphs[(True, True)]
        self._static_loss = self._losses[(True, True)]
        for p_idx, p in enumerate(self.params):
            o, n = self.offsets[p_idx]
            assert p.grad is not None and p.grad.data_ptr() == self.flat_grad.data_ptr() + 4 * o, \
                "a gradient left the flat buffer during capture"
        self._flag_dev = getattr(self.model, "_pending_flag", None)
        if self._flag_dev is not None:
            self.model._pending_flag = None
            self._flag_host = torch.zeros(1, dtype=self._flag_dev.dtype).pin_memory()
            self._flag_event = None

    _CAPTURE_FIELDS = ("_graphs", "_losses", "_graph", "_static", "_static_loss", "_flag_dev", "_flag_host", "_flag_event",
                       "_micro", "_opt_in_graph", "_stage", "_stage_src")

    def export_capture(self) -> dict:
        """the state ``capture`` created (graphs, their static inputs / loss tensors, range-flag plumbing), so that a
        caller can keep one capture per input shape and switch between them with ``import_capture``"""
        return {k: getattr(self, k, None) for k in self._CAPTURE_FIELDS}

    def import_capture(self, state: dict) -> None:
        for k in self._CAPTURE_FIELDS:
            setattr(self, k, state.get(k))
        self._flag_event = None

    def _capture_variant(self, zero: bool, update: bool):
        """one CUDA graph of [zero the flat gradient] + forward + loss + backward [+ all-reduce + Adam]"""
        graph = torch.cuda.CUDAGraph()
        if getattr(self, "_pool", None) is None:          # the variants never run concurrently: one memory pool
            self._pool = torch.cuda.graph_pool_handle()
        self._capturing = True
        # weight-gradient kernels add straight into the flat gradient buffer (NPFK_DIRECT_GRAD=0: through autograd)
        K.set_direct_grad(os.environ.get("NPFK_DIRECT_GRAD", "1") != "0")
        self._pending.clear()
        self._event_no, self._next_bucket = 0, 0
        opt = self._opt_in_graph and update
        fork = opt and self.world > 1
        if fork:
            self._event_cb = self._capture_event
            K.set_sink_callback(self._capture_event)
        scale = 1.0 / (self.world * (self._micro if not (zero and update) else 1))
        try:
            with torch.cuda.graph(graph, pool=self._pool):
                if zero:
                    self.flat_grad.zero_()
                if K.get_mode() == K.MODE_FAST:       # one rounding kernel for every weight of the step
                    K.set_tf32_flat(self.flat_param)
                self._start_arena()
                out = self.model(**self._static)
                loss = self.criterion(out, self._static["Y_trgt"])
                loss.backward()
                static_loss = loss.detach()
                if opt:
                    if fork:
                        assert self._event_no == self._n_events, "the captured step differs from its rehearsal"
                        while self._next_bucket < len(self.bucket_slices):
                            self._launch_bucket(self._next_bucket)
                            self._next_bucket += 1
                        for w in self._pending:
                            w.wait()
                    K.adam_step_dev(self.flat_param, self.flat_grad, self.exp_avg, self.exp_avg_sq, self.lr,
                                    self.betas[0], self.betas[1], self.eps, self._step_dev, scale)
        finally:
            self._capturing = False
            self._event_cb = None
            self._pending.clear()
            K.set_direct_grad(False)
            K.set_sink_callback(None)
            K.set_tf32_flat(None)
            K.set_zero_arena(None)
        self._losses[(zero, update)] = static_loss
        return graph

    def prefetch(self, batch: dict) -> None:
        """Start the H2D copy of the NEXT step's batch (pinned host tensors) on a side stream into a device staging
        set, so that it overlaps the step running now; ``step_graphed(batch)`` with the same dict object then only
        does a device-to-device copy into the graph's static inputs.  One batch can be in flight."""
        if getattr(self, "_stage", None) is None:
            self._stage = {k: torch.empty_like(v) for k, v in self._static.items()}
            self._copy_stream = torch.cuda.Stream()
            self._stage_ready = torch.cuda.Event()
            self._stage_free = torch.cuda.Event()
            self._stage_free.record()
            self._stage_src = None
        self._copy_stream.wait_event(self._stage_free)        # the previous staged batch has been consumed
        with torch.cuda.stream(self._copy_stream):
            for k, v in batch.items():
                self._stage[k].copy_(v, non_blocking=True)
            self._stage_ready.record()
        self._stage_src = batch

    def step_graphed(self, batch: dict, first: bool = True, last: bool = True) -> torch.Tensor:
        """``first`` / ``last``: position of this micro-batch in the optimisation step (``capture(micro_batches=n)``);
        the default is a whole step.  One optimisation step replaying the captured graph (forward, loss, backward, bucketed all-reduce, Adam) after
        the H2D/D2D copy of the batch into the static buffers.  Returns the graph's static loss tensor: it is
        OVERWRITTEN by the next replay, so read or clone it before the next call.  The [-1,1] range check of the reference (base.py:241-247) is
        read back asynchronously and raised at most one step late."""
        if self._flag_dev is not None and self._flag_event is not None and self._flag_event.query():
            self._flag_event = None
            if int(self._flag_host.item()):
                raise ValueError("Features during training should be in [-1,1] (detected on an earlier step).")
        if getattr(self, "_stage_src", None) is batch:            # prefetched: wait for its H2D, then D2D
            torch.cuda.current_stream().wait_event(self._stage_ready)
            for k in batch:
                self._static[k].copy_(self._stage[k], non_blocking=True)
            self._stage_free.record()
            self._stage_src = None
        else:
            for k, v in batch.items():
                if v is not self._static[k]:
                    self._static[k].copy_(v, non_blocking=True)
        key = (bool(first), bool(last))
        self._graphs[key].replay()
        if not last:
            return self._losses[key]
        if self._flag_dev is not None and self._flag_event is None:
            self._flag_host.copy_(self._flag_dev, non_blocking=True)
            self._flag_event = torch.cuda.Event()
            self._flag_event.record()
        self.step_count += 1
        if not self._opt_in_graph:
            if self.world > 1:
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
            self.adam_impl(self.flat_param, self.flat_grad, self.exp_avg, self.exp_avg_sq, self.lr, self.betas[0],
                           self.betas[1], self.eps, self.step_count, 1.0 / self.world)
        K.tf32_cache_clear()      # the weights changed without bumping tensor versions
        return self._losses[key]

    def _gather_state(self, flat):
        return {i: flat[o:o + n].clone() for i, (o, n) in self.offsets.items()}

    def _scatter_state(self, flat, parts):
        for i, (o, n) in self.offsets.items():
            flat[o:o + n].copy_(parts[i])

    def close(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks.clear()
