"""autograd bindings of the npfk C-ABI (include/npfk.h).

Every function here takes and returns CUDA fp32 tensors and launches hand-written sm_100a kernels on
torch's current stream.  There is no fallback: CPU tensors or a missing libnpfk.so raise.
"""
from __future__ import annotations

import ctypes
import weakref
from typing import Optional

import torch
from torch.autograd import Function

from ._lib import LIB, NpfkError

MODE_FP32 = 0
MODE_FAST = 1

_state = {"mode": MODE_FP32}


def set_mode(mode) -> None:
    """Select the arithmetic mode: "fp32" (<=1e-4 parity mode) or "fast" (tensor-core mode)."""
    if isinstance(mode, str):
        mode = {"fp32": MODE_FP32, "fast": MODE_FAST}[mode.lower()]
    _state["mode"] = int(mode)


def get_mode() -> int:
    return _state["mode"]


# ------------------------------------------------------------------------------------------ helpers
def _p(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _st():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*ts) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise NpfkError(
                "npf_gwwaveform_b200 runs on CUDA (sm_100a) tensors only; got a CPU tensor. "
                "There is no CPU fallback for the npf hot path."
            )


def _c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """contiguous fp32 view/copy"""
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def set_direct_grad(flag: bool) -> None:
    """Direct-gradient mode (FlatAdamTrainer.capture turns it on while recording): the "+=" weight-gradient kernels add
    straight into ``param.grad`` (the trainer's zeroed flat buffer) and the autograd node returns None for that
    parameter, instead of zero-filling a fresh buffer that autograd then adds to ``param.grad`` (two tiny kernels per
    parameter per step).  No AccumulateGrad hook fires for such a parameter."""
    _state["direct_grad"] = bool(flag)


def set_sink_callback(cb) -> None:
    """``cb(param)`` is called each time a backward node is about to add a gradient straight into ``param.grad``
    (direct-gradient mode); the kernel doing so is launched before the next such call.  FlatAdamTrainer uses it to
    know when a gradient bucket is complete while the step is being captured."""
    _state["sink_cb"] = cb


def _grad_sink(p, shape=None):
    g = _grad_sink_buf(p, shape)
    if g is not None and _state.get("sink_cb") is not None:
        _state["sink_cb"](p)
    return g


def _grad_sink_buf(p, shape=None):
    if not _state.get("direct_grad") or not isinstance(p, torch.nn.Parameter) or p.grad is None:
        return None
    g = p.grad
    if g.dtype != torch.float32 or not g.is_contiguous() or not g.is_cuda:
        return None
    return g if shape is None else g.view(shape)


# ------------------------------------------------------------------------------------------ side stream for wgrads
# A weight gradient is a leaf of the backward graph: nothing downstream in the same pass reads it.  For the small,
# latency-bound GEMMs of the conv nets (6 k rows; ~10 us each where the data would stream in 1-2 us) the dgrad / wgrad
# pair of a layer is issued on two streams so that the wgrad overlaps the rest of the backward chain; an autograd engine
# callback joins the side stream before backward() returns (and inside a CUDA-graph capture the fork / join become graph
# edges).  Measured on the C3 step (ConvCNP-UNet, 32 tasks): 1.4341 ms with the fork, 1.4350 ms without -- the one-wave
# wgrad kernels leave no SMs for a concurrent dgrad -- so it is OFF by default; NPFK_WGRAD_STREAM=1 enables it.
import os as _os

_side = {"stream": None, "pending": False, "on": _os.environ.get("NPFK_WGRAD_STREAM", "0") == "1"}


def join_side_stream():
    """make the current stream wait for every wgrad issued on the side stream so far (FlatAdamTrainer calls this before
    it all-reduces a gradient bucket; the autograd engine callback calls it at the end of backward())"""
    _join_side()


def _join_side():
    if _side["pending"]:
        torch.cuda.current_stream().wait_stream(_side["stream"])
        _side["pending"] = False


def _fork_side(*tensors):
    """side stream ordered after everything issued so far on the current stream; the caller launches on it inside
    ``with torch.cuda.stream(side)``.  `tensors` are kept alive for the side stream (caching-allocator bookkeeping)."""
    if _side["stream"] is None or _side["stream"].device != torch.cuda.current_stream().device:
        _side["stream"] = torch.cuda.Stream()
    side = _side["stream"]
    side.wait_stream(torch.cuda.current_stream())
    for t in tensors:
        if t is not None:
            t.record_stream(side)
    if not _side["pending"]:
        _side["pending"] = True
        torch.autograd.Variable._execution_engine.queue_callback(_join_side)
    return side


# ------------------------------------------------------------------------------------------ zero arena
# "+=" outputs of the backward kernels (weight / bias / length-scale gradients, BatchNorm sums, dV of the SetConvs) need
# zero-filled buffers.  Under FlatAdamTrainer they are slices of ONE arena that the trainer zeroes once per step, instead
# of one fill kernel per buffer (23 fills per ConvCNP-UNet step).  Outside a trainer step `zeros` is torch.zeros.
_arena = {"buf": None, "off": 0}


def set_zero_arena(buf: Optional[torch.Tensor]) -> None:
    """``buf``: a float32 tensor that the caller has JUST zeroed on the current stream (or None to switch the arena off).
    Every slice handed out lives until the caller zeroes the arena again."""
    _arena["buf"], _arena["off"] = buf, 0


def zeros(*shape, device=None) -> torch.Tensor:
    n = 1
    for d in shape:
        n *= int(d)
    buf = _arena["buf"]
    if buf is not None and (device is None or buf.device == torch.device(device)):
        off = _arena["off"]
        end = off + ((n + 31) & ~31)            # 128-byte aligned slices
        if end <= buf.numel():
            _arena["off"] = end
            return buf[off:off + n].view(*shape)
    return torch.zeros(*shape, device=device, dtype=torch.float32)


def _rows_view(x: torch.Tensor):
    """(tensor, ld) such that x can be read as M rows of K elements with row stride ld (no copy when possible)."""
    K = x.shape[-1]
    if x.dtype != torch.float32:
        x = x.float()
    if x.is_contiguous():
        return x, K
    ok = x.stride(-1) == 1 and x.dim() >= 2 and x.stride(-2) >= K
    if ok:
        for i in range(x.dim() - 2):
            if x.shape[i] != 1 and x.stride(i) != x.stride(i + 1) * x.shape[i + 1]:
                ok = False
                break
    if ok:
        return x, x.stride(-2)
    return x.contiguous(), K


def _weight_view(w: torch.Tensor):
    if w.dtype == torch.float32 and w.stride(1) == 1 and w.stride(0) >= w.shape[1]:
        return w, w.stride(0)
    w = w.float().contiguous()
    return w, w.shape[1]


# ------------------------------------------------------------------------------------------ Linear
_tf32_cache = {}


def _tf32_evict(base_id):
    for k in [k for k in _tf32_cache if k[0] == base_id]:
        _tf32_cache.pop(k, None)


def tf32_cache_clear() -> None:
    """Forget every rounded weight copy.  Needed after weights were changed behind autograd's back (a kernel writing
    through raw pointers does not bump ``Tensor._version``): ``adam_step`` and ``FlatAdamTrainer`` call it."""
    _tf32_cache.clear()


def set_tf32_flat(src: Optional[torch.Tensor]) -> None:
    """``src`` = a flat fp32 buffer holding many weights (FlatAdamTrainer.flat_param): rounds ALL of it with one
    kernel, here, and until ``set_tf32_flat(None)`` ``tf32_weight`` hands out views of that one copy for every
    contiguous weight living inside ``src``.  Used while a training step is captured into a CUDA graph: one rounding
    launch per replay instead of one per weight."""
    if src is None:
        _state["tf32_flat"] = None
        return
    require_cuda(src)
    out = torch.empty_like(src)
    LIB.call("npfk_round_tf32", _p(src), _p(out), src.numel(), _st())
    _state["tf32_flat"] = (src.data_ptr(), src.numel(), out)


def tf32_weight(w: torch.Tensor) -> torch.Tensor:
    """Round-to-nearest tf32 copy of a weight (tcgen05.mma.kind::tf32 truncates its fp32 operands).

    Cached per parameter object (weak reference; views are keyed on their base) and parameter version, outside
    CUDA-graph capture; inside a capture the rounding kernel is always recorded so replays see current weights."""
    flat = _state.get("tf32_flat")
    if flat is not None and w.dtype == torch.float32 and w.is_contiguous():
        off = w.data_ptr() - flat[0]
        if off >= 0 and off % 4 == 0 and off // 4 + w.numel() <= flat[1]:
            return flat[2][off // 4:off // 4 + w.numel()].view(w.shape)
    base = w._base if w._base is not None else w
    wv, _ = _weight_view(w.detach())
    if not wv.is_contiguous():
        wv = wv.contiguous()
    capturing = torch.cuda.is_current_stream_capturing()
    # data_ptr / device in the key: `p.data = ...` or module.to() re-point a parameter without bumping its version
    key = (id(base), w.data_ptr(), str(w.device), w.storage_offset(), tuple(w.shape), tuple(w.stride()))
    if not capturing:
        hit = _tf32_cache.get(key)
        if hit is not None and hit[0]() is base and hit[1] == base._version:
            return hit[2]
    out = torch.empty_like(wv)
    LIB.call("npfk_round_tf32", _p(wv), _p(out), wv.numel(), _st())
    if not capturing:
        base_id = id(base)
        _tf32_cache[key] = (weakref.ref(base, lambda _r, i=base_id: _tf32_evict(i)), base._version, out)
    return out


class _LinearFn(Function):
    @staticmethod
    def forward(ctx, x, w, b, res, relu: bool, dx_stash=None):
        require_cuda(x, w, b, res)
        ctx.wb = (w, b)
        ctx.dx_stash = dx_stash
        ctx.w_shape = w.shape
        if w.dim() == 3:          # pointwise Conv1d weight [N, K, 1]
            w = w.reshape(w.shape[0], w.shape[1])
        N, K = w.shape
        lead = x.shape[:-1]
        xv, ldx = _rows_view(x)
        mode = _state["mode"]
        if mode == MODE_FAST:
            wv = tf32_weight(w)
            ldw = wv.shape[1]
        else:
            wv, ldw = _weight_view(w)
        M = 1
        for s in lead:
            M *= s
        bv = _c(b)
        resv = _c(res)
        res_rows = 0
        if resv is not None:
            res_rows = resv.numel() // N
            if res_rows == 0 or M % max(res_rows, 1) != 0:
                raise NpfkError(f"linear: residual rows {res_rows} do not tile M={M}")
        y = torch.empty(*lead, N, device=x.device, dtype=torch.float32)
        if M > 0:
            LIB.call("npfk_linear_fwd", _p(xv), ldx, _p(wv), ldw, _p(bv), _p(resv), N, res_rows, _p(y), N, M, N, K,
                     int(relu), mode, _st())
        ctx.relu = relu
        ctx.mode = mode
        ctx.dims = (M, N, K, ldx, ldw, res_rows)
        ctx.has_b = b is not None
        ctx.has_res = res is not None
        ctx.res_shape = None if res is None else res.shape
        ctx.save_for_backward(xv, wv, y if relu else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        xv, wv, y = ctx.saved_tensors
        M, N, K, ldx, ldw, res_rows = ctx.dims
        dy = _c(dy)
        st = _st()
        if ctx.relu and M > 0:
            g = torch.empty_like(dy)
            LIB.call("npfk_relu_bwd", _p(dy), N, _p(y), N, _p(g), N, M, N, st)
        else:
            g = dy
        dx = dw = db = dres = None
        need_w = ctx.needs_input_grad[1]
        need_b = ctx.has_b and ctx.needs_input_grad[2]
        # another gradient of x parked by a later node of the forward graph (GradStash): the fused tensor-core kernel
        # adds this layer's dx onto it (TMA reduce-add); other shapes / modes add it afterwards
        parked, flags = None, 0
        if ctx.dx_stash is not None and ctx.dx_stash.t is not None:
            parked, ctx.dx_stash.t = ctx.dx_stash.t, None
        if ctx.needs_input_grad[0]:
            fusable = (parked is not None and need_w and ctx.mode == MODE_FAST and N == 128 and K == 128 and M >= 128
                       and parked.is_contiguous() and parked.numel() == M * K and parked.data_ptr() % 16 == 0
                       and ldx % 4 == 0 and xv.data_ptr() % 16 == 0 and g.data_ptr() % 16 == 0)
            if fusable:
                dx, flags, parked = parked.view(*dy.shape[:-1], K), 2, None
            else:
                dx = torch.empty(*dy.shape[:-1], K, device=dy.device, dtype=torch.float32)
        sink_w = _grad_sink(ctx.wb[0], (N, K)) if need_w else None
        sink_b = _grad_sink(ctx.wb[1]) if need_b else None
        if need_w or need_b:
            dw = sink_w if sink_w is not None else torch.zeros(N, K, device=dy.device, dtype=torch.float32)
            if need_b:
                db = sink_b if sink_b is not None else torch.zeros(N, device=dy.device, dtype=torch.float32)
        if dx is not None and dw is not None and M > 0:
            LIB.call("npfk_linear_bwd_fused", _p(g), N, _p(xv), ldx, _p(wv), ldw, flags, _p(dx), K, _p(dw), K, _p(db),
                     M, N, K, ctx.mode, st)
        elif M > 0:
            if dx is not None:
                LIB.call("npfk_linear_bwd_dx", _p(g), N, _p(wv), ldw, None, 0, _p(dx), K, M, N, K, ctx.mode, st)
            if dw is not None:
                direct = sink_w is not None and (db is None or sink_b is not None)
                if dx is not None and direct and _side["on"] and M <= 65536:
                    # small layer, gradients written straight into the trainer's flat buffer (nothing is handed back to
                    # autograd, whose AccumulateGrad would read it on the main stream): the wgrad runs beside the rest of
                    # the backward pass (see _fork_side)
                    side = _fork_side(g, xv, dw, db)
                    with torch.cuda.stream(side):
                        LIB.call("npfk_linear_bwd_dw", _p(g), N, _p(xv), ldx, _p(dw), K, _p(db), M, N, K, ctx.mode, _st())
                else:
                    LIB.call("npfk_linear_bwd_dw", _p(g), N, _p(xv), ldx, _p(dw), K, _p(db), M, N, K, ctx.mode, st)
        if ctx.has_res and ctx.needs_input_grad[3]:
            if res_rows == M:
                dres = g.reshape(ctx.res_shape)
            else:
                dres = torch.empty(ctx.res_shape, device=dy.device, dtype=torch.float32)
                LIB.call("npfk_rows_fold_sum", _p(g), N, _p(dres), N, res_rows, M // res_rows, N, st)
        if parked is not None and dx is not None:
            dx = dx + parked.view_as(dx)
        dw = None if (sink_w is not None or not need_w or dw is None) else dw.reshape(ctx.w_shape)
        return dx, dw, (None if sink_b is not None else db), dres, None, None


def linear(x, w, b=None, res=None, relu=False, dx_stash=None):
    """act(x @ w.T + b + res); `res` rows are tiled over the rows of x when it has fewer rows.
    ``dx_stash``: a GradStash whose parked tensor (another gradient of x) is added to this layer's dx."""
    return _LinearFn.apply(x, w, b, res, relu, dx_stash)


_CHAIN_DIM = 128      # the fused chain kernel (K1, csrc/tc_mlp_chain.cu) takes layers of 128 -> 128
_CHAIN_MAX = 8


def set_chain_fusion(flag: bool) -> None:
    """Developer switch for A/B timing: False runs every MLP layer as its own GEMM launch (the round-1 path)."""
    _state["chain_fusion"] = bool(flag)


def set_chain_bwd_fusion(flag: bool) -> None:
    """Developer switch for A/B timing: False back-propagates every layer through its own npfk_linear_bwd_fused launch
    instead of pairing 128 -> 128 layers in npfk_linear_bwd_fused2 (K1b)."""
    _state["chain_bwd_fusion"] = bool(flag)


def _ptr_array(ts):
    return (ctypes.c_void_p * len(ts))(*[None if t is None else t.data_ptr() for t in ts])


def _chain_fwd(x, ldx, M, ws, bs, relus, res, res_rows, outs, head=None, nll=None, res_layer=None):
    """one launch of npfk_mlp_chain_fwd over consecutive 128 -> 128 layers; `outs[l]` None = that activation stays on
    chip.  head = (w[2,128], b or None, suff[M,2]); nll = (y, y_rows, T, sum_lp)."""
    L = len(ws)
    relu_mask = sum(1 << l for l, r in enumerate(relus) if r)
    hw, hb, suff = head if head is not None else (None, None, None)
    y, y_rows, T, sum_lp = nll if nll is not None else (None, 0, 0, None)
    LIB.call("npfk_mlp_chain_fwd", _p(x), ldx, M, L, _ptr_array(ws), _ptr_array(bs), relu_mask,
             ((L - 1) if res_layer is None else res_layer) if res is not None else -1, _p(res), _CHAIN_DIM, res_rows, _ptr_array(outs), _p(hw), _p(hb), _p(suff),
             _p(y), y_rows, T, _p(sum_lp), MODE_FAST, _st())


class _MlpChainFn(Function):
    """A whole ReLU MLP (npf/architectures/mlp.py:95-109) as ONE autograd node.

    forward:  h_0 = x ; h_i = relu(h_{i-1} W_i^T + b_i) ; y = act(h_{L-1} W_L^T + b_L + res)
    backward: the ReLU mask of layer i-1 is applied in the dgrad epilogue of layer i, so no separate
    elementwise ReLU-backward passes exist inside the chain.  Two flags extend this across the sum-merge
    (encoders.py:175-183) where the output of one chain (after ReLU) is the input of the next:
      x_is_relu   the input x is itself a ReLU output -> dx is masked with x > 0 in the first layer's dgrad;
      dy_premasked  the consumer of y already masked the incoming gradient with y > 0 (it has x_is_relu set),
                  so the out_relu backward is the identity here.
    In fast mode every run of >= 2 consecutive 128 -> 128 layers is ONE launch of the fused chain kernel
    (npfk_mlp_chain_fwd: activations stay in tensor memory, only what backward needs is written), and a final 128 -> 2
    layer (the decoder head) is evaluated in that kernel's epilogue.  ``nll = (y, T)`` additionally makes the node
    return ``sum_t log N(y; loc, scale)`` per task as a second output (decode tail + sum_log_prob fused, north_star (4)).
    ``res_opts = (res_stash, res_layer)``: the residual is added to the pre-activation of layer ``res_layer`` (None = the last
    one; inside the chain the layer's ReLU follows the sum), and its gradient -- the masked gradient of that pre-activation --
    is either returned or, with a GradStash, handed to the node that collects the other gradients of the same tensor.  In
    backward, pairs of consecutive 128 -> 128 layers go through npfk_linear_bwd_fused2 (K1b) unless the gradient between
    them is the residual's.
    args = (x, res, w_1, b_1, ..., w_L, b_L); b_i may be None.
    """

    @staticmethod
    def forward(ctx, out_relu: bool, x_is_relu: bool, dy_premasked: bool, nll, grad_enabled: bool, res_opts, x, res, *wb):
        require_cuda(x, res, *wb)
        ctx.set_materialize_grads(False)
        n_layers = len(wb) // 2
        # res_opts = (GradStash that takes the residual's gradient or None, index of the layer whose pre-activation gets
        # + res or None = the last one)
        res_stash, res_at = res_opts if res_opts is not None else (None, None)
        if res_at is None:
            res_at = n_layers - 1
        mode = _state["mode"]
        lead = x.shape[:-1]
        M = 1
        for s in lead:
            M *= s
        st = _st()
        h, ldh = _rows_view(x)
        acts = [h]
        wvs, dims = [], []
        for i in range(n_layers):
            w = wb[2 * i]
            if mode == MODE_FAST:
                wv = tf32_weight(w)
                ldw = wv.shape[1]
            else:
                wv, ldw = _weight_view(w)
            wvs.append(wv)
            dims.append([w.shape[0], w.shape[1], 0, ldw, 0])
        training = bool(grad_enabled) and any(ctx.needs_input_grad)
        sq = [d[0] == _CHAIN_DIM and d[1] == _CHAIN_DIM and d[3] == _CHAIN_DIM for d in dims]
        fuse_ok = (mode == MODE_FAST and _state.get("chain_fusion", True) and M >= 128)
        resv, res_rows = None, 0
        if res is not None:
            if not 0 <= res_at < n_layers:
                raise NpfkError(f"mlp: residual layer {res_at} outside the chain of {n_layers} layers")
            resv = _c(res)
            res_rows = resv.numel() // dims[res_at][0]
            if res_rows == 0 or M % max(res_rows, 1) != 0:
                raise NpfkError(f"mlp: residual rows {res_rows} do not tile M={M}")
        head_fused, suff, sum_lp, nll_saved = False, None, None, None
        i = 0
        while i < n_layers:
            j = i
            if fuse_ok and ldh % 4 == 0 and h.data_ptr() % 16 == 0:
                while j < n_layers and sq[j] and j - i < _CHAIN_MAX:
                    j += 1
            # the decoder head: a final 128 -> 2 layer without ReLU or residual rides in the chain's epilogue
            with_head = (j > i and j == n_layers - 1 and dims[j][0] == 2 and dims[j][1] == _CHAIN_DIM
                         and (res is None or res_at != j)
                         and not out_relu and wvs[j].is_contiguous())
            if j - i >= 2 or (with_head and j - i >= 1):
                relus = [True] * (j - i)
                if j == n_layers:               # the run ends the chain: the output activation is the caller's
                    relus[-1] = bool(out_relu)
                r_here = resv if (resv is not None and i <= res_at < j) else None     # the residual lands in this run
                outs = []
                for l in range(i, j):
                    keep = training or l == j - 1
                    if with_head and l == j - 1 and not training:
                        keep = False            # inference: nothing but the head's output leaves the chip
                    outs.append(torch.empty(M, _CHAIN_DIM, device=x.device, dtype=torch.float32) if keep else None)
                head = nl = None
                if with_head:
                    suff = torch.empty(M, 2, device=x.device, dtype=torch.float32)
                    hb = wb[2 * j + 1]
                    head = (wvs[j], _c(hb), suff)
                    if nll is not None:
                        yv, T = nll
                        yv = _c(yv)
                        sum_lp = torch.zeros(M // T, device=x.device, dtype=torch.float32)
                        nl = (yv, yv.numel(), T, sum_lp)
                        nll_saved = (yv, T)
                _chain_fwd(h, ldh, M, wvs[i:j], [_c(wb[2 * l + 1]) for l in range(i, j)], relus, r_here, res_rows, outs,
                           head, nl, res_layer=res_at - i)
                for l in range(i, j):
                    dims[l][2] = ldh if l == i else _CHAIN_DIM
                    dims[l][4] = res_rows if (r_here is not None and l == res_at) else 0
                    if l > i:
                        acts.append(outs[l - i - 1])
                h, ldh = outs[-1], _CHAIN_DIM
                if with_head:
                    acts.append(h)
                    dims[j][2] = _CHAIN_DIM
                    head_fused = True
                    h, ldh = suff, 2
                    i = j + 1
                else:
                    if j < n_layers:
                        acts.append(h)
                    i = j
                continue
            N, Kd = dims[i][0], dims[i][1]
            last = i == n_layers - 1
            bv = _c(wb[2 * i + 1])
            y = torch.empty(M, N, device=x.device, dtype=torch.float32)
            here = resv is not None and i == res_at
            if M > 0:
                LIB.call("npfk_linear_fwd", _p(h), ldh, _p(wvs[i]), dims[i][3], _p(bv), _p(resv if here else None), N,
                         res_rows if here else 0, _p(y), N, M, N, Kd, int(out_relu if last else True), mode, st)
            dims[i][2] = ldh
            dims[i][4] = res_rows if here else 0
            h, ldh = y, N
            if not last:
                acts.append(y)
            i += 1
        ctx.cfg = (out_relu, x_is_relu, dy_premasked, mode, M, n_layers, [tuple(d) for d in dims], res is not None,
                   None if res is None else res.shape, [wb[2 * i + 1] is not None for i in range(n_layers)], x.shape,
                   head_fused, nll_saved, res_at)
        ctx.wb = wb
        ctx.res_stash = res_stash
        if training:
            ctx.save_for_backward(h if (out_relu and not dy_premasked) or head_fused else None, *acts, *wvs,
                                  nll_saved[0] if nll_saved is not None else None)
        out = h.reshape(*lead, dims[-1][0])
        if nll is not None:
            if sum_lp is None:      # the tail did not ride in the chain (fp32 mode, odd shapes): separate kernel
                sum_lp = None
            return out, sum_lp
        return out

    @staticmethod
    def backward(ctx, dy, dsum_lp=None):
        (out_relu, x_is_relu, dy_premasked, mode, M, n_layers, dims, has_res, res_shape, has_b, x_shape, head_fused,
         nll_saved, res_at) = ctx.cfg
        saved = ctx.saved_tensors
        y_out, acts, wvs, y_nll = saved[0], saved[1:1 + n_layers], saved[1 + n_layers:1 + 2 * n_layers], saved[-1]
        st = _st()
        N_out = dims[-1][0]
        dev = acts[0].device
        g = None if dy is None else _c(dy).reshape(M, N_out)
        if g is None and not (head_fused and dsum_lp is not None):
            g = torch.zeros(M, N_out, device=dev, dtype=torch.float32)
        if out_relu and not dy_premasked and M > 0:
            g2 = torch.empty_like(g)
            LIB.call("npfk_relu_bwd", _p(g), N_out, _p(y_out), N_out, _p(g2), N_out, M, N_out, st)
            g = g2
        need_dres = has_res and ctx.needs_input_grad[7]

        def residual_grad(gr):
            """gradient of the residual = the (masked) gradient of the pre-activation it was added to, summed over the rows
            that shared it"""
            res_rows, Nr = dims[res_at][4], dims[res_at][0]
            if res_rows == M:
                return gr.reshape(res_shape)
            out = torch.empty(res_shape, device=gr.device, dtype=torch.float32)
            LIB.call("npfk_rows_fold_sum", _p(gr), Nr, _p(out), Nr, res_rows, M // res_rows, Nr, st)
            return out

        dres = residual_grad(g) if (need_dres and res_at == n_layers - 1) else None
        grads_wb = [None] * (2 * n_layers)

        def need_w_of(i):
            return ctx.needs_input_grad[8 + 2 * i] or (has_b[i] and ctx.needs_input_grad[9 + 2 * i])

        sunk = []       # parameters whose gradients the next launch writes straight into the trainer's flat buffer

        def wgrad_buffers(i):
            """(dw, db) of layer i: the trainer's flat gradient buffer when it takes direct writes, else fresh zeros that are
            handed back to autograd.  The sink callbacks fire in ``launched()``, AFTER the kernel that writes the gradients
            is enqueued: a pair launch (K1b) covers four parameters, and a callback that ran before the launch could make
            the trainer all-reduce a bucket that ends inside the pair before its gradients exist."""
            N, Kd = dims[i][0], dims[i][1]
            sink_w = _grad_sink_buf(ctx.wb[2 * i], (N, Kd)) if ctx.needs_input_grad[8 + 2 * i] else None
            sink_b = _grad_sink_buf(ctx.wb[2 * i + 1]) if has_b[i] and ctx.needs_input_grad[9 + 2 * i] else None
            sunk.extend(prm for prm, buf in ((ctx.wb[2 * i], sink_w), (ctx.wb[2 * i + 1], sink_b)) if buf is not None)
            dw = sink_w if sink_w is not None else torch.zeros(N, Kd, device=dev, dtype=torch.float32)
            db = None
            if has_b[i]:
                db = sink_b if sink_b is not None else torch.zeros(N, device=dev, dtype=torch.float32)
            grads_wb[2 * i] = None if sink_w is not None else dw
            grads_wb[2 * i + 1] = None if sink_b is not None else db
            return dw, db

        def launched():
            cb = _state.get("sink_cb")
            if cb is not None:
                for prm in sunk:
                    cb(prm)
            sunk.clear()

        def square(i):
            return dims[i][0] == _CHAIN_DIM and dims[i][1] == _CHAIN_DIM and dims[i][3] == _CHAIN_DIM

        pair_ok = (mode == MODE_FAST and _state.get("chain_bwd_fusion", True) and M >= 128)
        i = n_layers - 1
        while i >= 0:
            N, Kd, ldh, ldw, _ = dims[i]
            hin = acts[i]
            if need_dres and i == res_at and i < n_layers - 1:
                dres = residual_grad(g)      # a residual inside the chain (decoder: resizer + flat module as one node)
            need_w = need_w_of(i)
            need_dx = i > 0 or ctx.needs_input_grad[6]
            masked = i > 0 or x_is_relu
            if head_fused and i == n_layers - 1 and M > 0:
                # decoder tail: d(log-likelihood)/d(suff), the head's dgrad (+ ReLU mask of the layer below) and its
                # weight / bias gradients in one pass over the last hidden activation
                dw, db = wgrad_buffers(i) if need_w else (None, None)
                dx = torch.empty(M, Kd, device=dev, dtype=torch.float32) if need_dx else None
                yv, T = nll_saved if nll_saved is not None else (None, 0)
                use_lp = dsum_lp is not None and yv is not None
                LIB.call("npfk_nll_head_bwd", _p(hin), ldh, _p(wvs[i]), _p(y_out), _p(y_nll if use_lp else None),
                         y_nll.numel() if use_lp else 0, T if use_lp else 0,
                         _p(_c(dsum_lp).reshape(-1)) if use_lp else None, _p(g), int(masked), _p(dx), _p(dw), _p(db), M, st)
                launched()
                g = dx
                if not need_dx:
                    break
                i -= 1
                continue
            # K1b: two consecutive 128 -> 128 layers in one kernel -- the gradient between them stays on the SM
            if (pair_ok and i >= 1 and square(i) and square(i - 1) and need_w and need_w_of(i - 1)
                    and not (need_dres and res_at == i - 1)      # that gradient must reach HBM: it is the residual's
                    and (i - 1 > 0 or ctx.needs_input_grad[6]) and ldh % 4 == 0 and dims[i - 1][2] % 4 == 0
                    and hin.data_ptr() % 16 == 0 and acts[i - 1].data_ptr() % 16 == 0 and g.data_ptr() % 16 == 0):
                dw1, db1 = wgrad_buffers(i)
                dw0, db0 = wgrad_buffers(i - 1)
                masked0 = (i - 1 > 0) or x_is_relu
                dx = torch.empty(M, _CHAIN_DIM, device=dev, dtype=torch.float32)
                LIB.call("npfk_linear_bwd_fused2", _p(g), N, _p(hin), ldh, _p(wvs[i]), ldw, _p(acts[i - 1]), dims[i - 1][2],
                         _p(wvs[i - 1]), dims[i - 1][3], int(masked0), _p(dx), _CHAIN_DIM, _p(dw1), _CHAIN_DIM, _p(db1),
                         _p(dw0), _CHAIN_DIM, _p(db0), M, st)
                launched()
                g = dx
                i -= 2
                continue
            dw = db = None
            if need_w:
                dw, db = wgrad_buffers(i)
            if need_w and need_dx and M > 0:
                # one pass over g and the saved input: dgrad (+ ReLU mask from hin) and wgrad together
                dx = torch.empty(M, Kd, device=dev, dtype=torch.float32)
                LIB.call("npfk_linear_bwd_fused", _p(g), N, _p(hin), ldh, _p(wvs[i]), ldw, int(masked), _p(dx), Kd,
                         _p(dw), Kd, _p(db), M, N, Kd, mode, st)
                launched()
                g = dx
                i -= 1
                continue
            if need_w and M > 0:
                LIB.call("npfk_linear_bwd_dw", _p(g), N, _p(hin), ldh, _p(dw), Kd, _p(db), M, N, Kd, mode, st)
            launched()
            if not need_dx:
                g = None
                break
            dx = torch.empty(M, Kd, device=dev, dtype=torch.float32)
            if M > 0:
                LIB.call("npfk_linear_bwd_dx", _p(g), N, _p(wvs[i]), ldw, _p(hin) if masked else None, ldh if masked else 0,
                         _p(dx), Kd, M, N, Kd, mode, st)
            g = dx
            i -= 1
        if dres is not None and ctx.res_stash is not None:
            # the residual is another use of a tensor whose producer collects its gradients through a GradStash (the
            # transformer block: c feeds the MLP and, as the residual of its last layer, LN2): hand it over there instead
            # of returning it, so autograd does not sum two [M, 128] tensors in a separate pass
            ctx.res_stash.t, dres = dres, None
        dx = g.reshape(x_shape) if (g is not None and ctx.needs_input_grad[6]) else None
        return (None, None, None, None, None, None, dx, dres, *grads_wb)


def mlp_chain(x, layers, res=None, out_relu=False, x_is_relu=False, dy_premasked=False, nll=None, res_stash=None,
              res_layer=None):
    """ReLU MLP over `layers` = [(weight, bias), ...] as a single autograd node (see _MlpChainFn).  ``nll=(y, T)``:
    returns ``(out, sum_lp)`` where sum_lp[m // T] = sum of log N(y; loc, scale) over the T rows of a task, or
    ``(out, None)`` when the tail could not ride in the fused chain (the caller then uses gauss_sum_log_prob).
    ``res`` is added to the pre-activation of layer ``res_layer`` (default: the last one; a hidden layer applies its ReLU
    after the sum -- the decoder's ``relu(x1 + resizer(x2))`` followed by its flat module is one such chain)."""
    flat = []
    for w, b in layers:
        flat += [w, b]
    # Function.forward cannot see the caller's grad mode (needs_input_grad ignores no_grad): pass it, so that inference
    # keeps nothing for a backward pass and writes only the chain's result
    return _MlpChainFn.apply(out_relu, x_is_relu, dy_premasked, nll, torch.is_grad_enabled(), (res_stash, res_layer), x,
                             res, *flat)


# ------------------------------------------------------------------------------------------ broadcast merge
class _BcastAddReluFn(Function):
    @staticmethod
    def forward(ctx, a, r):
        # a:[B,T,C]  r:[Z,B,C] -> [Z,B,T,C]
        require_cuda(a, r)
        a, r = _c(a), _c(r)
        B, T, C = a.shape
        Z = r.shape[0]
        out = torch.empty(Z, B, T, C, device=a.device, dtype=torch.float32)
        LIB.call("npfk_bcast_add_relu_fwd", _p(a), _p(r), _p(out), Z, B, T, C, _st())
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, dout):
        (out,) = ctx.saved_tensors
        Z, B, T, C = out.shape
        dout = _c(dout)
        da = torch.empty(B, T, C, device=out.device, dtype=torch.float32) if ctx.needs_input_grad[0] else None
        dr = torch.zeros(Z, B, C, device=out.device, dtype=torch.float32) if ctx.needs_input_grad[1] else None
        LIB.call("npfk_bcast_add_relu_bwd", _p(dout), _p(out), _p(da), _p(dr), Z, B, T, C, _st())
        return da, dr


def bcast_add_relu(a, r):
    return _BcastAddReluFn.apply(a, r)


class _SetMeanFn(Function):
    @staticmethod
    def forward(ctx, x):
        require_cuda(x)
        x = _c(x)
        B, N, C = x.shape
        out = torch.empty(B, 1, C, device=x.device, dtype=torch.float32)
        LIB.call("npfk_set_mean_fwd", _p(x), _p(out), B, N, C, _st())
        ctx.shape = (B, N, C)
        return out

    @staticmethod
    def backward(ctx, dout):
        B, N, C = ctx.shape
        dout = _c(dout)
        dx = torch.empty(B, N, C, device=dout.device, dtype=torch.float32)
        LIB.call("npfk_set_mean_bwd", _p(dout), _p(dx), B, N, C, _st())
        return dx


def set_mean(x):
    """mean over dim 1 with keepdim: [B,N,C] -> [B,1,C]"""
    return _SetMeanFn.apply(x)


# ------------------------------------------------------------------------------------------ attention
class _MhaFn(Function):
    @staticmethod
    def forward(ctx, q, k, v, n_heads: int, scale: float, delta_stash=None):
        require_cuda(q, k, v)
        ctx.delta_stash = delta_stash
        q, k, v = _c(q), _c(k), _c(v)
        B, Nq, E = q.shape
        Nk = k.shape[1]
        D = E // n_heads
        out = torch.empty_like(q)
        lse = torch.empty(B, n_heads, Nq, device=q.device, dtype=torch.float32)
        mode = _state["mode"]
        LIB.call("npfk_mha_fwd", _p(q), _p(k), _p(v), _p(out), _p(lse), B, Nq, Nk, n_heads, D, float(scale), mode, _st())
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.cfg = (B, Nq, Nk, n_heads, D, float(scale), mode)
        if delta_stash is not None:        # what the consumer of `out` needs to write delta = sum_d dO*O for us
            delta_stash.spec = (B, Nq, n_heads, D, mode, out.data_ptr())
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        B, Nq, Nk, H, D, scale, mode = ctx.cfg
        dout = _c(dout)
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        ws, ready = None, 0
        st_ = ctx.delta_stash
        if st_ is not None and st_.t is not None:
            ws_, ptr = st_.t
            st_.t = None
            if ptr == dout.data_ptr():     # delta was computed from exactly this dO
                ws, ready = ws_, 1
        if ws is None:
            ws = torch.empty(int(LIB.load().npfk_mha_bwd_ws_floats(B, Nq, H)), device=q.device, dtype=torch.float32)
        LIB.call("npfk_mha_bwd", _p(q), _p(k), _p(v), _p(out), _p(lse), _p(dout), _p(dq), _p(dk), _p(dv), _p(ws), B, Nq,
                 Nk, H, D, scale, ready, mode, _st())
        return dq, dk, dv, None, None, None


def mha(q, k, v, n_heads: int, scale: float, delta_stash=None):
    """softmax(q k^T * scale) v per head; heads are slices of the feature dim (no permutes).
    ``delta_stash``: GradStash shared with the ``add_layernorm`` that consumes the output as its ``a`` (attn_of=...)."""
    return _MhaFn.apply(q, k, v, n_heads, scale, delta_stash)


class GradStash:
    """Side channel between two ``add_layernorm`` nodes of one transformer block, c = LN1(..) ; out = LN2(c + f(c)):
    LN2's backward parks its gradient for the direct ``c`` input here instead of returning it, and LN1's backward adds
    it to its incoming gradient inside the kernel (``dout2``).  autograd then never sums the two 537 MB gradients of
    ``c`` with a separate kernel.  Valid because ``f(c)`` always back-propagates into ``c`` after LN2's backward ran."""

    __slots__ = ("t", "spec")

    def __init__(self):
        self.t = None
        self.spec = None


class _AddLayerNormFn(Function):
    @staticmethod
    def forward(ctx, a, b, gamma, beta, eps: float, park_a: Optional[GradStash], extra: Optional[GradStash],
                park_b: Optional[GradStash], attn_of: Optional[GradStash]):
        require_cuda(a, b, gamma, beta)
        ctx.params = (gamma, beta)
        a, b, gamma, beta = _c(a), _c(b), _c(gamma), _c(beta)
        C = a.shape[-1]
        rows = a.numel() // C
        out = torch.empty_like(a)
        mean = torch.empty(rows, device=a.device, dtype=torch.float32)
        rstd = torch.empty(rows, device=a.device, dtype=torch.float32)
        LIB.call("npfk_add_layernorm_fwd", _p(a), _p(b), _p(gamma), _p(beta), _p(out), _p(mean), _p(rstd), rows, C,
                 float(eps), _st())
        ctx.save_for_backward(a, b, gamma, mean, rstd)
        ctx.has_b = b is not None
        ctx.park_a, ctx.extra, ctx.park_b = park_a, extra, park_b
        # `a` is the untouched output of mha (same storage): the backward can write the attention backward's delta
        ctx.attn_of = attn_of if (attn_of is not None and attn_of.spec is not None
                                  and attn_of.spec[5] == a.data_ptr()) else None
        return out

    @staticmethod
    def backward(ctx, dout):
        a, b, gamma, mean, rstd = ctx.saved_tensors
        C = a.shape[-1]
        rows = a.numel() // C
        dout = _c(dout)
        dout2 = None
        if ctx.extra is not None and ctx.extra.t is not None:
            dout2, ctx.extra.t = ctx.extra.t, None
            if dout2.shape != dout.shape:
                raise NpfkError("add_layernorm: parked gradient does not match the incoming one")
        ds = torch.empty_like(a)
        pg, pb = ctx.params
        sink_g = _grad_sink(pg) if ctx.needs_input_grad[2] else None
        sink_b = _grad_sink(pb) if ctx.needs_input_grad[3] else None
        dg = sink_g if sink_g is not None else torch.zeros_like(gamma)
        db = sink_b if sink_b is not None else torch.zeros_like(gamma)
        delta, H, Nq, Nqp = None, 0, 0, 0
        if ctx.attn_of is not None:
            aB, aNq, aH, aD, amode, _ = ctx.attn_of.spec
            if (amode == MODE_FAST and aD == 16 and aH % 2 == 0 and aH * 16 == C and C <= 128 and rows == aB * aNq
                    and aNq % 8 == 0):
                n_ws = int(LIB.load().npfk_mha_bwd_ws_floats(aB, aNq, aH))
                ws = torch.empty(n_ws, device=a.device, dtype=torch.float32)
                H, Nq, Nqp = aH, aNq, n_ws // (2 * aB * aH)
                delta = ws[n_ws // 2:]
                ctx.attn_of.t = (ws, ds.data_ptr())
        LIB.call("npfk_add_layernorm_bwd", _p(dout), _p(dout2), _p(a), _p(b), _p(gamma), _p(mean), _p(rstd), _p(ds),
                 _p(dg), _p(db), rows, C, _p(delta), H, Nq, Nqp, _st())
        da = ds
        if ctx.park_a is not None and ctx.has_b and ctx.needs_input_grad[0] and ctx.needs_input_grad[1]:
            ctx.park_a.t, da = ds, None            # the producer of `a` picks it up as its dout2
        dbr = ds if ctx.has_b else None
        if ctx.park_b is not None and ctx.has_b and ctx.needs_input_grad[1]:
            # b's other consumer (a Linear with dx_stash=park_b, back-propagated AFTER everything that reads `da`)
            # adds its dx onto this buffer and returns the sum as b's only gradient
            ctx.park_b.t, dbr = ds, None
        return (da, dbr, (None if sink_g is not None else dg), (None if sink_b is not None else db), None, None, None,
                None, None)


def add_layernorm(a, b, gamma, beta, eps=1e-5, park_a=None, extra=None, park_b=None, attn_of=None):
    """LayerNorm(a + b) over the last dim.  ``park_a`` / ``extra`` / ``park_b``: see GradStash (the producing and the
    consuming node must share one stash).  ``attn_of``: the stash given to the ``mha`` call that produced ``a``; the
    backward then also writes the attention backward's delta = sum_d dO*O while ds (= dO) and a (= O) are in registers."""
    return _AddLayerNormFn.apply(a, b, gamma, beta, eps, park_a, extra, park_b, attn_of)


# ------------------------------------------------------------------------------------------ SetConv
class _SetConvRbfFn(Function):
    @staticmethod
    def forward(ctx, xk, xq, v, theta, p: float, unshared: bool):
        # xk:[B,Nk,1] or [Nk] (shared grid)  xq:[B,Nq,1] or [Nq]  v:[B,Nk,C]  -> [B,Nq,C+1]
        require_cuda(xk, xq, v, theta)
        v, theta = _c(v), _c(theta)
        B, Nk, C = v.shape
        xk, sxk = _positions(xk, B, Nk)
        xq, sxq = _positions(xq, B, None)
        Nq = xq.numel() if sxq == 0 else xq.numel() // B
        out = torch.empty(B, Nq, C + 1, device=v.device, dtype=torch.float32)
        LIB.call("npfk_setconv_fwd", _p(xk), sxk, _p(xq), sxq, _p(v), _p(theta), _p(out), C + 1, B, Nk, Nq, C, float(p),
                 int(unshared), _st())
        ctx.save_for_backward(xk, xq, v, theta, out)
        ctx.cfg = (sxk, sxq, B, Nk, Nq, C, float(p), int(unshared))
        return out

    @staticmethod
    def backward(ctx, dout):
        xk, xq, v, theta, out = ctx.saved_tensors
        sxk, sxq, B, Nk, Nq, C, p, unshared = ctx.cfg
        dout = _c(dout)
        # values that are data (context -> grid) need no dv: few-channel inputs then take the length-scale-only kernel
        dv = zeros(*v.shape, device=v.device) if (ctx.needs_input_grad[2] or C > 4) else None
        dtheta = zeros(*theta.shape, device=v.device)
        LIB.call("npfk_setconv_bwd", _p(xk), sxk, _p(xq), sxq, _p(v), _p(theta), _p(out), _p(dout), C + 1, _p(dv),
                 _p(dtheta), B, Nk, Nq, C, p, unshared, _st())
        return None, None, dv, dtheta, None, None


def _positions(x, B, n):
    """positions as a flat fp32 tensor + batch stride (0 when one grid is shared by the whole batch)"""
    x = x.detach()
    if x.dim() == 1:
        return _c(x), 0
    if x.dim() == 3 and x.shape[-1] == 1:
        if x.stride(0) == 0:      # expanded grid
            return _c(x[0, :, 0]), 0
        x = _c(x)
        return x, x.shape[1]
    raise NpfkError(f"SetConv positions must be [N] or [B,N,1] (x_dim == 1); got {tuple(x.shape)}")


def setconv_rbf(xk, xq, v, theta, p=2.0, unshared=False):
    """cat(sum_k softmax-weights * v_k, density) with the exponential RBF; positions get no gradient."""
    return _SetConvRbfFn.apply(xk, xq, v, theta, p, unshared)


class _SetConvInterpFn(Function):
    @staticmethod
    def forward(ctx, xk, xq, v, theta, wd, bias, p: float, unshared: bool):
        # v:[B,Nk,C] (resizer already applied to the values)  wd:[C]  bias:[C]|None  -> [B,Nq,C]
        require_cuda(xk, xq, v, theta, wd, bias)
        v, theta, wd, bias = _c(v), _c(theta), _c(wd), _c(bias)
        B, Nk, C = v.shape
        xk, sxk = _positions(xk, B, Nk)
        xq, sxq = _positions(xq, B, None)
        Nq = xq.numel() if sxq == 0 else xq.numel() // B
        out = torch.empty(B, Nq, C, device=v.device, dtype=torch.float32)
        stats = torch.empty(B, Nq, 2, device=v.device, dtype=torch.float32)
        LIB.call("npfk_setconv_interp_fwd", _p(xk), sxk, _p(xq), sxq, _p(v), _p(theta), _p(wd), _p(bias), _p(out),
                 _p(stats), B, Nk, Nq, C, float(p), int(unshared), _st())
        ctx.save_for_backward(xk, xq, v, theta, wd, stats)
        ctx.cfg = (sxk, sxq, B, Nk, Nq, C, float(p), int(unshared), bias is not None)
        return out

    @staticmethod
    def backward(ctx, dout):
        xk, xq, v, theta, wd, stats = ctx.saved_tensors
        sxk, sxq, B, Nk, Nq, C, p, unshared, has_bias = ctx.cfg
        dout = _c(dout)
        dv = zeros(*v.shape, device=v.device)
        dtheta = zeros(*theta.shape, device=v.device)
        dwd = zeros(*wd.shape, device=v.device)
        dbias = zeros(*wd.shape, device=v.device) if has_bias else None
        LIB.call("npfk_setconv_interp_bwd", _p(xk), sxk, _p(xq), sxq, _p(v), _p(theta), _p(wd), _p(stats), _p(dout),
                 _p(dv), _p(dwd), _p(dbias), _p(dtheta), B, Nk, Nq, C, p, unshared, _st())
        return None, None, dv, dtheta, dwd, dbias, None, None


def setconv_interp(xk, xq, v, theta, wd, bias=None, p=2.0, unshared=False):
    """sum_k w_qk v_k + density_q * wd + bias: SetConv with its resizer already applied to the values `v`
    (grid -> targets direction, many more queries than keys); positions get no gradient."""
    return _SetConvInterpFn.apply(xk, xq, v, theta, wd, bias, p, unshared)


# ------------------------------------------------------------------------------------------ conv blocks
def set_sync_bn(flag: bool, group=None) -> None:
    """``sync_bn`` option of SURVEY.md section 8e: in training mode the BatchNorm layers of the conv blocks all-reduce
    their per-channel (sum x, sum x^2, count) over ``group`` (default: the world) in forward and (sum dh, sum dh*xhat) in
    backward, so that data-parallel training normalises with the statistics of the GLOBAL batch, as the reference does
    when it sees that batch on one device (npf/architectures/cnn.py:189,193).  Off by default: per-replica statistics."""
    _state["sync_bn"] = (bool(flag), group)


def _sync_bn_group():
    flag, group = _state.get("sync_bn", (False, None))
    if not flag:
        return None
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) < 2:
        return None
    return (group,)          # a 1-tuple: `group` itself may be None (= the default group)


class _NormActDwConvFn(Function):
    """y = dwconv(act(norm(x))) + bias + res on channel-last [B,L,C]; BatchNorm statistics handled inside."""

    @staticmethod
    def forward(ctx, x, gamma, beta, w, bias, res, running_mean, running_var, training: bool, eps: float,
                momentum: float, relu: bool, num_batches_tracked=None):
        require_cuda(x, w)
        x, w = _c(x), _c(w)
        B, L, C = x.shape
        ks = w.shape[-1]
        w2 = w.reshape(C, ks)
        has_norm = gamma is not None
        st = _st()
        stats = None
        mean = var = None
        if has_norm:
            gamma, beta = _c(gamma), _c(beta)
            sync = _sync_bn_group() if training else None
            ctx.sync = None
            if training and sync is not None:
                import torch.distributed as dist

                # local statistics (no running-stat update), then one all-reduce of [sum x | sum x^2 | count]
                stats = torch.empty(2 * C, device=x.device, dtype=torch.float32)
                scratch = torch.empty(2 * C + 1, device=x.device, dtype=torch.float64)
                LIB.call("npfk_bn_stats", _p(x), B * L, C, _p(scratch), _p(stats), None, None, 0.0, None, st)
                n_loc = float(B * L)
                m_l, v_l = stats[:C].double(), stats[C:].double()
                pack = torch.cat([m_l * n_loc, (v_l + m_l * m_l) * n_loc, torch.full((1,), n_loc, device=x.device,
                                                                                    dtype=torch.float64)])
                dist.all_reduce(pack, group=sync[0])
                n_tot = pack[-1]
                m_g = pack[:C] / n_tot
                v_g = (pack[C:2 * C] / n_tot - m_g * m_g).clamp_min(0.0)
                mean, var = m_g.float(), v_g.float()
                with torch.no_grad():
                    if running_mean is not None:
                        running_mean.mul_(1.0 - momentum).add_(mean, alpha=momentum)
                        running_var.mul_(1.0 - momentum).add_((v_g * (n_tot / (n_tot - 1.0).clamp_min(1.0))).float(),
                                                               alpha=momentum)
                    if num_batches_tracked is not None:
                        num_batches_tracked.add_(1)
                ctx.sync = (sync[0], n_loc, n_tot.float())
            elif training:
                stats = torch.empty(2 * C, device=x.device, dtype=torch.float32)
                scratch = torch.empty(2 * C + 1, device=x.device, dtype=torch.float64)
                LIB.call("npfk_bn_stats", _p(x), B * L, C, _p(scratch), _p(stats), _p(running_mean), _p(running_var),
                         float(momentum), _p(num_batches_tracked), st)
                mean, var = stats[:C], stats[C:]
            else:
                mean, var = _c(running_mean), _c(running_var)
        bias_c, res_c = _c(bias), _c(res)
        y = torch.empty_like(x)
        LIB.call("npfk_dwconv_fwd", _p(x), _p(mean), _p(var), _p(gamma), _p(beta), float(eps), _p(w2), _p(bias_c),
                 _p(res_c), _p(y), B, L, C, ks, int(relu), st)
        ctx.params = (gamma, beta, w, bias)
        ctx.save_for_backward(x, gamma, beta, w2, mean, var)
        ctx.cfg = (B, L, C, ks, has_norm, bool(training), float(eps), bool(relu), bias is not None, res is not None,
                   w.shape)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, beta, w2, mean, var = ctx.saved_tensors
        B, L, C, ks, has_norm, training, eps, relu, has_bias, has_res, wshape = ctx.cfg
        dy = _c(dy)
        st = _st()
        dev = x.device
        dh = torch.empty_like(x)
        # every "+=" output of the two kernels below comes out of one zero-filled buffer (one fill instead of five)
        z = zeros(C * (ks + 5), device=dev)
        pg, pb, pw, pbias = ctx.params
        ng = ctx.needs_input_grad
        sinks = (_grad_sink(pg) if ng[1] else None, _grad_sink(pb) if ng[2] else None,
                 _grad_sink(pw, (C, ks)) if ng[3] else None, _grad_sink(pbias) if ng[4] else None)
        dw = sinks[2] if sinks[2] is not None else z[:C * ks].view(C, ks)
        dbias = (sinks[3] if sinks[3] is not None else z[C * ks:C * (ks + 1)]) if has_bias else None
        sums = z[C * (ks + 1):C * (ks + 3)] if has_norm else None
        LIB.call("npfk_dwconv_bwd", _p(dy), _p(x), _p(mean), _p(var), _p(gamma), _p(beta), eps, _p(w2), _p(dh), _p(dw),
                 _p(dbias), _p(sums), B, L, C, ks, int(relu), st)
        dgamma = dbeta = None
        dx = dh
        if has_norm:
            dgamma = sinks[0] if sinks[0] is not None else z[C * (ks + 3):C * (ks + 4)]
            dbeta = sinks[1] if sinks[1] is not None else z[C * (ks + 4):]
            if getattr(ctx, "sync", None) is not None:
                import torch.distributed as dist

                # dx needs the GLOBAL sums (sum dh, sum dh*xhat) / N_total; gamma / beta take this rank's sums (the
                # gradient all-reduce adds the ranks up).  npfk_bn_bwd divides by the local row count, so the global
                # sums are pre-scaled by n_local / N_total; its "+=" of the sums goes to a scratch pair
                group, n_loc, n_tot = ctx.sync
                local = sums.clone()
                glob = sums.clone()
                dist.all_reduce(glob, group=group)
                glob.mul_(n_loc / n_tot)
                junk = torch.empty(2 * C, device=dev, dtype=torch.float32)
                LIB.call("npfk_bn_bwd", _p(dh), _p(x), _p(mean), _p(var), _p(gamma), eps, _p(glob), _p(dx), _p(junk[:C]),
                         _p(junk[C:]), B * L, C, int(training), st)
                dbeta.add_(local[:C])
                dgamma.add_(local[C:])
            else:
                LIB.call("npfk_bn_bwd", _p(dh), _p(x), _p(mean), _p(var), _p(gamma), eps, _p(sums), _p(dx), _p(dgamma),
                         _p(dbeta), B * L, C, int(training), st)
        if has_norm:
            dgamma, dbeta = (None if sinks[0] is not None else dgamma), (None if sinks[1] is not None else dbeta)
        dw = None if sinks[2] is not None else dw.reshape(wshape)
        dbias = None if sinks[3] is not None else dbias
        return (dx, dgamma, dbeta, dw, dbias, (dy if has_res else None), None, None, None, None, None,
                None, None)


def norm_act_dwconv(x, gamma, beta, w, bias, res, running_mean, running_var, training, eps, momentum, relu=True,
                    num_batches_tracked=None):
    """``num_batches_tracked`` (BatchNorm1d's int64 buffer) is incremented on the device in training mode"""
    return _NormActDwConvFn.apply(x, gamma, beta, w, bias, res, running_mean, running_var, training, eps, momentum,
                                  relu, num_batches_tracked)


class _AvgPool2Fn(Function):
    @staticmethod
    def forward(ctx, x):
        require_cuda(x)
        x = _c(x)
        B, L, C = x.shape
        y = torch.empty(B, L // 2, C, device=x.device, dtype=torch.float32)
        LIB.call("npfk_avgpool2_fwd", _p(x), _p(y), B, L, C, _st())
        ctx.shape = (B, L, C)
        return y

    @staticmethod
    def backward(ctx, dy):
        B, L, C = ctx.shape
        dy = _c(dy)
        dx = torch.empty(B, L, C, device=dy.device, dtype=torch.float32)
        LIB.call("npfk_avgpool2_bwd", _p(dy), _p(dx), B, L, C, _st())
        return dx


def avgpool2(x):
    return _AvgPool2Fn.apply(x)


class _Upsample2CatFn(Function):
    """cat(upsample_linear_x2(x), skip) along channels, written straight into the concat buffer."""

    @staticmethod
    def forward(ctx, x, skip):
        require_cuda(x, skip)
        x = _c(x)
        B, L, C1 = x.shape
        st = _st()
        if skip is None:
            y = torch.empty(B, 2 * L, C1, device=x.device, dtype=torch.float32)
            LIB.call("npfk_upsample2_fwd", _p(x), _p(y), C1, B, L, C1, st)
            ctx.cfg = (B, L, C1, 0)
            return y
        skip = _c(skip)
        C2 = skip.shape[-1]
        if skip.shape[1] != 2 * L:
            raise NpfkError(f"upsample/concat length mismatch: {2 * L} vs {skip.shape[1]}")
        y = torch.empty(B, 2 * L, C1 + C2, device=x.device, dtype=torch.float32)
        LIB.call("npfk_upsample2_fwd", _p(x), _p(y), C1 + C2, B, L, C1, st)
        LIB.call("npfk_copy2d", _p(skip), C2, ctypes.c_void_p(y.data_ptr() + 4 * C1), C1 + C2, B * 2 * L, C2, st)
        ctx.cfg = (B, L, C1, C2)
        return y

    @staticmethod
    def backward(ctx, dy):
        B, L, C1, C2 = ctx.cfg
        dy = _c(dy)
        st = _st()
        dx = torch.zeros(B, L, C1, device=dy.device, dtype=torch.float32)
        LIB.call("npfk_upsample2_bwd", _p(dy), C1 + C2, _p(dx), B, L, C1, st)
        dskip = None
        if C2:
            dskip = torch.empty(B, 2 * L, C2, device=dy.device, dtype=torch.float32)
            LIB.call("npfk_copy2d", ctypes.c_void_p(dy.data_ptr() + 4 * C1), C1 + C2, _p(dskip), C2, B * 2 * L, C2, st)
        return dx, dskip


def upsample2_cat(x, skip=None):
    return _Upsample2CatFn.apply(x, skip)


# ------------------------------------------------------------------------------------------ distributions / losses
class _LocScaleFn(Function):
    @staticmethod
    def forward(ctx, suff):
        require_cuda(suff)
        suff = _c(suff)
        ydim = suff.shape[-1] // 2
        n = suff.numel() // (2 * ydim)
        loc = torch.empty(*suff.shape[:-1], ydim, device=suff.device, dtype=torch.float32)
        scale = torch.empty_like(loc)
        LIB.call("npfk_loc_scale_fwd", _p(suff), _p(loc), _p(scale), n, ydim, _st())
        ctx.save_for_backward(suff)
        return loc, scale

    @staticmethod
    def backward(ctx, dloc, dscale):
        (suff,) = ctx.saved_tensors
        ydim = suff.shape[-1] // 2
        n = suff.numel() // (2 * ydim)
        dsuff = torch.empty_like(suff)
        LIB.call("npfk_loc_scale_bwd", _p(suff), _p(_c(dloc)), _p(_c(dscale)), _p(dsuff), n, ydim, _st())
        return dsuff


def loc_scale(suff):
    """split [..., 2*ydim] into loc and scale = 0.01 + 0.99*softplus(.)"""
    return _LocScaleFn.apply(suff)


class _GaussNllFn(Function):
    @staticmethod
    def forward(ctx, suff, y):
        # suff:[Z,B,T,2*ydim]  y:[B,T,ydim] -> sum_t log p : [Z,B]
        require_cuda(suff, y)
        suff, y = _c(suff), _c(y)
        Z, B = suff.shape[0], suff.shape[1]
        ydim = suff.shape[-1] // 2
        T = suff.numel() // (Z * B * 2 * ydim)
        out = torch.zeros(Z, B, device=suff.device, dtype=torch.float32)
        LIB.call("npfk_gauss_nll_fwd", _p(suff), _p(y), _p(out), Z, B, T, ydim, _st())
        ctx.save_for_backward(suff, y)
        ctx.cfg = (Z, B, T, ydim)
        return out

    @staticmethod
    def backward(ctx, dout):
        suff, y = ctx.saved_tensors
        Z, B, T, ydim = ctx.cfg
        dsuff = torch.empty_like(suff)
        LIB.call("npfk_gauss_nll_bwd", _p(suff), _p(y), _p(_c(dout)), _p(dsuff), Z, B, T, ydim, _st())
        return dsuff, None


def gauss_sum_log_prob(suff, y):
    return _GaussNllFn.apply(suff, y)


class _LatentStatsFn(Function):
    @staticmethod
    def forward(ctx, suff):
        require_cuda(suff)
        suff = _c(suff)
        zd = suff.shape[-1] // 2
        n = suff.numel() // (2 * zd)
        loc = torch.empty(*suff.shape[:-1], zd, device=suff.device, dtype=torch.float32)
        scale = torch.empty_like(loc)
        LIB.call("npfk_latent_stats_fwd", _p(suff), _p(loc), _p(scale), n, zd, _st())
        ctx.save_for_backward(suff)
        return loc, scale

    @staticmethod
    def backward(ctx, dloc, dscale):
        (suff,) = ctx.saved_tensors
        zd = suff.shape[-1] // 2
        n = suff.numel() // (2 * zd)
        dsuff = torch.empty_like(suff)
        LIB.call("npfk_latent_stats_bwd", _p(suff), _p(_c(dloc)), _p(_c(dscale)), _p(dsuff), n, zd, _st())
        return dsuff


def latent_stats(suff):
    """split [..., 2*z] into loc and scale = 0.1 + 0.9*sigmoid(.)"""
    return _LatentStatsFn.apply(suff)


class _RSampleFn(Function):
    @staticmethod
    def forward(ctx, loc, scale, eps):
        require_cuda(loc, scale, eps)
        loc, scale, eps = _c(loc), _c(scale), _c(eps)
        S = eps.shape[0]
        n = loc.numel()
        z = torch.empty_like(eps)
        LIB.call("npfk_rsample_fwd", _p(loc), _p(scale), _p(eps), _p(z), S, n, _st())
        ctx.save_for_backward(eps)
        ctx.shape = loc.shape
        return z

    @staticmethod
    def backward(ctx, dz):
        (eps,) = ctx.saved_tensors
        S = eps.shape[0]
        dz = _c(dz)
        dloc = torch.empty(ctx.shape, device=dz.device, dtype=torch.float32)
        dscale = torch.empty_like(dloc)
        LIB.call("npfk_rsample_bwd", _p(dz), _p(eps), _p(dloc), _p(dscale), S, dloc.numel(), _st())
        return dloc, dscale, None


def rsample(loc, scale, eps):
    """z[s] = loc + scale * eps[s]"""
    return _RSampleFn.apply(loc, scale, eps)


class _KlDiagFn(Function):
    @staticmethod
    def forward(ctx, lq, sq, lp, sp):
        require_cuda(lq, sq, lp, sp)
        lq, sq, lp, sp = _c(lq), _c(sq), _c(lp), _c(sp)
        zd = lq.shape[-1]
        n = lq.numel() // zd
        kl = torch.empty(lq.shape[:-1], device=lq.device, dtype=torch.float32)
        LIB.call("npfk_kl_diag_fwd", _p(lq), _p(sq), _p(lp), _p(sp), _p(kl), n, zd, _st())
        ctx.save_for_backward(lq, sq, lp, sp)
        return kl

    @staticmethod
    def backward(ctx, dkl):
        lq, sq, lp, sp = ctx.saved_tensors
        zd = lq.shape[-1]
        n = lq.numel() // zd
        outs = [torch.empty_like(lq) for _ in range(4)]
        LIB.call("npfk_kl_diag_bwd", _p(lq), _p(sq), _p(lp), _p(sp), _p(_c(dkl)), *[_p(o) for o in outs], n, zd, _st())
        return tuple(outs)


def kl_diag_normal(lq, sq, lp, sp):
    """KL(N(lq,sq) || N(lp,sp)) summed over the last dim"""
    return _KlDiagFn.apply(lq, sq, lp, sp)


def range_flag(x, flag, lo=-1.0, hi=1.0) -> None:
    """flag |= any(x outside [lo,hi]); flag is a device int32 scalar tensor"""
    require_cuda(x, flag)
    x = _c(x.detach())
    LIB.call("npfk_range_flag", _p(x), x.numel(), float(lo), float(hi), _p(flag), _st())


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, step, grad_scale=1.0) -> None:
    """torch.optim.Adam update on flat fp32 buffers, in place"""
    require_cuda(param, grad, exp_avg, exp_avg_sq)
    LIB.call("npfk_adam_step", _p(param), _p(grad), _p(exp_avg), _p(exp_avg_sq), param.numel(), float(lr), float(beta1),
             float(beta2), float(eps), int(step), float(grad_scale), _st())
    tf32_cache_clear()          # the parameters living in `param` changed without a version bump


def adam_step_dev(param, grad, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, step_dev, grad_scale=1.0) -> None:
    """``adam_step`` with the step index in device memory (int32 tensor, incremented by the call): capturable"""
    require_cuda(param, grad, exp_avg, exp_avg_sq, step_dev)
    assert step_dev.dtype == torch.int32 and step_dev.numel() == 1
    LIB.call("npfk_adam_step_dev", _p(param), _p(grad), _p(exp_avg), _p(exp_avg_sq), param.numel(), float(lr),
             float(beta1), float(beta2), float(eps), _p(step_dev), float(grad_scale), _st())
    if not torch.cuda.is_current_stream_capturing():
        tf32_cache_clear()


def launch_count() -> int:
    return LIB.launch_count()


# ------------------------------------------------------------------------------------------ per-kernel timing (bench.py)
class KernelTimer:
    """Brackets selected C-ABI calls with CUDA events on the launching stream (torch's current stream) and
    accumulates algorithmic work per call.  Used by bench.py for the roofline of the dominant kernel."""

    def __init__(self, names):
        self.names = set(names)
        self.records = []          # (name, work, ev0, ev1)
        self._orig = None

    def _work(self, name, a):
        """(executed flops, algorithmic HBM bytes) of one call; argument positions follow include/npfk.h"""
        if name == "npfk_linear_fwd":
            M, N, Kd = a[10], a[11], a[12]
            res_rows = a[7] if a[5] else 0
            return 2.0 * M * N * Kd, 4.0 * (M * Kd + M * N + (M * N if res_rows == M else 0))
        if name == "npfk_linear_bwd_dx":
            M, N, Kd = a[8], a[9], a[10]
            return 2.0 * M * N * Kd, 4.0 * (M * N + M * Kd * (2 if a[4] else 1))
        if name == "npfk_linear_bwd_dw":
            M, N, Kd = a[7], a[8], a[9]
            return 2.0 * M * N * Kd, 4.0 * (M * N + M * Kd)
        if name == "npfk_linear_bwd_fused":
            M, N, Kd = a[12], a[13], a[14]
            return 4.0 * M * N * Kd, 4.0 * (M * N + 2 * M * Kd)
        if name == "npfk_linear_bwd_fused2":
            M = a[19]       # g, x1, x0 read, dx written: four [M,128] tensors; two dgrads + two wgrads
            return 8.0 * M * 128 * 128, 4.0 * 4 * M * 128
        if name == "npfk_mlp_chain_fwd":
            # algorithmic HBM bytes: the input tile, the residual, every activation the caller asked for, the head's [M,2]
            M, L = a[2], a[3]
            n_out = sum(1 for v in a[11] if v)
            head = 1 if a[12] else 0
            return (2.0 * M * 128 * 128 * L + 4.0 * M * 128 * head,
                    4.0 * M * (128 * (1 + (1 if a[8] else 0) + n_out) + 3 * head))
        if name == "npfk_nll_head_bwd":
            M = a[13]
            return 8.0 * M * 128, 4.0 * M * (128 * (2 if a[10] else 1) + 4)
        if name == "npfk_mha_fwd":
            B, Nq, Nk, H, D = a[5:10]
            return 4.0 * B * H * Nq * Nk * D, 4.0 * B * H * D * (2 * Nq + 2 * Nk)
        if name == "npfk_mha_bwd":
            B, Nq, Nk, H, D = a[10:15]
            # S^T, dP^T, dV, dK, dQ: five GEMMs of 2*Nq*Nk*D each per (batch, head)
            return 10.0 * B * H * Nq * Nk * D, 4.0 * B * H * D * (4 * Nq + 4 * Nk)
        if name == "npfk_add_layernorm_fwd":
            rows, C = a[7], a[8]
            return 0.0, 4.0 * rows * C * (3 if a[1] else 2)
        if name == "npfk_add_layernorm_bwd":
            rows, C = a[10], a[11]
            return 0.0, 4.0 * rows * C * (3 + (1 if a[3] else 0) + (1 if a[1] else 0))
        if name == "npfk_gauss_nll_fwd":
            Z, B, T, yd = a[3:7]
            return 0.0, 4.0 * (Z * B * T * 2 * yd + B * T * yd)
        if name == "npfk_gauss_nll_bwd":
            Z, B, T, yd = a[4:8]
            return 0.0, 4.0 * (Z * B * T * 4 * yd + B * T * yd)
        if name == "npfk_setconv_interp_fwd":
            B, Nk, Nq, C = a[10:14]
            return 0.0, 4.0 * B * (Nq * C + Nk * C + 3 * Nq)
        if name == "npfk_setconv_interp_bwd":
            B, Nk, Nq, C = a[13:17]
            return 0.0, 4.0 * B * (Nq * C + 2 * Nk * C + 3 * Nq)
        return 0.0, 0.0

    def __enter__(self):
        self._orig = LIB.call

        def timed(name, *args):
            if name not in self.names:
                return self._orig(name, *args)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            self._orig(name, *args)
            e1.record()
            flops, nbytes = self._work(name, args)
            self.records.append((name, flops, nbytes, e0, e1))

        LIB.call = timed
        return self

    def __exit__(self, *exc):
        LIB.call = self._orig

    def summary(self):
        """{name: dict(calls, ms_total, ms_avg, flops_total, bytes_total)} -- call after torch.cuda.synchronize()"""
        out = {}
        for name, flops, nbytes, e0, e1 in self.records:
            d = out.setdefault(name, dict(calls=0, ms_total=0.0, flops_total=0.0, bytes_total=0.0))
            d["calls"] += 1
            d["ms_total"] += e0.elapsed_time(e1)
            d["flops_total"] += flops
            d["bytes_total"] += nbytes
        for d in out.values():
            d["ms_avg"] = d["ms_total"] / max(d["calls"], 1)
        return out
