"""Cross-attention plug-ins (``Attender(kq_size, value_size, out_size, **kw)``).

Mirrors the reference's attender registry and module layouts (npf/architectures/attention.py:16-86,
172-220, 375-588).  "scaledot", "multihead" and "transformer" run on the fused flash-style npfk kernel;
heads are read in place (no permute copies, no [B*H, Nq, Nk] logits in HBM).
"""
import math
import os

import torch
import torch.nn as nn

from .. import kernels as K
from ..utils.initialization import weights_init
from .blocks import MLP

_USE_STASH = os.environ.get("NPFK_LN_STASH", "1") != "0"   # developer switch for A/B timing

__all__ = ["get_attender", "DotAttender", "MultiheadAttender", "TransformerAttender"]


def get_attender(attention, kq_size, value_size, out_size, **kwargs):
    """String -> attender module; a callable is instantiated as ``attention(kq_size, value_size, out_size)``."""
    if not isinstance(attention, str):
        return attention(kq_size, value_size, out_size, **kwargs)
    name = attention.lower()
    if name == "scaledot":
        return DotAttender(kq_size, value_size, out_size, is_scale=True, **kwargs)
    if name == "multihead":
        return MultiheadAttender(kq_size, value_size, out_size, **kwargs)
    if name == "transformer":
        return TransformerAttender(kq_size, value_size, out_size, **kwargs)
    if name in ("multiplicative", "additive", "cosine", "manhattan", "euclidean", "weighted_dist"):
        raise NotImplementedError(
            f"attention={attention!r} is outside the B200 hot path (no 1-D waveform config uses it); "
            "pass a callable Attender to plug in your own module."
        )
    raise ValueError("Unknown attention method {}".format(attention))


class DotAttender(nn.Module):
    """softmax(q k^T / sqrt(d)) v  (single head), optional resizer when value_size != out_size."""

    def __init__(self, kq_size, value_size, out_size, is_scale=True, is_normalize=True, dropout=0):
        super().__init__()
        if not is_normalize:
            raise NotImplementedError("is_normalize=False is not served by the fused attention kernel")
        self.kq_size, self.value_size, self.out_size = kq_size, value_size, out_size
        self.is_scale, self.is_normalize = is_scale, is_normalize
        self.dropout = nn.Dropout(p=dropout) if dropout > 0 else nn.Identity()
        if dropout > 0:
            raise NotImplementedError("attention dropout is not implemented in the fused kernel")
        self.is_resize = value_size != out_size
        if self.is_resize:
            self.resizer = nn.Linear(value_size, out_size)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def attend(self, keys, queries, values, n_heads=1, delta_stash=None):
        d = queries.shape[-1] // n_heads
        scale = 1.0 / math.sqrt(d) if self.is_scale else 1.0
        if values.shape[-1] != keys.shape[-1]:
            raise NotImplementedError("fused attention needs value_size == kq_size")
        return K.mha(queries, keys, values, n_heads, scale, delta_stash)

    def forward(self, keys, queries, values, **kwargs):
        ctx = self.attend(keys, queries, values)
        if self.is_resize:
            ctx = K.linear(ctx, self.resizer.weight, self.resizer.bias)
        return ctx


class MultiheadAttender(nn.Module):
    """K = W_k k, Q = W_q q + b, V = W_v v; n_heads heads of kq_size // n_heads; optional output Linear."""

    def __init__(self, kq_size, value_size, out_size, n_heads=8, is_post_process=True, dropout=0,
                 is_relative_pos=False):
        super().__init__()
        if is_relative_pos:
            raise NotImplementedError("relative positional encodings are outside the 1-D hot path")
        self.is_relative_pos = is_relative_pos
        self.key_transform = nn.Linear(kq_size, kq_size, bias=False)
        self.query_transform = nn.Linear(kq_size, kq_size, bias=True)
        self.value_transform = nn.Linear(value_size, value_size, bias=False)
        self.dot = DotAttender(kq_size, value_size, out_size, is_scale=True, dropout=dropout)
        self.n_heads = n_heads
        self.kq_head_size = kq_size // n_heads
        self.value_head_size = kq_size // n_heads
        self.kq_size, self.value_size, self.out_size = kq_size, value_size, out_size
        self.post_processor = nn.Linear(value_size, out_size) if (is_post_process or value_size != out_size) else None
        assert kq_size % n_heads == 0, "{} % {} != 0".format(kq_size, n_heads)
        assert value_size % n_heads == 0, "{} % {} != 0".format(value_size, n_heads)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)
        std = math.sqrt(2.0 / (self.kq_size + self.kq_head_size))
        nn.init.normal_(self.key_transform.weight, mean=0, std=std)
        nn.init.normal_(self.query_transform.weight, mean=0, std=std)
        std = math.sqrt(2.0 / (self.value_size + self.value_head_size))
        nn.init.normal_(self.value_transform.weight, mean=0, std=std)

    def forward(self, keys, queries, values, q_stash=None, delta_stash=None, **kwargs):
        k = K.linear(keys, self.key_transform.weight)
        q = K.linear(queries, self.query_transform.weight, self.query_transform.bias, dx_stash=q_stash)
        v = K.linear(values, self.value_transform.weight)
        ctx = self.dot.attend(k, q, v, n_heads=self.n_heads,
                              delta_stash=delta_stash if self.post_processor is None else None)
        if self.post_processor is not None:
            ctx = K.linear(ctx, self.post_processor.weight, self.post_processor.bias)
        return ctx


class TransformerAttender(MultiheadAttender):
    """Multi-head attention + ``LN(ctx + queries)`` + ``LN(c + MLP(c))``; no output Linear."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, is_post_process=False, **kwargs)
        assert self.kq_size == self.out_size
        self.layer_norm1 = nn.LayerNorm(self.out_size)
        self.layer_norm2 = nn.LayerNorm(self.out_size)
        self.mlp = MLP(self.out_size, self.out_size, hidden_size=self.out_size, activation=nn.ReLU())
        self.reset_parameters()

    def forward(self, keys, queries, values, **kwargs):
        # queries feed the query transform and LN1's residual: LN1's backward parks its gradient for them in `q_stash`
        # and the query transform's backward adds its dx onto that buffer (TMA reduce-add in the fused kernel)
        q_stash = K.GradStash() if torch.is_grad_enabled() and _USE_STASH else None
        # LN1's backward produces dO of the attention with O in registers: it also writes delta = sum_d dO*O
        d_stash = K.GradStash() if torch.is_grad_enabled() and _USE_STASH else None
        ctx = super().forward(keys, queries, values, q_stash=q_stash, delta_stash=d_stash, **kwargs)
        ln1, ln2 = self.layer_norm1, self.layer_norm2
        # c feeds LN2 directly and through the MLP: LN2's backward parks its gradient for c in `stash` and LN1's
        # backward adds it to the MLP's while reading it (no separate fan-out add of two [B, Nt, 128] tensors)
        stash = K.GradStash() if torch.is_grad_enabled() and _USE_STASH else None
        c = K.add_layernorm(ctx, queries, ln1.weight, ln1.bias, ln1.eps, extra=stash, park_b=q_stash,
                            attn_of=d_stash)   # residual on the raw queries
        return K.add_layernorm(c, self.mlp(c), ln2.weight, ln2.bias, ln2.eps, park_a=stash)
