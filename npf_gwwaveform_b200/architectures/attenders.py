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
_FUSE_LN2_RESIDUAL = os.environ.get("NPFK_LN2_RESIDUAL", "1") != "0"   # developer switch for A/B timing

__all__ = ["get_attender", "DotAttender", "MultiheadAttender", "TransformerAttender", "ScoreAttender",
           "MultiplicativeAttender", "AdditiveAttender", "CosineAttender", "DistanceAttender"]


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
    # the remaining registry names (attention.py:60-86) are not on the waveform hot path: they run as plain PyTorch
    # module composition on the GPU (plug-in row), with the reference's module layouts
    if name == "multiplicative":
        return MultiplicativeAttender(kq_size, value_size, out_size, **kwargs)
    if name == "additive":
        return AdditiveAttender(kq_size, value_size, out_size, **kwargs)
    if name == "cosine":
        return CosineAttender(kq_size, value_size, out_size, **kwargs)
    if name == "manhattan":
        return DistanceAttender(kq_size, value_size, out_size, p=1, **kwargs)
    if name == "euclidean":
        return DistanceAttender(kq_size, value_size, out_size, p=2, **kwargs)
    if name == "weighted_dist":
        return DistanceAttender(kq_size, value_size, out_size, is_weight=True, p=1, **kwargs)
    raise ValueError("Unknown attention method {}".format(attention))


class ScoreAttender(nn.Module):
    """Generic attender (attention.py:92-170): ``context = dropout(attn(score(keys, queries))) @ values`` with
    ``attn`` = softmax over keys when ``is_normalize`` else the raw logits, plus a resizer when value_size != out_size.
    Subclasses give ``score(keys, queries) -> [B, Nq, Nk]``.  Plain PyTorch ops: this is the module-composition path for
    everything the fused kernel does not recognise."""

    def __init__(self, kq_size, value_size, out_size, is_normalize=True, dropout=0):
        super().__init__()
        self.kq_size, self.value_size, self.out_size = kq_size, value_size, out_size
        self.is_normalize = is_normalize
        self.dropout = nn.Dropout(p=dropout) if dropout > 0 else nn.Identity()
        self.is_resize = value_size != out_size
        if self.is_resize:
            self.resizer = nn.Linear(value_size, out_size)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def score(self, keys, queries, **kwargs):
        raise NotImplementedError

    def forward(self, keys, queries, values, **kwargs):
        logits = self.score(keys, queries, **kwargs)
        attn = logits.softmax(dim=-1) if self.is_normalize else logits
        context = torch.bmm(self.dropout(attn), values)
        return self.resizer(context) if self.is_resize else context


def _dot_score(keys, queries, is_scale):
    logits = torch.einsum("bkd,bqd->bqk", keys, queries)
    return logits / math.sqrt(queries.size(-1)) if is_scale else logits


class MultiplicativeAttender(ScoreAttender):
    """Luong attention (attention.py:223-254): ``k . (W q)``; the inner ``dot`` keeps the reference's module tree."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.linear = nn.Linear(self.kq_size, self.kq_size, bias=False)
        self.dot = DotAttender(*args, is_scale=False)
        self.reset_parameters()

    def score(self, keys, queries, **kwargs):
        return _dot_score(keys, self.linear(queries), False)


class AdditiveAttender(ScoreAttender):
    """Bahdanau attention (attention.py:257-292): a tanh MLP over cat(k, q) for every (query, key) pair."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.mlp = MLP(self.kq_size * 2, 1, hidden_size=self.kq_size, activation=nn.Tanh())
        self.reset_parameters()

    def score(self, keys, queries, **kwargs):
        B, Nq, D = queries.shape
        Nk = keys.size(1)
        # kept exactly as the reference has it (attention.py:284-290): the queries are expanded along dim 1 too, so the
        # expression only type-checks when Nq == Nk and raises the same RuntimeError otherwise
        k = keys.unsqueeze(1).expand(B, Nq, Nk, D)
        q = queries.unsqueeze(1).expand(B, Nq, Nk, D)
        return self.mlp(torch.cat((k, q), dim=-1)).squeeze(-1)


class CosineAttender(ScoreAttender):
    """attention.py:295-311.  The reference reshapes with ``view`` (no transpose) before ``CosineSimilarity(dim=1)``;
    the same memory reinterpretation is kept so the numbers agree."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.similarity = nn.CosineSimilarity(dim=1)

    def score(self, keys, queries, **kwargs):
        B, Nq, D = queries.size()
        Nk = keys.size(1)
        return self.similarity(keys.contiguous().view(B, D, 1, Nk), queries.contiguous().view(B, D, Nq, 1))


class DistanceAttender(ScoreAttender):
    """negative squared p-norm of (optionally dimension-wise weighted) k - q (attention.py:314-372)"""

    def __init__(self, *args, p=1, is_weight=False, **kwargs):
        super().__init__(*args, **kwargs)
        self.p, self.is_weight = p, is_weight
        if is_weight:
            self.weighter = nn.Linear(self.kq_size, self.kq_size)
        self.reset_parameters()

    def score(self, keys, queries, **kwargs):
        diff = keys.unsqueeze(1) - queries.unsqueeze(2)
        if self.is_weight:
            diff = self.weighter(diff)
        return -torch.norm(diff, p=self.p, dim=-1) ** 2


class DotAttender(nn.Module):
    """softmax(q k^T / sqrt(d)) v  (single head), optional resizer when value_size != out_size."""

    def __init__(self, kq_size, value_size, out_size, is_scale=True, is_normalize=True, dropout=0):
        super().__init__()
        self.kq_size, self.value_size, self.out_size = kq_size, value_size, out_size
        self.is_scale, self.is_normalize = is_scale, is_normalize
        self.dropout = nn.Dropout(p=dropout) if dropout > 0 else nn.Identity()
        self.is_resize = value_size != out_size
        if self.is_resize:
            self.resizer = nn.Linear(value_size, out_size)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def _fusable(self, keys, values):
        """softmax-normalised, no attention dropout, value width == key width: what the flash kernel computes"""
        return (self.is_normalize and isinstance(self.dropout, nn.Identity) and values.shape[-1] == keys.shape[-1]
                and keys.is_cuda)

    def score(self, keys, queries):
        return _dot_score(keys, queries, self.is_scale)

    def attend(self, keys, queries, values, n_heads=1, delta_stash=None):
        d = queries.shape[-1] // n_heads
        if self._fusable(keys, values):
            scale = 1.0 / math.sqrt(d) if self.is_scale else 1.0
            return K.mha(queries, keys, values, n_heads, scale, delta_stash)
        # module composition (attention.py:129-164 with the head split of :507-527): is_normalize=False, attention
        # dropout, or value heads of another width than the key heads
        B, Nq, Nk = queries.shape[0], queries.shape[1], keys.shape[1]
        dv = values.shape[-1] // n_heads
        qh = queries.reshape(B, Nq, n_heads, d).transpose(1, 2)
        kh = keys.reshape(B, Nk, n_heads, d).transpose(1, 2)
        vh = values.reshape(B, Nk, n_heads, dv).transpose(1, 2)
        logits = qh @ kh.transpose(-1, -2)
        if self.is_scale:
            logits = logits / math.sqrt(d)
        attn = self.dropout(logits.softmax(dim=-1) if self.is_normalize else logits)
        return (attn @ vh).transpose(1, 2).reshape(B, Nq, n_heads * dv)

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
        if _FUSE_LN2_RESIDUAL and type(self.mlp) is MLP and self.mlp.is_plain_relu_chain():
            # c + MLP(c) leaves the MLP's last layer (residual added in its epilogue) instead of being formed by LN2 from
            # two tensors: LN2 reads one tensor instead of two, forward and backward.  The same bits (one fp32 addition of
            # the same two operands); the gradient of the residual use of c goes through the stash as before.
            return K.add_layernorm(self.mlp(c, res=c, res_stash=stash), None, ln2.weight, ln2.bias, ln2.eps)
        return K.add_layernorm(c, self.mlp(c), ln2.weight, ln2.bias, ln2.eps, park_a=stash)
