"""MLP and the two-input wrappers used as XEncoder / XYEncoder / Decoder / LatentEncoder plug-ins.

Same constructor signatures, attribute names and ``state_dict`` layout as the reference
(npf/architectures/mlp.py:12-115, npf/architectures/encoders.py:105-213); the arithmetic runs in the
npfk kernels (fused bias + ReLU (+ residual) epilogues) instead of addmm + separate elementwise kernels.
"""
import os
import warnings

import torch
import torch.nn as nn

from .. import kernels as K
from ..utils.initialization import hidden_linear_init_, linear_init, output_linear_init_, weights_init

__all__ = ["MLP", "MergeFlatInputs", "merge_flat_input", "DiscardIthArg", "discard_ith_arg"]

_ONE_CHAIN_MERGE = os.environ.get("NPFK_ONE_CHAIN_MERGE", "1") != "0"   # developer switch for A/B timing


class MLP(nn.Module):
    """to_hidden -> act -> [linears -> act]* -> out.

    Parameters mirror the reference: input_size, output_size, hidden_size=32, n_hidden_layers=1,
    activation=nn.ReLU(), is_bias=True, dropout=0, is_force_hid_smaller=False, is_res=False.
    """

    def __init__(self, input_size, output_size, hidden_size=32, n_hidden_layers=1, activation=nn.ReLU(),
                 is_bias=True, dropout=0, is_force_hid_smaller=False, is_res=False):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.n_hidden_layers, self.is_res = n_hidden_layers, is_res
        lo, hi = min(output_size, input_size), max(output_size, input_size)
        if is_force_hid_smaller and hidden_size > hi:
            warnings.warn(f"hidden_size={hidden_size} larger than output={output_size} and input={input_size}. "
                          f"Setting it to {hi}.")
            hidden_size = hi
        elif hidden_size < lo:
            warnings.warn(f"hidden_size={hidden_size} smaller than output={output_size} and input={input_size}. "
                          f"Setting it to {lo}.")
            hidden_size = lo
        self.hidden_size = hidden_size
        self.dropout = nn.Dropout(p=dropout) if dropout > 0 else nn.Identity()
        self.activation = activation
        self.to_hidden = nn.Linear(input_size, hidden_size, bias=is_bias)
        self.linears = nn.ModuleList(nn.Linear(hidden_size, hidden_size, bias=is_bias)
                                     for _ in range(n_hidden_layers - 1))
        self.out = nn.Linear(hidden_size, output_size, bias=is_bias)
        self.reset_parameters()

    def reset_parameters(self):
        # see utils/initialization.py for why hidden and output layers differ
        if isinstance(self.activation, nn.ReLU):
            hidden_linear_init_(self.to_hidden)
            for lin in self.linears:
                hidden_linear_init_(lin)
        else:
            linear_init(self.to_hidden, activation=self.activation)
            for lin in self.linears:
                linear_init(lin, activation=self.activation)
        output_linear_init_(self.out)

    def _hidden(self, layer, x):
        if isinstance(self.activation, nn.ReLU):
            return K.linear(x, layer.weight, layer.bias, relu=True)
        return self.activation(K.linear(x, layer.weight, layer.bias))   # user activation: module composition

    def is_plain_relu_chain(self):
        """True when the whole MLP can run as one K.mlp_chain node (ReLU, no dropout, no residual links)."""
        return isinstance(self.activation, nn.ReLU) and isinstance(self.dropout, nn.Identity) and not self.is_res

    def layer_list(self):
        """[(weight, bias), ...] in forward order (what kernels.mlp_chain takes)"""
        layers = [(self.to_hidden.weight, self.to_hidden.bias)]
        layers += [(lin.weight, lin.bias) for lin in self.linears]
        layers.append((self.out.weight, self.out.bias))
        return layers

    def forward(self, x, res=None, out_relu=False, x_is_relu=False, dy_premasked=False, nll=None, res_stash=None):
        """``res`` / ``out_relu`` fuse ``relu(res + MLP(x))`` (the sum-merge) into the last layer's epilogue.
        ``x_is_relu`` / ``dy_premasked``: see kernels._MlpChainFn (ReLU backward fused across the sum-merge).
        ``res_stash``: a kernels.GradStash that takes the gradient of ``res`` instead of autograd (plain ReLU chains only).
        ``nll=(y, T)`` (decoders only; see ``accepts_nll``): returns ``(out, sum_lp or None)``, the Gaussian
        log-likelihood of y under the predictive distribution summed over the T rows of each task, evaluated in the
        chain kernel's epilogue."""
        if self.is_plain_relu_chain():
            return K.mlp_chain(x, self.layer_list(), res=res, out_relu=out_relu, x_is_relu=x_is_relu, dy_premasked=dy_premasked,
                               nll=nll, res_stash=res_stash)
        assert not (x_is_relu or dy_premasked or nll is not None or res_stash is not None)
        h = self.dropout(self._hidden(self.to_hidden, x))
        for lin in self.linears:
            o = self._hidden(lin, h)
            if self.is_res:
                o = o + h
            h = self.dropout(o)
        return K.linear(h, self.out.weight, self.out.bias, res=res, relu=out_relu)


class DiscardIthArg(nn.Module):
    """Drops the i-th positional argument of the constructor and of forward before delegating to ``To``."""

    def __init__(self, *args, i=0, To=nn.Identity, **kwargs):
        super().__init__()
        self.i = i
        self.destination = To(*self.filter_args(*args), **kwargs)

    def filter_args(self, *args):
        return [a for j, a in enumerate(args) if j != self.i]

    @property
    def accepts_nll(self):
        return type(self.destination) is MLP and self.destination.is_plain_relu_chain()

    def forward(self, *args, **kwargs):
        args = self.filter_args(*args)
        if type(self.destination) is nn.Linear and len(args) == 1 and not kwargs and args[0].is_cuda:
            # ConvLNP's default decoder (convnp.py:250): a plain Linear runs in the npfk kernels too
            return K.linear(args[0], self.destination.weight, self.destination.bias)
        return self.destination(*args, **kwargs)


def discard_ith_arg(module, i, **kwargs):
    def make(*args, **kwargs2):
        return DiscardIthArg(*args, i=i, To=module, **kwargs, **kwargs2)

    return make


def _is_broadcast_over(x2, dim):
    return x2.shape[dim] == 1 or x2.stride(dim) == 0


class MergeFlatInputs(nn.Module):
    """FlatModule over two flat inputs: concatenation, or ``relu(x1 + MLP(x2))`` when ``is_sum_merge``."""

    def __init__(self, FlatModule, x1_dim, x2_dim, n_out, is_sum_merge=False, **kwargs):
        super().__init__()
        self.is_sum_merge = is_sum_merge
        if is_sum_merge:
            dim = x1_dim
            self.resizer = MLP(x2_dim, dim)
        else:
            dim = x1_dim + x2_dim
        self.flat_module = FlatModule(dim, n_out, **kwargs)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    @property
    def accepts_nll(self):
        """the decode tail (head + log-likelihood) can ride in flat_module's fused chain: forward(..., nll=(y, T))"""
        return type(self.flat_module) is MLP and self.flat_module.is_plain_relu_chain()

    def _sum_merge(self, x1, x2, nll=None):
        # Case A: x2 is one vector per (sample, task) replicated over the set dimension (CNP / LNP decoders,
        # reference np.py:103-110 expands it).  Run the resizer once per task and broadcast in the merge kernel.
        if x2.dim() == 4 and x1.dim() == 3 and _is_broadcast_over(x2, 2) and x1.shape[1] != 1:
            r = self.resizer(x2[:, :, 0, :])                       # [Z,B,C]
            return K.bcast_add_relu(x1, r)     # (the caller applies flat_module, with nll when asked)
        # Case B: same rows (x2 possibly with a leading n_z dim that x1 is tiled over): fuse into the epilogue
        if x2.shape[-x1.dim():-1] == x1.shape[:-1] and x2.dim() in (x1.dim(), x1.dim() + 1):
            if self._one_chain():
                # resizer and flat module as ONE chain node: relu(x1 + resizer(x2)) is a hidden layer with a residual, the
                # merged activation is written once (for backward) and never read back in forward; in fast mode a single
                # kernel launch (K1) up to the head and the log-likelihood tail
                layers = self.resizer.layer_list()
                return K.mlp_chain(x2, layers + self.flat_module.layer_list(), res=x1, res_layer=len(layers) - 1,
                                   nll=nll), True
            if self._fused_relu_bwd():
                # merged feeds only flat_module, whose first dgrad masks with merged > 0: the resizer skips its own
                merged = self.resizer(x2, res=x1, out_relu=True, dy_premasked=True)
                return self.flat_module(merged, x_is_relu=True, nll=nll), True
            return self.resizer(x2, res=x1, out_relu=True)
        return torch.relu(x1 + self.resizer(x2))                     # exotic broadcasting: plain composition

    def _one_chain(self):
        fm = self.flat_module
        return (_ONE_CHAIN_MERGE and type(fm) is MLP and fm.is_plain_relu_chain() and type(self.resizer) is MLP
                and self.resizer.is_plain_relu_chain())

    def _fused_relu_bwd(self):
        fm = self.flat_module
        return (type(fm) is MLP and fm.is_plain_relu_chain() and self.resizer.is_plain_relu_chain()
                and torch.is_grad_enabled())

    def forward(self, x1, x2, nll=None):
        if self.is_sum_merge:
            merged = self._sum_merge(x1, x2, nll)
            if isinstance(merged, tuple):
                return merged[0]
        else:
            if x2.dim() == x1.dim() + 1:
                x1 = x1.unsqueeze(0).expand(x2.shape[0], *x1.shape)
            merged = torch.cat((x1, x2), dim=-1)
        if nll is not None:
            return self.flat_module(merged, nll=nll)
        return self.flat_module(merged)


def merge_flat_input(module, is_sum_merge=False, **kwargs):
    def make(x_shape, flat_dim, n_out, **kwargs2):
        assert isinstance(x_shape, int)
        return MergeFlatInputs(module, x_shape, flat_dim, n_out, is_sum_merge=is_sum_merge, **kwargs2, **kwargs)

    return make
