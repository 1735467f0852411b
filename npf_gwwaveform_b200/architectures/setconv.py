"""SetConv and its radial basis functions (``Interpolator(x_dim, in_dim, out_dim)`` plug-in).

Mirrors npf/architectures/setcnn.py: ``ExpRBF`` (:86-142), the fork's ``UnsharedExpRBF`` (:14-83) and
``SetConv`` (:194-268).  The [B,Nq,Nk,C] weight*value product of the reference is never materialised.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import kernels as K
from ..utils.initialization import weights_init

__all__ = ["SetConv", "ExpRBF", "UnsharedExpRBF"]


def _length_scale_init(max_dist, max_dist_weight, p):
    # exp(-(max_dist/sigma)^p) = max_dist_weight  =>  sigma ; parameter = inverse softplus of sigma
    sigma = max_dist / ((-math.log(max_dist_weight)) ** (1 / p))
    return math.log(math.exp(sigma) - 1)


class ExpRBF(nn.Module):
    """exp(-(dist/sigma)^p), weights normalised by a softmax over keys; sigma = 1e-5 + softplus(param)."""

    n_scales = 1

    def __init__(self, x_dim, max_dist=1 / 256, max_dist_weight=0.9, p=2, **kwargs):
        super().__init__()
        self.max_dist, self.max_dist_weight, self.p = max_dist, max_dist_weight, p
        self.length_scale_param = nn.Parameter(torch.zeros(self.n_scales))
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)
        v = _length_scale_init(self.max_dist, self.max_dist_weight, self.p)
        self.length_scale_param = nn.Parameter(torch.tensor([v] * self.n_scales))

    def sigma(self):
        return 1e-5 + F.softplus(self.length_scale_param)

    def forward(self, diff):
        """The reference's plug-in contract (setcnn.py:126-142): ``diff[B, Nq, Nk, x_dim] -> (weight[B, Nq, Nk, 1],
        density[B, Nq, 1])``.  SetConv does not call this for ExpRBF / UnsharedExpRBF (the fused kernels evaluate the same
        expression without materialising [B, Nq, Nk, 1]); it exists so that code written against the reference's
        RadialBasisFunc interface, or a subclass overriding it, keeps working as ordinary module composition."""
        logit = -(torch.norm(diff, p=self.p, dim=-1, keepdim=True) / self.sigma()).pow(self.p)
        return torch.softmax(logit, dim=-2), torch.exp(logit).sum(dim=-2)


class UnsharedExpRBF(ExpRBF):
    """The fork's variant: separate length scales for weights and density, weights = exp / (density + 1e-8)."""

    n_scales = 2

    def __init__(self, x_dim, max_dist=1 / 256, max_dist_weight=0.99, p=2, **kwargs):
        super().__init__(x_dim, max_dist=max_dist, max_dist_weight=max_dist_weight, p=p, **kwargs)

    def forward(self, diff):
        """setcnn.py:65-83: channel 0 of exp(-(d/sigma)^p) uses sigma[0] and is divided by the density, which is the
        key-sum of channel 1 (sigma[1])."""
        e = torch.exp(-(torch.norm(diff, p=self.p, dim=-1, keepdim=True) / self.sigma()).pow(self.p))
        density = e[..., 1:].sum(dim=-2)
        return e[..., 0:1] / (density.unsqueeze(2) + 1e-8), density


class SetConv(nn.Module):
    """resizer(cat(sum_k w(x_q - x_k) v_k, density)); keys/queries are 1-D positions."""

    def __init__(self, x_dim, in_channels, out_channels, RadialBasisFunc=ExpRBF, **kwargs):
        super().__init__()
        assert x_dim == 1, "Currently only supports single spatial dimension `x_dim==1`"
        self.radial_basis_func = RadialBasisFunc(x_dim, **kwargs)
        self.resizer = nn.Linear(in_channels + 1, out_channels)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def _fused(self):
        """the kernels evaluate exactly ExpRBF.forward / UnsharedExpRBF.forward; anything else (a user RadialBasisFunc, or
        a subclass that overrides ``forward``) runs as module composition"""
        t = type(self.radial_basis_func)
        return t is ExpRBF or t is UnsharedExpRBF or (
            isinstance(self.radial_basis_func, ExpRBF) and t.forward in (ExpRBF.forward, UnsharedExpRBF.forward))

    def forward_generic(self, keys, queries, values):
        """setcnn.py:254-268 as plain PyTorch: any ``RadialBasisFunc.forward(diff) -> (weight, density)``"""
        weight, density = self.radial_basis_func(keys.unsqueeze(1) - queries.unsqueeze(2))
        targets = (weight * values.unsqueeze(1)).sum(dim=2)
        return self.resizer(torch.cat([targets, density], dim=-1))

    def forward(self, keys, queries, values):
        if not self._fused():
            return self.forward_generic(keys, queries, values)
        rbf = self.radial_basis_func
        n_in = values.shape[-1]
        if queries.shape[-2] > keys.shape[-2] and self.resizer.in_features == n_in + 1:
            # many queries per key (grid -> targets): the resizer is linear, so apply it to the keys' values and
            # interpolate the result; out = sum_k w_qk (W_T v_k) + density_q * w_d + b is written once per query
            w = self.resizer.weight
            v = K.linear(values, w[:, :n_in])
            return K.setconv_interp(keys, queries, v, rbf.length_scale_param, w[:, n_in], self.resizer.bias,
                                    float(rbf.p), rbf.n_scales == 2)
        t = K.setconv_rbf(keys, queries, values, rbf.length_scale_param, float(rbf.p), rbf.n_scales == 2)
        return K.linear(t, self.resizer.weight, self.resizer.bias)
