"""SetConv and its radial basis functions (``Interpolator(x_dim, in_dim, out_dim)`` plug-in).

Mirrors npf/architectures/setcnn.py: ``ExpRBF`` (:86-142), the fork's ``UnsharedExpRBF`` (:14-83) and
``SetConv`` (:194-268).  The [B,Nq,Nk,C] weight*value product of the reference is never materialised.
"""
import math

import torch
import torch.nn as nn

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


class UnsharedExpRBF(ExpRBF):
    """The fork's variant: separate length scales for weights and density, weights = exp / (density + 1e-8)."""

    n_scales = 2

    def __init__(self, x_dim, max_dist=1 / 256, max_dist_weight=0.99, p=2, **kwargs):
        super().__init__(x_dim, max_dist=max_dist, max_dist_weight=max_dist_weight, p=p, **kwargs)


class SetConv(nn.Module):
    """resizer(cat(sum_k w(x_q - x_k) v_k, density)); keys/queries are 1-D positions."""

    def __init__(self, x_dim, in_channels, out_channels, RadialBasisFunc=ExpRBF, **kwargs):
        super().__init__()
        assert x_dim == 1, "Currently only supports single spatial dimension `x_dim==1`"
        self.radial_basis_func = RadialBasisFunc(x_dim, **kwargs)
        if not isinstance(self.radial_basis_func, ExpRBF):
            raise NotImplementedError("the fused SetConv kernel implements ExpRBF and UnsharedExpRBF")
        self.resizer = nn.Linear(in_channels + 1, out_channels)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def forward(self, keys, queries, values):
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
