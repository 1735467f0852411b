"""Small host-side helpers that the reference exposes from npf.utils.helpers and that sit on the 1-D hot path."""
import operator
from functools import reduce

import torch
from torch.distributions import Normal
from torch.distributions.independent import Independent

__all__ = ["MultivariateNormalDiag", "sum_from_nth_dim", "isin_range", "channels_to_2nd_dim", "channels_to_last_dim",
           "prod", "rescale_range", "make_depth_sep_conv"]


def MultivariateNormalDiag(loc, scale_diag):
    """Independent(Normal(loc, scale), 1) -- reference npf/utils/helpers.py:125-129.

    Argument validation is switched off: it would force a device->host sync on every forward.
    """
    if loc.dim() < 1:
        raise ValueError("loc must be at least one-dimensional.")
    return Independent(Normal(loc, scale_diag, validate_args=False), 1, validate_args=False)


def sum_from_nth_dim(t, dim):
    """Sum all dims from `dim` on (npf/utils/helpers.py:15-17)."""
    return t.reshape(*t.shape[:dim], -1).sum(-1)


def isin_range(x, valid_range):
    """Elementwise range test reduced with all() (npf/utils/helpers.py:55-57); host/torch version for user code."""
    return ((x >= valid_range[0]) & (x <= valid_range[1])).all()


def channels_to_2nd_dim(X):
    return X.permute(*([0, X.dim() - 1] + list(range(1, X.dim() - 1))))


def channels_to_last_dim(X):
    return X.permute(*([0] + list(range(2, X.dim())) + [1]))


def prod(iterable):
    return reduce(operator.mul, iterable, 1)


def rescale_range(X, old_range, new_range):
    old_min, new_min = old_range[0], new_range[0]
    return (((X - old_min) * (new_range[1] - new_min)) / (old_range[1] - old_min)) + new_min


def make_depth_sep_conv(Conv):
    """Factory kept for API compatibility (npf/utils/helpers.py:354-403); builds the B200 depth-separable conv."""
    from ..architectures.convnets import DepthSepConv1d

    def build(in_channels, out_channels, kernel_size, bias=True, **kwargs):
        return DepthSepConv1d(in_channels, out_channels, kernel_size, bias=bias, Conv=Conv, **kwargs)

    return build
