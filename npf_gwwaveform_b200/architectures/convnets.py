"""1-D ResNet / UNet between the induced points of ConvCNP (``CNN(r_dim)`` plug-in, channel-last in and out).

Mirrors npf/architectures/cnn.py: ``ResConvBlock`` (:126-215), ``CNN`` (:307-380), ``UnetCNN`` (:383-516) for
``Conv=nn.Conv1d``.  All tensors stay channel-last [B, L, C]: BatchNorm + ReLU + depthwise conv (+ residual) is
one npfk kernel, 1x1 convolutions are row-major Linear layers, pool / upsample / concat are fused copies.
"""
import torch
import torch.nn as nn

from .. import kernels as K
from ..utils.initialization import weights_init

__all__ = ["DepthSepConv1d", "ResConvBlock", "CNN", "UnetCNN"]


def _check_conv(Conv):
    if Conv is not nn.Conv1d:
        raise NotImplementedError("the B200 hot path implements the 1-D models (Conv=nn.Conv1d) only")


def _norm_args(norm):
    """(gamma, beta, running_mean, running_var, eps, momentum) or Nones for Identity."""
    if isinstance(norm, nn.Identity):
        return None, None, None, None, 1e-5, 0.1
    if isinstance(norm, nn.BatchNorm1d) and norm.affine and norm.track_running_stats and norm.momentum is not None:
        return norm.weight, norm.bias, norm.running_mean, norm.running_var, norm.eps, norm.momentum
    raise NotImplementedError(f"Normalization {type(norm).__name__} is not implemented (BatchNorm1d or Identity)")


def _norm_relu_dwconv(x, norm, conv, res, training, relu=True):
    gamma, beta, rm, rv, eps, mom = _norm_args(norm)
    nbt = norm.num_batches_tracked if gamma is not None and training else None
    return K.norm_act_dwconv(x, gamma, beta, conv.weight, conv.bias, res, rm, rv, training, eps, mom, relu, nbt)


def _pointwise(x, conv):
    return K.linear(x, conv.weight, conv.bias)      # [N, K, 1] weight taken as is (no view node in the graph)


class DepthSepConv1d(nn.Module):
    """depthwise (groups = channels) followed by a pointwise 1x1 convolution; channel-last input."""

    def __init__(self, in_channels, out_channels, kernel_size, bias=True, Conv=nn.Conv1d, padding=None, **kwargs):
        super().__init__()
        _check_conv(Conv)
        padding = kernel_size // 2 if padding is None else padding
        if padding != kernel_size // 2 or kwargs:
            raise NotImplementedError("only 'same' zero padding, stride 1, dilation 1 are implemented")
        self.depthwise = nn.Conv1d(in_channels, in_channels, kernel_size, padding=padding, groups=in_channels, bias=bias)
        self.pointwise = nn.Conv1d(in_channels, out_channels, 1, bias=bias)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def forward(self, x, norm=None, training=False, relu=False):
        h = _norm_relu_dwconv(x, norm if norm is not None else nn.Identity(), self.depthwise, None, training, relu)
        return _pointwise(h, self.pointwise)


class ResConvBlock(nn.Module):
    """Pre-activation residual block with depth-separable convolutions (channel-last [B, L, C])."""

    def __init__(self, in_chan, out_chan, Conv, kernel_size=5, activation=nn.ReLU(), Normalization=nn.Identity,
                 is_bias=True, n_conv_layers=1):
        super().__init__()
        _check_conv(Conv)
        if not isinstance(activation, nn.ReLU):
            raise NotImplementedError("the fused conv block implements ReLU")
        self.activation = activation
        self.n_conv_layers = n_conv_layers
        assert self.n_conv_layers in [1, 2]
        if kernel_size % 2 == 0:
            raise ValueError("`kernel_size={}`, but should be odd.".format(kernel_size))
        padding = kernel_size // 2
        if self.n_conv_layers == 2:
            self.norm1 = Normalization(in_chan)
            self.conv1 = DepthSepConv1d(in_chan, in_chan, kernel_size, bias=is_bias, Conv=Conv)
        self.norm2 = Normalization(in_chan)
        self.conv2_depthwise = nn.Conv1d(in_chan, in_chan, kernel_size, padding=padding, groups=in_chan, bias=is_bias)
        self.conv2_pointwise = nn.Conv1d(in_chan, out_chan, 1, bias=is_bias)
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def forward(self, X):
        out = X
        if self.n_conv_layers == 2:
            out = self.conv1(out, norm=self.norm1, training=self.training, relu=True)
        out = _norm_relu_dwconv(out, self.norm2, self.conv2_depthwise, X, self.training)   # + residual
        return _pointwise(out, self.conv2_pointwise)


class CNN(nn.Module):
    """Sequential conv blocks; ``n_channels`` int or list of n_blocks+1 channel counts."""

    def __init__(self, n_channels, ConvBlock, n_blocks=3, is_chan_last=False, **kwargs):
        super().__init__()
        self.n_blocks = n_blocks
        self.is_chan_last = is_chan_last
        self.in_out_channels = self._get_in_out_channels(n_channels, n_blocks)
        self.conv_blocks = nn.ModuleList(ConvBlock(i, o, **kwargs) for i, o in self.in_out_channels)
        self.is_return_rep = False
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    def _get_in_out_channels(self, n_channels, n_blocks):
        chans = [n_channels] * (n_blocks + 1) if isinstance(n_channels, int) else list(n_channels)
        assert len(chans) == (n_blocks + 1), "{} != {}".format(len(chans), n_blocks + 1)
        return list(zip(chans, chans[1:]))

    def forward(self, X):
        if not self.is_chan_last:          # NCL callers: one transposition at the boundary
            X = X.permute(0, 2, 1).contiguous()
        X, representation = self.apply_convs(X)
        if not self.is_chan_last:
            X = X.permute(0, 2, 1).contiguous()
        if self.is_return_rep:
            return X, representation
        return X

    def apply_convs(self, X):
        for block in self.conv_blocks:
            X = block(X)
        return X, None


class UnetCNN(CNN):
    """UNet: down (block + AvgPool 2) x n//2, bottleneck, up (linear x2 upsample + skip concat + block) x n//2."""

    def __init__(self, n_channels, ConvBlock, Pool, upsample_mode, max_nchannels=256, pooling_size=2,
                 is_force_same_bottleneck=False, is_return_rep=False, **kwargs):
        self.max_nchannels = max_nchannels
        super().__init__(n_channels, ConvBlock, **kwargs)
        if Pool is not nn.AvgPool1d or pooling_size != 2 or upsample_mode != "linear":
            raise NotImplementedError("UnetCNN on B200 implements Pool=nn.AvgPool1d, pooling_size=2, upsample_mode='linear'")
        if is_force_same_bottleneck:
            raise NotImplementedError("is_force_same_bottleneck is outside the 1-D hot path")
        self.pooling_size = pooling_size
        self.pooling = Pool(self.pooling_size)
        self.upsample_mode = upsample_mode
        self.is_force_same_bottleneck = is_force_same_bottleneck
        self.is_return_rep = is_return_rep

    def apply_convs(self, X):
        n_down = self.n_blocks // 2
        skips = []
        for i in range(n_down):
            X = self.conv_blocks[i](X)
            skips.append(X)
            X = K.avgpool2(X)
        X = self.conv_blocks[n_down](X)
        representation = K.set_mean(X).squeeze(1) if self.is_return_rep else None
        for i in range(n_down + 1, self.n_blocks):
            X = K.upsample2_cat(X, skips[n_down - i])
            X = self.conv_blocks[i](X)
        return X, representation

    def _get_in_out_channels(self, n_channels, n_blocks):
        assert n_blocks % 2 == 1, "n_blocks={} not odd".format(n_blocks)
        half = [2 ** i * n_channels for i in range(n_blocks // 2 + 1)]
        chans = half + half[::-1]
        chans = chans[:1] + [min(c, self.max_nchannels) for c in chans[1:-1]] + chans[-1:]
        pairs = super()._get_in_out_channels(chans, n_blocks)
        first_up = len(pairs) // 2 + 1
        return pairs[:first_up] + [(i * 2, o) for i, o in pairs[first_up:]]   # skip concat doubles the inputs
