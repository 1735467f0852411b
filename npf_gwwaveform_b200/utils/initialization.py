"""Parameter initialisation with the same *effective* distributions as the reference.

Reference quirks that are kept on purpose (npf/utils/initialization.py):
  * ``weights_init(module)`` never touches a module that itself defines ``reset_parameters`` (it tests the
    root, not the child, :15-22), so plain Linear / Conv layers keep PyTorch's default initialisation;
  * ``linear_init(layer, activation=nn.ReLU())`` resolves the activation to a *class* (:46-48), matches no
    branch, and therefore only zeroes the bias (:79-94);
  * ``linear_init(layer)`` (string default "relu") applies Kaiming-uniform and zeroes the bias.
"""
import torch
from torch import nn

__all__ = ["weights_init", "linear_init", "init_param_", "hidden_linear_init_", "output_linear_init_"]


def hidden_linear_init_(layer: nn.Linear) -> None:
    """What MLP does to its hidden layers: PyTorch-default weight, zero bias."""
    if layer.bias is not None:
        with torch.no_grad():
            layer.bias.zero_()


def output_linear_init_(layer: nn.Linear) -> None:
    """What MLP does to its output layer: Kaiming-uniform(relu) weight, zero bias."""
    if layer.bias is not None:
        with torch.no_grad():
            layer.bias.zero_()
    nn.init.kaiming_uniform_(layer.weight, nonlinearity="relu")


_NAME_OF = {nn.LeakyReLU: "leaky_relu", nn.ReLU: "relu", nn.Tanh: "tanh", nn.Sigmoid: "sigmoid", nn.Softmax: "sigmoid"}


def linear_init(module, activation="relu"):
    """API-compatible with the reference's ``linear_init`` (module instances as ``activation`` only zero the bias)."""
    if module.bias is not None:
        module.bias.data.zero_()
    if activation is None:
        return nn.init.xavier_uniform_(module.weight)
    if not isinstance(activation, str):
        if not any(isinstance(activation, k) for k in _NAME_OF):
            raise ValueError("Unkown given activation type : {}".format(activation))
        return None
    if activation == "leaky_relu":
        return nn.init.kaiming_uniform_(module.weight, a=0, nonlinearity="leaky_relu")
    if activation == "relu":
        return nn.init.kaiming_uniform_(module.weight, nonlinearity="relu")
    if activation in ("sigmoid", "tanh"):
        return nn.init.xavier_uniform_(module.weight, gain=nn.init.calculate_gain(activation))
    return None


def weights_init(module, **kwargs):
    """Marks the module as reset.  Descendants are only initialised when the root has no ``reset_parameters``."""
    module.is_resetted = True
    if hasattr(module, "reset_parameters"):
        return
    for m in module.modules():
        if isinstance(m, torch.nn.modules.conv._ConvNd):
            nn.init.kaiming_normal_(m.weight, mode="fan_out", **kwargs)
        elif isinstance(m, nn.Linear):
            linear_init(m, **kwargs)
        elif isinstance(m, nn.BatchNorm2d):
            m.weight.data.fill_(1)
            m.bias.data.zero_()


def init_param_(param, activation=None, is_positive=False, bound=0.05, shift=0):
    gain = 1 if activation is None else nn.init.calculate_gain(activation if isinstance(activation, str) else "relu")
    if is_positive:
        nn.init.uniform_(param, 1e-5 + shift, bound * gain + shift)
        return
    nn.init.uniform_(param, -bound * gain + shift, bound * gain + shift)
