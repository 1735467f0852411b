"""Losses of the neural-process family (drop-in for npf/losses.py: CNPFLoss, ELBOLossLNPF, NLLLossLNPF).

``criterion(pred_outputs, Y_trgt)`` with ``pred_outputs = model(...)``.  When the predictive distribution was
produced by this package's decode() the Gaussian log-likelihood and its sum over targets run as one kernel
straight from the decoder output (npfk_gauss_nll); any other distribution goes through ``log_prob``.
"""
import abc
import math

import torch
import torch.nn as nn
from torch.distributions.kl import kl_divergence

from . import kernels as K
from .utils.helpers import sum_from_nth_dim

__all__ = ["CNPFLoss", "ELBOLossLNPF", "NLLLossLNPF", "SUMOLossLNPF", "sum_log_prob"]


def sum_log_prob(prob, sample):
    """log-probability summed over everything but (n_z, batch): [n_z, B] (npf/losses.py:18-24)."""
    pre = getattr(prob, "_npfk_sum_lp", None)
    if pre is not None and sample.is_cuda and pre[1:] == (sample.data_ptr(), sample._version, tuple(sample.shape)):
        return pre[0]          # already reduced in the decoder kernel's epilogue, for exactly this tensor
    suff = getattr(prob, "_npfk_suffstat", None)
    if suff is not None and sample.is_cuda:
        return K.gauss_sum_log_prob(suff, sample)
    return sum_from_nth_dim(prob.log_prob(sample), 2)


def _kl_sum(q, p):
    """sum over latents of KL(q || p) -> [B]; fused for diagonal Gaussians."""
    bq, bp = getattr(q, "base_dist", None), getattr(p, "base_dist", None)
    if bq is not None and bp is not None and all(hasattr(b, a) for b in (bq, bp) for a in ("loc", "scale")) \
            and bq.loc.is_cuda:
        return sum_from_nth_dim(K.kl_diag_normal(bq.loc, bq.scale, bp.loc, bp.scale), 1)
    return sum_from_nth_dim(kl_divergence(q, p), 1)


class BaseLossNPF(nn.Module, abc.ABC):
    """reduction in {None, "mean", "sum"}; eval mode always evaluates the NPML / logsumexp estimator."""

    def __init__(self, reduction="mean", is_force_mle_eval=True):
        super().__init__()
        self.reduction = reduction
        self.is_force_mle_eval = is_force_mle_eval

    def forward(self, pred_outputs, Y_trgt):
        p_yCc, z_samples, q_zCc, q_zCct = pred_outputs
        if self.training:
            loss = self.get_loss(p_yCc, z_samples, q_zCc, q_zCct, Y_trgt)
        else:
            if self.is_force_mle_eval:
                q_zCct = None
            loss = NLLLossLNPF.get_loss(self, p_yCc, z_samples, q_zCc, q_zCct, Y_trgt)
        if self.reduction is None:
            return loss
        if self.reduction == "mean":
            return loss.mean(0)
        if self.reduction == "sum":
            return loss.sum(0)
        raise ValueError(f"Unknown {self.reduction}")

    @abc.abstractmethod
    def get_loss(self, p_yCc, z_samples, q_zCc, q_zCct, Y_trgt):
        """-> loss per task, [B]"""


class CNPFLoss(BaseLossNPF):
    """Negative log-likelihood of the conditional family (npf/losses.py:112-123)."""

    def get_loss(self, p_yCc, _, q_zCc, ___, Y_trgt):
        assert q_zCc is None
        return -sum_log_prob(p_yCc, Y_trgt).squeeze(0)


class ELBOLossLNPF(BaseLossNPF):
    """-(E_z sum_t log p(y_t|z) - KL(q(z|cntxt,trgt) || q(z|cntxt))) (npf/losses.py:126-150)."""

    def get_loss(self, p_yCc, _, q_zCc, q_zCct, Y_trgt):
        e_log_p = sum_log_prob(p_yCc, Y_trgt).mean(0)
        return -(e_log_p - _kl_sum(q_zCct, q_zCc))


class NLLLossLNPF(BaseLossNPF):
    """-(logsumexp_z sum_t log p(y_t|z) - log n_z), importance weighted when q_zCct is given (losses.py:153-203)."""

    def get_loss(self, p_yCc, z_samples, q_zCc, q_zCct, Y_trgt):
        n_z = p_yCc.batch_shape[0]
        log_w = sum_log_prob(p_yCc, Y_trgt)
        if q_zCct is not None:
            log_w = log_w + sum_from_nth_dim(q_zCc.log_prob(z_samples), 2) \
                - sum_from_nth_dim(q_zCct.log_prob(z_samples), 2)
        return -(torch.logsumexp(log_w, 0) - math.log(n_z))


class SUMOLossLNPF(BaseLossNPF):
    """Out of scope for the B200 hot path (commented out of the reference's book; python-loop logcumsumexp)."""

    def __init__(self, *args, **kwargs):
        raise NotImplementedError("SUMOLossLNPF is not part of the 1-D hot path; use NLLLossLNPF or ELBOLossLNPF")

    def get_loss(self, *args):  # pragma: no cover
        raise NotImplementedError
