"""The neural-process family on B200: CNP / LNP / AttnCNP / AttnLNP / ConvCNP.

Drop-in for the reference's model classes (npf/neuralproc/base.py, np.py, attnnp.py, convnp.py): same
constructors, plug-in factories, attribute names, ``state_dict`` layout and
``forward(X_cntxt, Y_cntxt, X_trgt, Y_trgt=None) -> (p_yCc, z_samples, q_zCc, q_zCct)`` contract.
All arithmetic on the path runs in the npfk CUDA kernels; CPU tensors raise (no fallback).
"""
import abc
import logging
from functools import partial

import torch
import torch.nn as nn

from .. import kernels as K
from ..architectures.attenders import get_attender
from ..architectures.blocks import MLP, discard_ith_arg, merge_flat_input
from ..architectures.convnets import CNN, ResConvBlock
from ..architectures.setconv import SetConv
from ..utils.helpers import MultivariateNormalDiag
from ..utils.initialization import weights_init

logger = logging.getLogger(__name__)

__all__ = ["NeuralProcessFamily", "LatentNeuralProcessFamily", "CNP", "LNP", "AttnCNP", "AttnLNP", "ConvCNP",
           "ConvLNP", "set_lazy_validation"]

_cfg = {"lazy_validation": False}


def set_lazy_validation(flag: bool) -> None:
    """False (default): like the reference, a training forward synchronises to raise ValueError for X outside
    [-1,1].  True: the range flag is still computed on device every training forward, but only read back at the
    *next* forward (no per-step pipeline stall); the error then surfaces one call late."""
    _cfg["lazy_validation"] = bool(flag)


class _DefaultScaleTransform:
    """0.01 + 0.99 * softplus(s) (base.py:116).  A class, so decode() can recognise the default and fuse it."""

    def __call__(self, s):      # direct calls only; decode() fuses this into npfk_loc_scale / npfk_gauss_nll
        return 0.01 + 0.99 * torch.nn.functional.softplus(s)


class _DefaultLatentScaleTransform:
    """0.1 + 0.9 * sigmoid(s) (base.py:432)."""

    def __call__(self, s):      # direct calls only; infer_latent_dist() fuses this into npfk_latent_stats
        return 0.1 + 0.9 * torch.sigmoid(s)


class NeuralProcessFamily(nn.Module, abc.ABC):
    _valid_paths = ["deterministic", "latent", "both"]

    def __init__(self, x_dim, y_dim, encoded_path, r_dim=128, x_transf_dim=-1, is_heteroskedastic=True,
                 XEncoder=None, Decoder=None, PredictiveDistribution=MultivariateNormalDiag,
                 p_y_loc_transformer=nn.Identity(), p_y_scale_transformer=None):
        super().__init__()
        self.x_dim, self.y_dim, self.r_dim = x_dim, y_dim, r_dim
        self.is_heteroskedastic = is_heteroskedastic
        self.x_transf_dim = x_dim if x_transf_dim is None else (r_dim if x_transf_dim == -1 else x_transf_dim)
        self.encoded_path = encoded_path.lower()
        if self.encoded_path not in self._valid_paths:
            raise ValueError(f"Unknown encoded_path={self.encoded_path}.")
        XEncoder = self.dflt_Modules["XEncoder"] if XEncoder is None else XEncoder
        Decoder = self.dflt_Modules["Decoder"] if Decoder is None else Decoder
        self.x_encoder = XEncoder(self.x_dim, self.x_transf_dim)
        self.decoder = Decoder(self.x_transf_dim, self.r_dim, self.y_dim * 2)   # loc and scale
        self.PredictiveDistribution = PredictiveDistribution
        self.p_y_loc_transformer = p_y_loc_transformer
        self.p_y_scale_transformer = _DefaultScaleTransform() if p_y_scale_transformer is None else p_y_scale_transformer
        self._range_flag = None
        self._pending_flag = None
        self.reset_parameters()

    def reset_parameters(self):
        weights_init(self)

    @property
    def dflt_Modules(self):
        sub_decoder = partial(MLP, n_hidden_layers=4, hidden_size=self.r_dim)
        return {
            "XEncoder": partial(MLP, n_hidden_layers=1, hidden_size=self.r_dim),
            "SubDecoder": sub_decoder,
            "Decoder": merge_flat_input(sub_decoder, is_sum_merge=True),
        }

    # ------------------------------------------------------------------ forward template (base.py:177-239)
    def forward(self, X_cntxt, Y_cntxt, X_trgt, Y_trgt=None):
        K.require_cuda(X_cntxt, Y_cntxt, X_trgt, Y_trgt)
        self._validate_inputs(X_cntxt, Y_cntxt, X_trgt, Y_trgt)
        X_cntxt = self.x_encoder(X_cntxt)
        X_trgt = self.x_encoder(X_trgt)
        R = self.encode_globally(X_cntxt, Y_cntxt)
        if self.encoded_path in ("latent", "both"):
            z_samples, q_zCc, q_zCct = self.latent_path(X_cntxt, R, X_trgt, Y_trgt)
        else:
            z_samples, q_zCc, q_zCct = None, None, None
        if self.encoded_path == "latent":
            R = None
        R_trgt = self.trgt_dependent_representation(X_cntxt, z_samples, R, X_trgt)
        p_yCc = self.decode(X_trgt, R_trgt, Y_trgt)
        return p_yCc, z_samples, q_zCc, q_zCct

    def _validate_inputs(self, X_cntxt, Y_cntxt, X_trgt, Y_trgt):
        """ValueError for features outside [-1,1], training mode only (base.py:241-247)."""
        capturing = X_trgt.is_cuda and torch.cuda.is_current_stream_capturing()
        if self._pending_flag is not None and not capturing:     # lazy mode: result of the previous forward
            flag, self._pending_flag = self._pending_flag, None
            if int(flag.item()):
                raise ValueError("Features during training should be in [-1,1] (detected on the previous forward).")
        if not self.training:
            return
        flag = torch.zeros(1, dtype=torch.int32, device=X_trgt.device)
        if X_cntxt.numel():
            K.range_flag(X_cntxt, flag)
        K.range_flag(X_trgt, flag)
        if _cfg["lazy_validation"] or capturing:     # inside a CUDA-graph capture nothing may synchronise:
            self._pending_flag = flag                # the owner of the graph reads the flag after a replay
        elif int(flag.item()):
            raise ValueError(
                f"Features during training should be in [-1,1]. {X_cntxt.min()} <= X_cntxt <= {X_cntxt.max()} ; "
                f"{X_trgt.min()} <= X_trgt <= {X_trgt.max()}."
            )

    @abc.abstractmethod
    def encode_globally(self, X_cntxt, Y_cntxt):
        """[B,Nc,x_transf], [B,Nc,y] -> R [B, n_rep, r_dim]"""

    @abc.abstractmethod
    def trgt_dependent_representation(self, X_cntxt, z_samples, R, X_trgt):
        """-> R_trgt [n_z, B, Nt, r_dim]"""

    def latent_path(self, X_cntxt, R, X_trgt, Y_trgt):
        raise NotImplementedError(
            f"`latent_path` not implemented. Cannot use encoded_path={self.encoded_path} in such case."
        )

    # ------------------------------------------------------------------ decode (base.py:327-367)
    def decode(self, X_trgt, R_trgt, Y_trgt=None):
        """``Y_trgt`` (optional; the reference's decode has no such argument): when the targets are known, the
        Gaussian log-likelihood summed over each task's targets is evaluated in the decoder kernel's epilogue and
        attached to the returned distribution, where ``sum_log_prob`` picks it up (north_star (4))."""
        fused = isinstance(self.p_y_loc_transformer, nn.Identity) and isinstance(
            self.p_y_scale_transformer, _DefaultScaleTransform)
        lp = None
        if (fused and self.is_heteroskedastic and self.y_dim == 1 and Y_trgt is not None and R_trgt.dim() == 4
                and getattr(self.decoder, "accepts_nll", False)):
            suff, lp = self.decoder(X_trgt, R_trgt, nll=(Y_trgt, R_trgt.shape[2]))
        else:
            suff = self.decoder(X_trgt, R_trgt)                  # [n_z, B, Nt, 2*y_dim]
        if fused:
            loc, scale = K.loc_scale(suff)
        else:
            loc, scale = suff.split(self.y_dim, dim=-1)
            loc = self.p_y_loc_transformer(loc)
            scale = self.p_y_scale_transformer(scale)
        if not self.is_heteroskedastic:
            n_z, B, *n_trgt, y_dim = scale.shape
            pooled = K.set_mean(scale.reshape(n_z * B, -1, y_dim))
            scale = pooled.view(n_z, B, *([1] * len(n_trgt)), y_dim).expand(n_z, B, *n_trgt, y_dim)
        p_yCc = self.PredictiveDistribution(loc, scale)
        if fused and self.is_heteroskedastic:
            # lets the npf losses run Normal.log_prob + the target sum as one kernel straight from the logits
            p_yCc._npfk_suffstat = suff
            if lp is not None:
                p_yCc._npfk_sum_lp = (lp.view(R_trgt.shape[0], R_trgt.shape[1]), Y_trgt.data_ptr(), Y_trgt._version,
                                      tuple(Y_trgt.shape))
        return p_yCc

    def set_extrapolation(self, min_max):
        pass


class LatentNeuralProcessFamily(NeuralProcessFamily):
    _valid_paths = ["latent", "both"]

    def __init__(self, *args, is_q_zCct=False, n_z_samples_train=32, n_z_samples_test=32, LatentEncoder=None,
                 LatentDistribution=MultivariateNormalDiag, q_z_loc_transformer=nn.Identity(),
                 q_z_scale_transformer=None, z_dim=None, **kwargs):
        super().__init__(*args, **kwargs)
        self.is_q_zCct = is_q_zCct
        self.n_z_samples_train = n_z_samples_train
        self.n_z_samples_test = n_z_samples_test
        self.z_dim = self.r_dim if z_dim is None else z_dim
        LatentEncoder = self.dflt_Modules["LatentEncoder"] if LatentEncoder is None else LatentEncoder
        self.latent_encoder = LatentEncoder(self.r_dim, self.z_dim * 2)
        if self.encoded_path == "both":
            self.r_z_merger = nn.Linear(self.r_dim + self.z_dim, self.r_dim)
        self.LatentDistribution = LatentDistribution
        self.q_z_loc_transformer = q_z_loc_transformer
        self.q_z_scale_transformer = (_DefaultLatentScaleTransform() if q_z_scale_transformer is None
                                      else q_z_scale_transformer)
        if self.z_dim != self.r_dim and self.encoded_path == "latent":
            self.reshaper_z = nn.Linear(self.z_dim, self.r_dim)
        self.eps_generator = None     # optional torch.Generator (device RNG) for reproducible latent draws
        self.reset_parameters()

    @property
    def dflt_Modules(self):
        mods = NeuralProcessFamily.dflt_Modules.__get__(self)
        mods["LatentEncoder"] = partial(MLP, n_hidden_layers=1, hidden_size=self.r_dim)
        return mods

    def forward(self, *args, **kwargs):
        n = self.n_z_samples_train if self.training else self.n_z_samples_test
        self.n_z_samples = n.rvs() if hasattr(n, "rvs") else n       # scipy rv = random number of samples
        return super().forward(*args, **kwargs)

    def latent_path(self, X_cntxt, R, X_trgt, Y_trgt):
        q_zCc = self.infer_latent_dist(X_cntxt, R)
        if self.is_q_zCct and Y_trgt is not None:
            q_zCct = self.infer_latent_dist(X_trgt, self.encode_globally(X_trgt, Y_trgt))
            sampling = q_zCct
        else:
            q_zCct, sampling = None, q_zCc
        z_samples = self._rsample(sampling, int(self.n_z_samples))
        return z_samples, q_zCc, q_zCct

    def _rsample(self, dist, n):
        base = getattr(dist, "base_dist", None)
        if base is None or not hasattr(base, "loc") or not hasattr(base, "scale"):
            return dist.rsample([n])                    # user-supplied distribution: its own sampler
        eps = self.draw_eps(n, base.loc)
        return K.rsample(base.loc, base.scale, eps)

    def draw_eps(self, n, like):
        """Standard-normal noise [n, *like.shape] on the device (override / patch to inject a fixed epsilon)."""
        return torch.randn(n, *like.shape, device=like.device, dtype=torch.float32, generator=self.eps_generator)

    def infer_latent_dist(self, X, R):
        suff = self.latent_encoder(self.rep_to_lat_input(R))
        fused = isinstance(self.q_z_loc_transformer, nn.Identity) and isinstance(
            self.q_z_scale_transformer, _DefaultLatentScaleTransform)
        if fused:
            loc, scale = K.latent_stats(suff)
        else:
            loc, scale = suff.split(self.z_dim, dim=-1)
            loc, scale = self.q_z_loc_transformer(loc), self.q_z_scale_transformer(scale)
        return self.LatentDistribution(loc, scale)

    def rep_to_lat_input(self, R):
        return R

    def merge_r_z(self, R, z_samples):
        """relu(r_z_merger(cat(R, z))) with the weight split as [W_r | W_z] (base.py:554-575):
        W_r R is computed once per target, W_z z once per (sample, task); nothing of size n_z*B*Nt*2r is built."""
        W, b = self.r_z_merger.weight, self.r_z_merger.bias
        W_r, W_z = W[:, : self.r_dim], W[:, self.r_dim:]
        if R.dim() == 3 and z_samples.dim() == 4 and (z_samples.shape[2] == 1 or z_samples.stride(2) == 0) \
                and R.shape[1] != 1:
            zp = K.linear(z_samples[:, :, 0, :], W_z, b)                 # [n_z, B, r]
            return K.bcast_add_relu(K.linear(R, W_r), zp)                 # [n_z, B, Nt, r]
        if R.shape != z_samples.shape:
            R = R.unsqueeze(0).expand(*z_samples.shape[:-1], self.r_dim)
        return K.linear(z_samples, W_z, b, res=K.linear(R, W_r), relu=True)


# =============================================================================== members
def _xy_encoder_default(self):
    sub = partial(MLP, n_hidden_layers=2, is_force_hid_smaller=True, hidden_size=self.r_dim)
    return merge_flat_input(sub, is_sum_merge=True)


class CNP(NeuralProcessFamily):
    """Conditional neural process: mean-pooled context representation (np.py:19-110)."""

    _valid_paths = ["deterministic"]

    def __init__(self, x_dim, y_dim, XYEncoder=None, **kwargs):
        kwargs["encoded_path"] = kwargs.get("encoded_path", "deterministic")
        super().__init__(x_dim, y_dim, **kwargs)
        XYEncoder = self.dflt_Modules["XYEncoder"] if XYEncoder is None else XYEncoder
        self.xy_encoder = XYEncoder(self.x_transf_dim, self.y_dim, self.r_dim)
        self.reset_parameters()

    @property
    def dflt_Modules(self):
        mods = NeuralProcessFamily.dflt_Modules.__get__(self)
        mods["XYEncoder"] = _xy_encoder_default(self)
        return mods

    def encode_globally(self, X_cntxt, Y_cntxt):
        B, n_cntxt, _ = X_cntxt.shape
        if n_cntxt == 0:      # no context: zero representation
            return torch.zeros(B, 1, self.r_dim, device=X_cntxt.device)
        return K.set_mean(self.xy_encoder(X_cntxt, Y_cntxt))

    def trgt_dependent_representation(self, _, __, R, X_trgt):
        B, n_trgt, _ = X_trgt.shape
        return R.expand(B, n_trgt, self.r_dim).unsqueeze(0)


class LNP(LatentNeuralProcessFamily, CNP):
    """(Latent) neural process (np.py:113-163)."""

    def __init__(self, x_dim, y_dim, encoded_path="latent", **kwargs):
        super().__init__(x_dim, y_dim, encoded_path=encoded_path, **kwargs)

    @property
    def dflt_Modules(self):
        mods = CNP.dflt_Modules.__get__(self)
        mods.update(LatentNeuralProcessFamily.dflt_Modules.__get__(self))
        return mods

    def trgt_dependent_representation(self, _, z_samples, R, X_trgt):
        B, n_trgt, _ = X_trgt.shape
        n_z = z_samples.size(0)
        if self.encoded_path == "both":
            R_trgt = self.merge_r_z(R, z_samples)
        else:
            R_trgt = z_samples
            if self.z_dim != self.r_dim:
                R_trgt = K.linear(R_trgt, self.reshaper_z.weight, self.reshaper_z.bias)
        return R_trgt.expand(n_z, B, n_trgt, self.r_dim)


class AttnCNP(NeuralProcessFamily):
    """Attentive CNP: per-context representations + target->context cross attention (attnnp.py:27-131)."""

    _valid_paths = ["deterministic"]

    def __init__(self, x_dim, y_dim, XYEncoder=None, attention="scaledot", attention_kwargs={},
                 self_attention_kwargs={}, is_self_attn=False, **kwargs):
        kwargs["encoded_path"] = kwargs.get("encoded_path", "deterministic")
        super().__init__(x_dim, y_dim, **kwargs)
        self.is_self_attn = is_self_attn
        if is_self_attn:
            raise NotImplementedError("is_self_attn=True (image models) is outside the 1-D hot path")
        XYEncoder = self.dflt_Modules["XYEncoder"] if XYEncoder is None else XYEncoder
        self.xy_encoder = XYEncoder(self.x_transf_dim, self.y_dim, self.r_dim)
        self.attender = get_attender(attention, self.x_transf_dim, self.r_dim, self.r_dim, **attention_kwargs)
        self.reset_parameters()

    dflt_Modules = CNP.dflt_Modules

    def encode_globally(self, X_cntxt, Y_cntxt):
        B, n_cntxt, _ = X_cntxt.shape
        if n_cntxt == 0:
            return torch.zeros(B, 0, self.r_dim, device=X_cntxt.device)
        return self.xy_encoder(X_cntxt, Y_cntxt)

    def trgt_dependent_representation(self, X_cntxt, _, R, X_trgt):
        B, n_cntxt, _ = X_cntxt.shape
        if n_cntxt == 0:
            R_trgt = torch.zeros(B, X_trgt.size(1), self.r_dim, device=R.device)
        else:
            R_trgt = self.attender(X_cntxt, X_trgt, R)       # keys, queries, values
        return R_trgt.unsqueeze(0)


class AttnLNP(LatentNeuralProcessFamily, AttnCNP):
    """Attentive latent NP: deterministic attention path merged with a global latent (attnnp.py:134-202)."""

    _valid_paths = ["both"]

    def __init__(self, x_dim, y_dim, **kwargs):
        super().__init__(x_dim, y_dim, encoded_path="both", **kwargs)

    @property
    def dflt_Modules(self):
        mods = AttnCNP.dflt_Modules.__get__(self)
        mods.update(LatentNeuralProcessFamily.dflt_Modules.__get__(self))
        return mods

    def rep_to_lat_input(self, R):
        B, n_cntxt, _ = R.shape
        if n_cntxt == 0:
            return torch.zeros(B, 1, self.r_dim, device=R.device)
        return K.set_mean(R)

    def trgt_dependent_representation(self, X_cntxt, z_samples, R, X_trgt):
        B, n_trgt, _ = X_trgt.shape
        n_z = z_samples.size(0)
        z_samples = z_samples.expand(n_z, B, n_trgt, self.z_dim)
        R_det = AttnCNP.trgt_dependent_representation(self, X_cntxt, None, R, X_trgt).squeeze(0)
        return self.merge_r_z(R_det, z_samples)


class ConvCNP(NeuralProcessFamily):
    """Convolutional CNP: SetConv to a regular grid, 1-D CNN, SetConv to the targets (convnp.py:26-181)."""

    _valid_paths = ["deterministic"]

    def __init__(self, x_dim, y_dim, density_induced=128, Interpolator=SetConv,
                 CNN=partial(CNN, ConvBlock=ResConvBlock, Conv=nn.Conv1d, n_blocks=3, Normalization=nn.Identity,
                             is_chan_last=True, kernel_size=11),
                 **kwargs):
        if "Decoder" in kwargs and kwargs["Decoder"] != nn.Identity:
            logger.warning("`Decoder` was given to `ConvCNP`. To be translation equivariant you should disregard the "
                           "first argument for example using `discard_ith_arg(Decoder, i=0)`, which is done by "
                           "default when you DO NOT provide the Decoder.")
        kwargs["encoded_path"] = kwargs.get("encoded_path", "deterministic")
        super().__init__(x_dim, y_dim, x_transf_dim=None, XEncoder=nn.Identity, **kwargs)
        self.density_induced = density_induced
        # half a unit of margin on each side of [-1,1] against boundary effects; a plain attribute, not a buffer
        self.X_induced = torch.linspace(-1.5, 1.5, int(self.density_induced * 3))
        self.CNN = CNN
        self.cntxt_to_induced = Interpolator(self.x_dim, self.y_dim, self.r_dim)
        self.induced_to_induced = CNN(self.r_dim)
        self.induced_to_trgt = Interpolator(self.x_dim, self.r_dim, self.r_dim)
        self.reset_parameters()

    @property
    def n_induced(self):
        return len(self.X_induced)

    @property
    def dflt_Modules(self):
        mods = NeuralProcessFamily.dflt_Modules.__get__(self)
        mods["Decoder"] = discard_ith_arg(mods["SubDecoder"], i=0)
        return mods

    def _get_X_induced(self, X):
        B = X.shape[0]
        self.X_induced = self.X_induced.to(X.device)          # moved once, then cached on the device
        return self.X_induced.view(1, -1, 1).expand(B, self.n_induced, self.x_dim)

    def encode_globally(self, X_cntxt, Y_cntxt):
        B, n_cntxt, _ = X_cntxt.shape
        X_induced = self._get_X_induced(X_cntxt)
        if n_cntxt == 0:
            R_induced = torch.zeros(B, self.n_induced, self.r_dim, device=X_cntxt.device)
        else:
            R_induced = self.cntxt_to_induced(X_cntxt, X_induced, Y_cntxt)
        return self.induced_to_induced(R_induced)

    def trgt_dependent_representation(self, X_cntxt, z_samples, R_induced, X_trgt):
        X_induced = self._get_X_induced(X_cntxt)
        return self.induced_to_trgt(X_induced, X_trgt, R_induced).unsqueeze(0)

    def set_extrapolation(self, min_max):
        lo, hi = min_max[0] - 0.5, min_max[1] + 0.5
        self.X_induced = torch.linspace(lo, hi, int(self.density_induced * (hi - lo)))


class ConvLNP(LatentNeuralProcessFamily, ConvCNP):
    """Convolutional latent NP (convnp.py:184-335): one latent per grid point, a second CNN after sampling
    ("mixing" the samples), optional global latent, SetConv to the targets on the n_z * B batch.

    Same constructor as the reference: ``CNNPostZ=None`` (defaults to ``CNN``), ``encoded_path`` in
    {"latent", "both"}, ``is_global``, plus the ConvCNP / latent-family keyword arguments.  Every stage runs in the
    npfk kernels already used by ConvCNP and the latent family (SetConv, conv blocks, latent statistics, rsample,
    hoisted SetConv interpolation, skinny Linear decoder)."""

    _valid_paths = ["latent", "both"]

    def __init__(self, x_dim, y_dim, CNNPostZ=None, encoded_path="latent", is_global=False, **kwargs):
        super().__init__(x_dim, y_dim, encoded_path=encoded_path, **kwargs)
        self.is_global = is_global
        if CNNPostZ is None:
            CNNPostZ = self.CNN
        self.induced_to_induced_post_sampling = CNNPostZ(self.r_dim)
        self.reset_parameters()

    @property
    def dflt_Modules(self):
        mods = ConvCNP.dflt_Modules.__get__(self)
        mods.update(LatentNeuralProcessFamily.dflt_Modules.__get__(self))
        # smaller decoder than ConvCNP: the representation already went through the post-sampling CNN
        mods["Decoder"] = discard_ith_arg(nn.Linear, i=0)
        return mods

    def rep_to_lat_input(self, R):
        if self.encoded_path == "latent":
            return R                                         # one latent per induced point: [B, n_induced, r]
        return K.set_mean(R.reshape(R.shape[0], -1, self.r_dim))    # "both": a global latent

    def trgt_dependent_representation(self, X_cntxt, z_samples, R_induced, X_trgt):
        B, n_trgt, _ = X_trgt.shape
        n_z = z_samples.shape[0]
        self.X_induced = self.X_induced.to(X_trgt.device)
        X_induced = self.X_induced.view(1, -1, 1).expand(n_z * B, self.n_induced, self.x_dim)   # shared grid (stride 0)
        X_trgt_rep = X_trgt.unsqueeze(0).expand(n_z, B, n_trgt, self.x_dim).reshape(n_z * B, n_trgt, self.x_dim)
        if self.encoded_path == "latent":
            z = z_samples.reshape(n_z * B, self.n_induced, self.z_dim)
            if self.z_dim != self.r_dim:
                z = K.linear(z, self.reshaper_z.weight, self.reshaper_z.bias)
            z = self.induced_to_induced_post_sampling(z)    # mixing after sampling: coherent samples
            if self.is_global:
                z = self.add_global_latent(z)
            R_trgt = self.induced_to_trgt(X_induced, X_trgt_rep, z)
        else:
            z = z_samples.expand(n_z, B, self.n_induced, self.z_dim)
            R = self.merge_r_z(R_induced, z).reshape(n_z * B, self.n_induced, self.r_dim)
            R = self.induced_to_induced_post_sampling(R)
            R_trgt = self.induced_to_trgt(X_induced, X_trgt_rep, R)
        return R_trgt.view(n_z, B, n_trgt, self.r_dim)

    def add_global_latent(self, z_samples):
        """second half of the channels := their mean over the grid (helpers.py:21-34 pool_and_replicate_middle)"""
        half = z_samples.shape[-1] // 2
        local, glob = z_samples.split(half, dim=-1)
        glob = glob.mean(dim=1, keepdim=True).expand(-1, z_samples.shape[1], -1)
        return torch.cat([local, glob], dim=-1)
