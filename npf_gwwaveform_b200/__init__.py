"""npf_gwwaveform_b200 -- B200-native (sm_100a) implementation of the npf neural-process hot path.

Drop-in for ``import npf``: same model classes, plug-in factories and losses, backed by the hand-written
CUDA kernels of ``libnpfk.so`` (C-ABI in include/npfk.h).  There is no CPU fallback.
"""
from . import kernels
from .architectures import *  # noqa: F401,F403
from .kernels import get_mode, launch_count, set_mode, set_sync_bn
from .losses import *  # noqa: F401,F403
from .neuralproc import *  # noqa: F401,F403

__version__ = "0.1.0"
