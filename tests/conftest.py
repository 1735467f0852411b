import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (sm_100a); run with -m gpu on the B200 box")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(tag):
    import torch

    return torch.load(os.path.join(GOLDEN, f"{tag}.pt"), map_location="cpu", weights_only=False)


GOLDEN_TAGS = [
    "cnp_r32", "attncnp_r32", "attncnp_multihead_r32", "attncnp_scaledot_r32", "convcnp_r32",
    "convcnp_unet_r32", "convcnp_unshared_r32", "lnp_r32", "lnp_both_r32", "attnlnp_r32", "attnlnp_nll_r32",
    "attncnp_pretrained", "convcnp_pretrained", "convlnp_nll_r32", "convlnp_both_r32",
]
