import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _fp32_mlp_by_default():
    """Parity tests at the 1e-4 bar run the fp32 CUDA-core MLP; tensor-core tests opt in explicitly."""
    from nerf_kbs_b200 import ops
    ops.set_mlp_precision("fp32")
    yield
    ops.set_mlp_precision("tc")
