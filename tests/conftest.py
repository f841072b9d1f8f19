import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) device; run with -m gpu on the GPU box")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(tag):
    return torch.load(os.path.join(GOLDEN, tag + ".pt"))


def seeded_model(name, **kw):
    """torch.manual_seed(0) + create_model reproduces the reference's default-initialised weights (see
    test_boundary.py::test_default_init_matches_reference)."""
    import unet_zoo_b200
    torch.manual_seed(0)
    return unet_zoo_b200.create_model(name, **kw)


def rel_err(a, b):
    """max|a-b| / max|b| - the logits criterion of BASELINE.json's north_star."""
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))
