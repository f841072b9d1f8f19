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


def conditioned_model(name, **kw):
    """(model, state_dict copy) on the well-conditioned parity state: default init under seed 0 + randomised BatchNorm
    affine parameters (oracle.reference_port.randomize_bn_affine).  The golden fixtures hold the
    unmodified reference's outputs on this state under gold["cond"]."""
    import copy

    from oracle import reference_port as O
    model = seeded_model(name, **kw)
    sd = O.randomize_bn_affine(copy.deepcopy(model.state_dict()))
    model.load_state_dict(sd)
    return model, copy.deepcopy(sd)


def grad_errors(grads, ref_grads):
    """(rel. L2 error of all conv / convT weight gradients concatenated, sorted [(rel. L2, key)] per tensor)."""
    keys = [k for k, g in ref_grads.items() if g.dim() == 4]
    ours = torch.cat([grads[k].double().reshape(-1).cpu() for k in keys])
    ref = torch.cat([ref_grads[k].double().reshape(-1) for k in keys])
    per = sorted(((float((grads[k].double().cpu() - ref_grads[k].double()).norm() /
                         ref_grads[k].double().norm().clamp_min(1e-30)), k) for k in keys), reverse=True)
    return float((ours - ref).norm() / ref.norm()), per


def parity_report(tag, **kv):
    """Append measured parity numbers to $UZ_PARITY_REPORT (json lines) - how the fixed bars were calibrated."""
    import json
    path = os.environ.get("UZ_PARITY_REPORT_LINES")
    if path:
        with open(path, "a") as f:
            f.write(json.dumps({"tag": tag, **kv}) + "\n")
