"""fp32 parity mode on the GPU (`precision="fp32"`, csrc/f32.cu through the C ABI) against the golden fixtures that
the UNMODIFIED reference produced (oracle/make_golden.py) and against the CPU oracle.

BASELINE.json's bar for this mode: logits within 1e-4 relative (max|a-b| / max|b|), >= 99.9 % agreement of the
thresholded masks; configs[0] is `unet`, batch 1, 256x256 fp32 forward (the reference's README example)."""
import copy

import pytest
import torch

from conftest import load_golden, rel_err, seeded_model
from oracle import reference_port as O

pytestmark = pytest.mark.gpu
LOGIT_TOL_FP32 = 1e-4
CASES = [("unet", {}), ("attention_unet", {}), ("nested_unet", {}), ("nested_unet_ds", {"deep_supervision": True}),
         ("resunet", {}), ("transatt_unet", {"attn_dropout": 0.0})]


def _as_list(o):
    return list(o) if isinstance(o, (list, tuple)) else [o]


@pytest.mark.parametrize("tag,kw", CASES)
def test_fp32_train_and_eval_forward_match_the_reference_goldens(tag, kw):
    name = tag.replace("_ds", "")
    gold = load_golden(tag)
    model = seeded_model(name, precision="fp32", **kw).cuda()
    x, _ = O.synthetic_batch(*gold["recipe"]["batch"])
    model.train()
    with torch.no_grad():                                   # batch statistics + running-stat update, like the fixture
        outs = [o.cpu() for o in _as_list(model(x.cuda()))]
    assert len(outs) == len(gold["logits_train"])
    for o, g in zip(outs, gold["logits_train"]):
        assert o.shape == g.shape and o.dtype == torch.float32
        assert rel_err(o, g) < LOGIT_TOL_FP32, rel_err(o, g)
    got = model.state_dict()
    for k, v in gold["bn_after"].items():
        if "running_" in k:
            assert rel_err(got[k].cpu(), v) < 1e-5, k
        else:
            assert int(got[k]) == int(v), k
    model.eval()
    with torch.no_grad():
        outs = [o.cpu() for o in _as_list(model(x.cuda()))]
    for o, g in zip(outs, gold["logits_eval"]):
        assert rel_err(o, g) < LOGIT_TOL_FP32, rel_err(o, g)
        agree = float((O.threshold_mask(o) == O.threshold_mask(g)).float().mean())
        assert agree >= 0.999, agree


def test_fp32_readme_example_config0():
    """BASELINE configs[0]: unet in_channels=3 num_classes=1, batch 1, 256x256 fp32 forward (README.md:83-90)."""
    model = seeded_model("unet", precision="fp32")
    sd = copy.deepcopy(model.state_dict())
    g = torch.Generator().manual_seed(7)
    x = torch.randn(1, 3, 256, 256, generator=g)
    ref = O.FORWARDS["unet"](sd, x, False)
    model.cuda().eval()
    with torch.no_grad():
        out = model(x.cuda()).cpu()
    assert tuple(out.shape) == (1, 1, 256, 256)
    assert rel_err(out, ref) < LOGIT_TOL_FP32, rel_err(out, ref)
    assert float((O.threshold_mask(out) == O.threshold_mask(ref)).float().mean()) >= 0.999


@pytest.mark.parametrize("name,kw,shape", [("unet", {"in_channels": 1, "num_classes": 3}, (3, 1, 48, 80)),
                                           ("resunet", {"in_channels": 4, "num_classes": 2}, (1, 4, 32, 64)),
                                           ("attention_unet", {"num_classes": 5}, (2, 3, 32, 32)),
                                           ("nested_unet", {"in_channels": 9, "num_classes": 2}, (1, 9, 64, 32))])
def test_fp32_other_channel_counts_and_ragged_shapes(name, kw, shape):
    """Channel / class counts other than 3 -> 1 (multi-class logits leave through the NHWC -> NCHW kernel),
    non-square inputs, batch statistics in train mode."""
    model = seeded_model(name, precision="fp32", **kw)
    sd = copy.deepcopy(model.state_dict())
    x = torch.randn(*shape, generator=torch.Generator().manual_seed(11))
    model.cuda()
    for training in (True, False):
        ref = _as_list(O.FORWARDS[name](copy.deepcopy(sd), x, training))
        model.load_state_dict(sd)
        model.train(training)
        with torch.no_grad():
            outs = [o.cpu() for o in _as_list(model(x.cuda()))]
        for o, r in zip(outs, ref):
            assert o.shape == r.shape
            assert rel_err(o, r) < LOGIT_TOL_FP32, (training, rel_err(o, r))


def test_fp32_mode_is_forward_only_and_switchable():
    model = seeded_model("unet", precision="fp32").cuda().train()
    x, mask = O.synthetic_batch(2, 3, 32, 32)
    with pytest.raises(RuntimeError, match="forward-only"):
        model(x.cuda())
    with torch.no_grad():
        a = model(x.cuda())
    assert not a.requires_grad
    model.set_precision("bf16")                              # same parameters, tensor-core arithmetic, trains
    out = model(x.cuda())
    torch.nn.functional.binary_cross_entropy_with_logits(out, mask.cuda()).backward()
    assert model.out.conv.weight.grad is not None
    with pytest.raises(ValueError):
        model.set_precision("fp16")
