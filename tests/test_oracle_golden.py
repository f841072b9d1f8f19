"""The oracle (CPU restatement, oracle/reference_port.py) against fixtures produced by the reference modules
themselves (oracle/make_golden.py).  CPU only."""
import copy

import pytest
import torch

from conftest import load_golden, rel_err, seeded_model
from oracle import reference_port as O

CASES = [("unet", {}), ("attention_unet", {}), ("nested_unet", {}), ("resunet", {}),
         ("nested_unet_ds", {"deep_supervision": True}), ("transatt_unet", {"attn_dropout": 0.0}),
         ("u2net", {}), ("u2netp", {}),      # seven outputs [main, side1..6]; the default-init fixture stores the first two
         ("multiresunet", {})]               # no BatchNorm affine parameters: the "conditioned" state is the default one


@pytest.mark.parametrize("tag,kw", CASES)
def test_oracle_matches_reference_outputs(tag, kw):
    torch.set_num_threads(8)
    name = tag.replace("_ds", "")
    gold = load_golden(tag)
    sd = copy.deepcopy(seeded_model(name, **kw).state_dict())
    n, c, h, w = gold["recipe"]["batch"]
    x, mask = O.synthetic_batch(n, c, h, w, gold["recipe"]["seed_x"], gold["recipe"]["seed_m"])
    out, loss, grads = O.train_step(name, sd, x, mask, **kw)
    outs = out if isinstance(out, list) else [out]
    assert len(outs) == (7 if name.startswith("u2net") else len(gold["logits_train"]))
    for o, g in zip(outs, gold["logits_train"]):
        assert o.shape == g.shape
        assert rel_err(o.detach(), g) < 1e-5
    assert abs(float(loss) - gold["loss"]) < 1e-5
    for k, gn in gold["grad_norm"].items():
        assert abs(float(grads[k].double().norm()) - gn) <= 2e-4 * gn + 1e-7, k
    for k, g in gold["grad_small"].items():
        assert float((grads[k] - g).abs().max()) <= 2e-4 * float(g.abs().max()) + 1e-6, k
    for k, v in gold["bn_after"].items():
        assert torch.allclose(sd[k].double(), v.double(), rtol=1e-5, atol=1e-6), k
    with torch.no_grad():
        oe = O.FORWARDS[name](sd, x, False, **kw)
    oes = oe if isinstance(oe, list) else [oe]
    for o, g in zip(oes, gold["logits_eval"]):
        assert rel_err(o, g) < 1e-5


@pytest.mark.parametrize("tag,kw", CASES)
def test_oracle_matches_reference_on_conditioned_state(tag, kw):
    """Same pin on the well-conditioned state the bf16 parity tests use (randomised BatchNorm affine parameters,
    Gaussian projection loss): train logits, BCE, gradients, running statistics, eval logits of the unmodified reference."""
    torch.set_num_threads(8)
    name = tag.replace("_ds", "")
    gold = load_golden(tag)
    cond = gold["cond"]
    sd = O.randomize_bn_affine(copy.deepcopy(seeded_model(name, **kw).state_dict()))
    n, c, h, w = gold["recipe"]["batch"]
    x, mask = O.synthetic_batch(n, c, h, w, gold["recipe"]["seed_x"], gold["recipe"]["seed_m"])
    out, _, grads = O.train_step(name, sd, x, mask, loss_fn=O.gaussian_projection_loss, **kw)
    outs = out if isinstance(out, list) else [out]
    for o, g in zip(outs, cond["logits_train"]):
        assert rel_err(o.detach(), g) < 1e-5
    assert abs(float(sum(O.bce_with_logits(o.detach(), mask) for o in outs)) - cond["bce"]) < 1e-5
    for k, gn in cond["grad_norm"].items():
        # + 1e-7: conv biases in front of a train-mode BatchNorm have an identically zero gradient (fp32 noise ~1e-8)
        assert abs(float(grads[k].double().norm()) - gn) <= 5e-4 * gn + 1e-7, k
    for k, v in cond["bn_after"].items():
        assert torch.allclose(sd[k].double(), v.double(), rtol=1e-5, atol=1e-6), k
    with torch.no_grad():
        oe = O.FORWARDS[name](sd, x, False, **kw)
    for o, g in zip(oe if isinstance(oe, list) else [oe], cond["logits_eval"]):
        assert rel_err(o, g) < 1e-5


def test_transatt_unet_attention_dropout_path():
    """The reference AS IS (nn.Dropout(0.1) on the channel-attention matrix, transatt_unet.py:88,103) under the fixture's
    RNG seed: the oracle's dropout restatement draws the same mask on CPU."""
    gold = load_golden("transatt_unet")
    sd = copy.deepcopy(seeded_model("transatt_unet").state_dict())
    n, c, h, w = gold["recipe"]["batch"]
    x, _ = O.synthetic_batch(n, c, h, w, gold["recipe"]["seed_x"], gold["recipe"]["seed_m"])
    torch.manual_seed(gold["dropout_seed"])
    with torch.no_grad():
        out = O.transatt_unet_forward(sd, x, True)               # default attn_dropout = 0.1
    assert rel_err(out, gold["logits_train_dropout"]) < 1e-5
    assert rel_err(gold["logits_train"][0], gold["logits_train_dropout"]) > 1e-4     # the mask does matter


def test_unet_transformer_oracle_groundwork():
    """SURVEY 8(f) rank-4 model `unet_transformer`, oracle side: the oracle's own default-init recipe
    reproduces the reference's weights (fingerprint of all 209 state_dict entries) and its restatement of
    U_Transformer.forward (self-attention, three cross-attention decoders) reproduces the reference's train / eval
    outputs, loss, gradients and running statistics."""
    torch.set_num_threads(8)
    gold = load_golden("unet_transformer")
    sd = O.unet_transformer_state_dict(gold["recipe"]["seed"])
    assert list(sd) == list(gold["fingerprint"]) and len(sd) == 209
    for k, v in sd.items():
        fp = gold["fingerprint"][k]
        assert float(v.double().sum()) == fp[0] and float(v.double().abs().sum()) == fp[1], k
    n, c, h, w = gold["recipe"]["batch"]
    x, mask = O.synthetic_batch(n, c, h, w, gold["recipe"]["seed_x"], gold["recipe"]["seed_m"])
    out, loss, grads = O.train_step("unet_transformer", sd, x, mask)
    assert rel_err(out.detach(), gold["logits_train"][0]) < 1e-5
    assert abs(float(loss) - gold["loss"]) < 1e-5
    for k, gn in gold["grad_norm"].items():
        assert abs(float(grads[k].double().norm()) - gn) <= 2e-4 * gn + 1e-7, k
    for k, g in gold["grad_small"].items():
        assert float((grads[k] - g).abs().max()) <= 2e-4 * float(g.abs().max()) + 1e-6, k
    for k, v in gold["bn_after"].items():
        assert torch.allclose(sd[k].double(), v.double(), rtol=1e-5, atol=1e-6), k
    with torch.no_grad():
        oe = O.unet_transformer_forward(sd, x, False)
    assert rel_err(oe, gold["logits_eval"][0]) < 1e-5
    # the product's parameter containers reproduce the same default initialisation (tests/test_boundary.py checks the
    # fingerprint; here: the oracle's own recipe and the product agree entry by entry)
    psd = seeded_model("unet_transformer").state_dict()
    fresh = O.unet_transformer_state_dict(gold["recipe"]["seed"])
    assert list(psd) == list(fresh)
    for k, v in psd.items():
        assert torch.equal(v, fresh[k]), k


def test_threshold_is_bit_exact_with_sign_test():
    """sigmoid(z) > 0.5 (metrics.py:11-12) agrees with z > 0 away from the fp32 saturation point; the parity tests use
    the reference formulation itself."""
    z = torch.linspace(-8, 8, 100001)
    assert torch.equal(O.threshold_mask(z), z > 0)


def test_dice_edge_cases():
    z = torch.full((1, 1, 4, 4), -5.0)
    assert float(O.dice_coefficient(z, torch.zeros(1, 1, 4, 4))) == 1.0     # empty union -> 1.0 (metrics.py:18-19)
    t = torch.ones(1, 1, 4, 4)
    assert abs(float(O.dice_coefficient(-z, t)) - 1.0) < 1e-6
    assert float(O.dice_coefficient(z, t)) < 1e-6


def test_readme_example_shape():
    """README.md:74-90: unet on randn(1,3,256,256) -> (1,1,256,256)."""
    assert load_golden("unet")["readme_shape"] == [1, 1, 256, 256]
    sd = seeded_model("unet").state_dict()
    with torch.no_grad():
        assert list(O.unet_forward(sd, torch.randn(1, 3, 64, 64), False).shape) == [1, 1, 64, 64]
