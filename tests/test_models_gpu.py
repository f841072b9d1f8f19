"""Whole-model parity of `attention_unet`, `nested_unet` (with and without deep supervision) and `resunet` on the GPU:
the CUDA path through create_model / nn.Module against the CPU oracle and the golden fixtures generated from the
unmodified reference (oracle/make_golden.py).  Tolerances are BASELINE.json's bf16 bars: logits within 2e-2 relative
(max|a-b| / max|b|), eval-mode thresholded-mask agreement >= 99.9 %.

Train mode at random init is ill-conditioned for attention_unet / nested_unet (1-channel BatchNorm on the gate logit,
batch statistics over 32 values at the 4x4 bottleneck): torch's OWN bf16 autocast of the reference math deviates from
fp32 by 26 % / 5 % there (measured, tools/debug_models.py).  The train-mode bar is therefore max(2e-2, 1.3 x that
floor), computed in the test from the oracle; the eval-mode bar is the plain 2e-2 (measured 0.1-0.5 %)."""
import copy

import pytest
import torch

from conftest import load_golden, rel_err, seeded_model
from oracle import reference_port as O

pytestmark = pytest.mark.gpu
LOGIT_TOL_BF16 = 2e-2
CASES = [("attention_unet", {}), ("nested_unet", {}), ("nested_unet_ds", {"deep_supervision": True}), ("resunet", {}),
         ("transatt_unet", {"attn_dropout": 0.0})]     # SURVEY 8(f) rank 4; attention dropout off: CPU / CUDA RNGs differ


def _as_list(o):
    return list(o) if isinstance(o, (list, tuple)) else [o]


def _autocast_floor(name, sd, x, training, kw):
    """Outputs of the oracle under torch's CPU bf16 autocast: the noise floor of bf16 compute on the reference math."""
    with torch.no_grad(), torch.autocast("cpu", dtype=torch.bfloat16):
        return [o.float() for o in _as_list(O.FORWARDS[name](copy.deepcopy(sd), x, training, **kw))]


def _step(model, x, mask):
    model.cuda().train()
    outs = _as_list(model(x.cuda()))
    loss = sum(O.bce_with_logits(o, mask.cuda()) for o in outs)
    loss.backward()
    torch.cuda.synchronize()
    return [o.detach().cpu() for o in outs], float(loss)


@pytest.mark.parametrize("tag,kw", CASES)
def test_train_step_against_golden_and_oracle(tag, kw):
    name = tag.replace("_ds", "")
    gold = load_golden(tag)
    model = seeded_model(name, **kw)
    sd = copy.deepcopy(model.state_dict())
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    outs, loss = _step(model, x, mask)
    assert len(outs) == len(gold["logits_train"])
    floor = _autocast_floor(name, sd, x, True, kw)
    tols = [max(LOGIT_TOL_BF16, 1.3 * rel_err(f, g)) for f, g in zip(floor, gold["logits_train"])]
    for o, g, tol in zip(outs, gold["logits_train"], tols):
        assert o.shape == g.shape and o.dtype == torch.float32
        assert rel_err(o, g) < tol, (rel_err(o, g), tol)                   # vs the reference's own output
    assert abs(loss - gold["loss"]) < 3e-3 * len(outs)
    ref_out, _, _ = O.train_step(name, sd, x, mask, **kw)
    for o, r, tol in zip(outs, _as_list(ref_out), tols):
        assert rel_err(o, r.detach()) < tol
    got = model.state_dict()
    for k, v in gold["bn_after"].items():                                 # running statistics after one step
        if "running_" in k:
            assert rel_err(got[k].cpu(), v) < 2e-2, k
        else:
            assert int(got[k]) == int(v), k
    named = dict(model.named_parameters())
    for k, p in named.items():
        assert p.grad is not None and p.grad.shape == p.shape and bool(torch.isfinite(p.grad).all()), k
    # small, well-conditioned gradients (classifier head) against the reference's fp32 values
    for k, g in gold["grad_small"].items():
        if k.startswith(("final", "conv_1x1", "output_layer", "outc")):
            got_g = named[k].grad.cpu()
            assert float((got_g - g).norm() / g.norm().clamp_min(1e-20)) < max(3e-2, max(tols)), k


@pytest.mark.parametrize("tag,kw", CASES)
def test_gradients_no_noisier_than_bf16_autocast_of_the_reference(tag, kw):
    """Same criterion as for `unet`: per weight tensor, rel. L2 error vs fp32 within 1.3x of what torch's own bf16
    autocast of the reference math shows (+2 %), on a well-conditioned loss."""
    name = tag.replace("_ds", "")
    model = seeded_model(name, **kw)
    sd = copy.deepcopy(model.state_dict())
    x, mask = O.synthetic_batch(2, 3, 64, 64)
    model.cuda().train()
    outs = _as_list(model(x.cuda()))
    sum(O.projection_loss(o, mask.cuda()) for o in outs).backward()
    torch.cuda.synchronize()
    _, _, g32 = O.train_step(name, copy.deepcopy(sd), x, mask, loss_fn=O.projection_loss, **kw)
    with torch.autocast("cpu", dtype=torch.bfloat16):
        _, _, g16 = O.train_step(name, copy.deepcopy(sd), x, mask, loss_fn=O.projection_loss, **kw)
    named = dict(model.named_parameters())
    worst = []
    for k, ref in g32.items():
        if not k.endswith("weight") or ref.dim() < 2:
            continue
        den = ref.double().norm().clamp_min(1e-20)
        floor = float((g16[k].double() - ref.double()).norm() / den)
        ours = float((named[k].grad.cpu().double() - ref.double()).norm() / den)
        if ours > 1.3 * floor + 0.02:
            worst.append((k, ours, floor))
    assert not worst, worst[:8]


@pytest.mark.parametrize("tag,kw", CASES)
def test_eval_forward_and_mask_agreement(tag, kw):
    name = tag.replace("_ds", "")
    gold = load_golden(tag)
    model = seeded_model(name, **kw)
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    _step(model, x, mask)                                    # one step of running-stat updates, like the fixture
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        outs = [o.cpu() for o in _as_list(model(x.cuda()))]
        refs = _as_list(O.FORWARDS[name](sd, x, False, **kw))
    floor = _autocast_floor(name, sd, x, False, kw)
    for o, r, g, fl in zip(outs, refs, gold["logits_eval"], floor):
        assert rel_err(o, r) < LOGIT_TOL_BF16
        assert rel_err(o, g) < 3e-2                          # fixture used the reference's own fp32 statistics
        agree = float((O.threshold_mask(o) == O.threshold_mask(r)).float().mean())
        # 99.9 % bar; resunet's random-init logits straddle the threshold so that torch's own bf16 autocast of the
        # reference reaches only 99.85 % - never demand more than that floor (minus 0.1 %)
        floor_agree = float((O.threshold_mask(fl) == O.threshold_mask(r)).float().mean())
        assert agree >= min(0.999, floor_agree - 0.001), (agree, floor_agree)


def test_resunet_inference_shapes_of_the_sweep():
    """BASELINE config #5: resunet inference over batch / resolution (a few corners of the sweep)."""
    model = seeded_model("resunet").cuda().eval()
    for n, s in ((1, 256), (4, 384), (2, 512)):
        with torch.no_grad():
            y = model(torch.randn(n, 3, s, s, device="cuda"))
        assert tuple(y.shape) == (n, 1, s, s) and bool(torch.isfinite(y).all())


def test_transatt_unet_graphed_step_with_fused_tail_matches_eager_steps():
    """The stock-torch bottleneck island inside one CUDA graph: three replayed steps of transatt_unet (fused criterion,
    FlatAdamW - the island's `pos.*` / `pam.*` gradients must be slices of the same flat buffer) reproduce three eager
    steps.  Attention dropout off: its Philox offsets differ between capture and eager issue order."""
    from unet_zoo_b200.loss import BCEDiceWithLogits
    from unet_zoo_b200.optim import FlatAdamW
    from unet_zoo_b200.train_step import GraphedTrainStep
    x, m = O.synthetic_batch(2, 3, 64, 64)
    x, m = x.cuda(), m.cuda()
    results = []
    for graphed in (False, True):
        model = seeded_model("transatt_unet", attn_dropout=0.0).cuda().train()
        with torch.no_grad():
            model.pam.gamma.fill_(0.5)                       # zero at init: let the PAM branch carry gradient
        params = list(model.parameters())
        opt = FlatAdamW(params, lr=1e-3, weight_decay=1e-5, max_norm=1.0)
        crit = BCEDiceWithLogits()
        losses = []
        if graphed:
            sd0 = copy.deepcopy(model.state_dict())
            step = GraphedTrainStep(model, crit, opt, x, m, max_norm=None, params=params, warmup=1)
            model.load_state_dict(sd0)
            opt.exp_avg.zero_()
            opt.exp_avg_sq.zero_()
            opt.state.zero_()
            for _ in range(3):
                losses.append(float(step(x, m)))
        else:
            for _ in range(3):
                opt.zero_grad(set_to_none=True)
                loss = crit(model(x), m)
                loss.backward()
                opt.step()
                losses.append(float(loss))
            assert float(model.pam.query_conv.weight.grad.abs().sum()) > 0      # the island's parameters do train
            assert float(model.pos.row_embed.weight.grad[:4].abs().sum()) > 0
            assert float(model.pos.row_embed.weight.grad[4:].abs().sum()) == 0   # rows beyond the 4x4 map: untouched
        torch.cuda.synchronize()
        results.append((losses, {k: v.detach().clone() for k, v in model.state_dict().items()}))
    (l0, s0), (l1, s1) = results
    assert max(abs(a - b) for a, b in zip(l0, l1)) < 1e-5, (l0, l1)
    for k in s0:
        if s0[k].dtype.is_floating_point:
            assert torch.allclose(s0[k], s1[k], rtol=0, atol=2e-5), k
