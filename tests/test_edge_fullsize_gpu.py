"""Edge cases of the hot path against the CPU oracle (ragged / minimal shapes, other channel and class counts, error
behaviour) and size-independent properties at BASELINE.json's full size (unet, batch 32, 512x512), where the oracle
cannot run in seconds: determinism, batch-split equivalence, analytic head-bias gradient, zero conv-bias gradient."""
import copy

import pytest
import torch

from conftest import conditioned_model, parity_report, rel_err, seeded_model
from oracle import reference_port as O

pytestmark = pytest.mark.gpu
LOGIT_TOL_BF16 = 2e-2


def _oracle_logits(name, model, x, training, **kw):
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        out = O.FORWARDS[name](sd, x, training, **kw)
    return out if isinstance(out, (list, tuple)) else [out]


@pytest.mark.parametrize("name,kw,shape", [
    ("unet", {}, (1, 3, 16, 16)),                                  # smallest legal image: 1x1 bottleneck
    ("unet", {}, (3, 3, 48, 80)),                                  # odd batch, non-square, not a multiple of the tile
    ("unet", {"in_channels": 1}, (2, 1, 32, 32)),
    ("unet", {"in_channels": 9, "num_classes": 3}, (2, 9, 32, 48)),   # 9*cin > 64: NHWC stem path; fused head K = 3
    ("unet", {"num_classes": 6}, (1, 3, 32, 32)),                  # K > 4: unfused head
    ("attention_unet", {"num_classes": 2}, (1, 3, 32, 64)),
    ("nested_unet", {"in_channels": 4}, (2, 4, 48, 32)),
    ("resunet", {}, (3, 3, 32, 48)),
    ("resunet", {"in_channels": 12}, (1, 12, 16, 24)),
])
def test_eval_forward_edge_shapes_match_oracle(name, kw, shape):
    model = seeded_model(name, **kw)
    g = torch.Generator().manual_seed(7)
    x = torch.randn(*shape, generator=g)
    with torch.no_grad():                                          # non-trivial running statistics
        for m in model.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.normal_(0, 0.1, generator=g)
                m.running_var.uniform_(0.5, 1.5, generator=g)
    model.cuda().eval()
    with torch.no_grad():
        outs = model(x.cuda())
    outs = [o.cpu() for o in (outs if isinstance(outs, (list, tuple)) else [outs])]
    refs = _oracle_logits(name, model, x, False)
    for o, r in zip(outs, refs):
        assert o.shape == r.shape and o.dtype == torch.float32
        assert rel_err(o, r) < LOGIT_TOL_BF16


@pytest.mark.parametrize("name,kw,shape", [("unet", {"num_classes": 2}, (2, 3, 32, 48)),
                                           ("unet", {"in_channels": 1}, (1, 1, 64, 32)),
                                           ("unet", {"in_channels": 10}, (2, 10, 32, 32)),
                                           ("attention_unet", {}, (2, 3, 32, 32))])
def test_train_step_edge_shapes_match_oracle(name, kw, shape):
    model, sd = conditioned_model(name, **kw)                  # well-conditioned state: plain bars (see conftest)
    g = torch.Generator().manual_seed(11)
    x = torch.randn(*shape, generator=g)
    K = kw.get("num_classes", 1)
    mask = (torch.rand(shape[0], K, *shape[2:], generator=g) > 0.5).float()
    model.cuda().train()
    out = model(x.cuda())
    O.bce_with_logits(out, mask.cuda()).backward()
    torch.cuda.synchronize()
    ref_out, ref_loss, ref_grads = O.train_step(name, sd, x, mask)
    tol = LOGIT_TOL_BF16
    assert rel_err(out.detach().cpu(), ref_out.detach()) < tol
    named = dict(model.named_parameters())
    head = [k for k in ref_grads if k.startswith(("out.", "conv_1x1"))]
    assert head
    for k in head:
        gref = ref_grads[k]
        err = float((named[k].grad.cpu() - gref).norm() / gref.norm().clamp_min(1e-20))
        assert err < 3e-2, (k, err)
    assert all(bool(torch.isfinite(p.grad).all()) for p in model.parameters())
    # the stem weight gradient (im2col stem, or NHWC stem with zero-padded channels) against fp32: the deepest gradient
    # of the model, fixed bar (bf16 rounding accumulated over the whole backward pass)
    k0 = next(k for k in ref_grads if k.endswith("weight") and ref_grads[k].dim() == 4)
    den = ref_grads[k0].double().norm().clamp_min(1e-20)
    ours_g = float((named[k0].grad.cpu().double() - ref_grads[k0].double()).norm() / den)
    parity_report("edge_train/" + name + str(shape), logits=rel_err(out.detach().cpu(), ref_out.detach()), stem_grad=ours_g)
    assert named[k0].grad.shape == ref_grads[k0].shape
    assert ours_g < 0.35, (k0, ours_g)


def test_bad_inputs_raise_like_the_contract_says():
    model = seeded_model("unet").cuda()
    with pytest.raises(RuntimeError):
        model(torch.randn(1, 3, 8, 40, device="cuda"))             # four poolings need H, W >= 16
    with pytest.raises(RuntimeError, match="multiples of 16"):      # the reference's attention_unet fails here too
        seeded_model("attention_unet").cuda()(torch.randn(1, 3, 40, 40, device="cuda"))
    with pytest.raises(RuntimeError):
        model(torch.randn(1, 4, 32, 32, device="cuda"))            # wrong channel count
    with pytest.raises(RuntimeError):
        model(torch.randn(1, 3, 32, 32))                           # CPU tensor: no CPU fallback
    with pytest.raises(RuntimeError):
        model(torch.randn(3, 32, 32, device="cuda"))               # not NCHW


# ------------------------------------------------------------------------------------------ full size (BASELINE configs[1])
@pytest.fixture(scope="module")
def full():
    model = seeded_model("unet").cuda()
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(32, 3, 512, 512, generator=g).cuda()
    mask = (torch.rand(32, 1, 512, 512, generator=g) > 0.5).float().cuda()
    return model, x, mask


def test_fullsize_eval_is_deterministic_and_batch_separable(full):
    model, x, _ = full
    model.eval()
    with torch.no_grad():
        a = model(x)
        b = model(x)
        parts = torch.cat([model(x[i:i + 8]) for i in range(0, 32, 8)])
    assert a.shape == (32, 1, 512, 512) and bool(torch.isfinite(a).all())
    assert torch.equal(a, b)                                        # same launch twice: bit-exact
    assert torch.equal(a, parts)                                    # images are independent in eval mode: bit-exact
    # thresholding is a sign test on the fp32 logits (metrics.py:11): identical masks
    assert torch.equal(O.threshold_mask(a), O.threshold_mask(parts))


def test_fullsize_train_step_properties(full):
    model, x, mask = full
    model.train()
    grads = []
    for _ in range(2):
        model.zero_grad(set_to_none=True)
        out = model(x)
        loss = torch.nn.functional.binary_cross_entropy_with_logits(out, mask)
        loss.backward()
        grads.append({k: p.grad.clone() for k, p in model.named_parameters()})
    torch.cuda.synchronize()
    assert bool(torch.isfinite(loss))
    for k in grads[0]:                                              # static schedules, no atomics: bit-exact repeat
        assert torch.equal(grads[0][k], grads[1][k]), k
    # head bias gradient = sum dL/dlogits = sum (sigmoid(z) - y) / numel, computable from the logits alone
    want = ((torch.sigmoid(out.detach().double()) - mask.double()) / mask.numel()).sum()
    got = grads[0]["out.conv.bias"].double().sum()
    assert abs(float(got - want)) < 1e-6 + 1e-4 * abs(float(want))
    # a conv bias in front of a train-mode BatchNorm has zero gradient (BASELINE a1)
    for k, gk in grads[0].items():
        if k.endswith("conv_op.0.bias") or k.endswith("conv_op.3.bias"):
            assert float(gk.abs().max()) < 1e-4, k
