"""Training-step tail on the GPU (csrc/tail.cu through the C ABI): fused BCE-with-logits + Dice + loss gradient
against nn.BCEWithLogitsLoss / the oracle's dice_coefficient (metrics.py:7-24), and clip_grad_norm_ + AdamW over flat
buffers against torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW (training_loop.py:120-121)."""
import copy

import pytest
import torch

from conftest import seeded_model
from oracle import reference_port as O
from unet_zoo_b200.loss import BCEDiceWithLogits
from unet_zoo_b200.optim import FlatAdamW

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(2, 1, 64, 64), (3, 1, 17, 23), (1, 2, 5, 7), (32, 1, 512, 512)])
def test_bce_dice_matches_torch_and_oracle(shape):
    g = torch.Generator().manual_seed(5)
    z = (torch.randn(*shape, generator=g) * 3).cuda().requires_grad_(True)
    y = (torch.rand(*shape, generator=g) > 0.6).float().cuda()
    crit = BCEDiceWithLogits()
    loss = crit(z, y)
    (loss * 1.7).backward()                                    # upstream gradient other than 1
    z2 = z.detach().clone().requires_grad_(True)
    ref = torch.nn.BCEWithLogitsLoss()(z2, y)
    (ref * 1.7).backward()
    assert abs(float(loss) - float(ref)) < 2e-6 * max(1.0, abs(float(ref)))
    assert float((z.grad - z2.grad).abs().max()) <= 1e-6 * float(z2.grad.abs().max()) + 1e-12
    dice_ref = O.dice_coefficient(z.detach().cpu(), y.cpu())
    assert abs(float(crit.last_dice) - float(dice_ref)) < 1e-6
    # thresholding is bit-exact with torch.sigmoid(z) > 0.5 on the same logits: identical counts
    mask = torch.sigmoid(z.detach()) > 0.5
    inter, pred, targ = [float(v) for v in crit.last_stats]
    assert pred == float(mask.sum()) and targ == float(y.sum()) and inter == float((mask.float() * y).sum())


def test_bce_dice_edge_cases():
    crit = BCEDiceWithLogits()
    z = torch.full((1, 1, 4, 4), -5.0, device="cuda")
    crit(z, torch.zeros_like(z))
    assert float(crit.last_dice) == 1.0                        # empty union -> 1.0 (metrics.py:18-19), no host branch
    crit(-z, torch.ones_like(z))
    assert abs(float(crit.last_dice) - 1.0) < 1e-6
    crit(z, torch.ones_like(z))
    assert float(crit.last_dice) < 1e-6
    big = torch.tensor([[[[80.0, -80.0, 0.0, 1e-9]]]], device="cuda")      # saturated / threshold-straddling logits
    t = torch.tensor([[[[1.0, 0.0, 1.0, 0.0]]]], device="cuda")
    loss = crit(big, t)
    ref = torch.nn.BCEWithLogitsLoss()(big, t)
    assert torch.isfinite(loss) and abs(float(loss) - float(ref)) < 1e-6
    assert float(crit.last_stats[1]) == float((torch.sigmoid(big) > 0.5).sum())
    with pytest.raises(RuntimeError):
        crit(torch.zeros(2, 2), torch.zeros(2, 2))             # CPU tensors: no CPU path


def test_flat_adamw_matches_clip_plus_torch_adamw():
    """Three training steps of unet: FlatAdamW(max_norm=1) vs clip_grad_norm_(1.0) + torch.optim.AdamW fed with the
    same gradients."""
    model = seeded_model("unet").cuda().train()
    opt = FlatAdamW(model.parameters(), lr=1e-3, weight_decay=1e-2, max_norm=1.0)
    shadow = [p.detach().clone().requires_grad_(True) for p in model.parameters()]
    ref_opt = torch.optim.AdamW(shadow, lr=1e-3, weight_decay=1e-2)
    crit = BCEDiceWithLogits()
    x, mask = O.synthetic_batch(2, 3, 64, 64)
    x, mask = x.cuda(), mask.cuda()
    for step in range(3):
        opt.zero_grad()
        loss = crit(model(x), mask)
        loss.backward()
        for s, p in zip(shadow, model.parameters()):
            s.grad = p.grad.detach().clone()
        ref_norm = torch.nn.utils.clip_grad_norm_(shadow, 1.0)
        ref_opt.step()
        opt.step()
        assert abs(float(opt.total_norm) - float(ref_norm)) < 1e-5 * float(ref_norm)
        assert float(opt.step_count) == step + 1
        for s, p in zip(shadow, model.parameters()):
            assert float((p.detach() - s.detach()).abs().max()) <= 2e-6 * float(s.detach().abs().max()) + 1e-9
            # like clip_grad_norm_, the gradients the caller sees afterwards are the clipped ones
            assert float((p.grad - s.grad).abs().max()) <= 1e-5 * float(s.grad.abs().max()) + 1e-12
    sd = model.state_dict()
    assert len(sd) == 136 and all(v.is_contiguous() for v in sd.values())     # flat re-seating keeps the checkpoint format


def test_flat_adamw_rejects_foreign_gradients_and_cpu():
    with pytest.raises(RuntimeError, match="CUDA"):
        FlatAdamW([torch.nn.Parameter(torch.zeros(4))])
    ps = [torch.nn.Parameter(torch.randn(5, device="cuda")), torch.nn.Parameter(torch.randn(7, device="cuda"))]
    opt = FlatAdamW(ps, lr=1e-3)
    for p in ps:
        p.grad = torch.randn_like(p)                           # separate allocations: not one flat buffer
    with pytest.raises(RuntimeError, match="flat buffer"):
        opt.step()


def test_graphed_step_with_fused_tail_matches_eager():
    from unet_zoo_b200.train_step import GraphedTrainStep
    x, mask = O.synthetic_batch(2, 3, 64, 64)
    x, mask = x.cuda(), mask.cuda()
    losses = {}
    for mode in ("eager", "graph"):
        model = seeded_model("unet").cuda().train()
        opt = FlatAdamW(model.parameters(), lr=1e-4, weight_decay=1e-5, max_norm=1.0)
        crit = BCEDiceWithLogits()
        out = []
        if mode == "graph":
            sd0 = copy.deepcopy(model.state_dict())
            step = GraphedTrainStep(model, crit, opt, x, mask, max_norm=None, warmup=1)
            model.load_state_dict(sd0)                         # undo the warm-up / capture steps
            opt.exp_avg.zero_(), opt.exp_avg_sq.zero_(), opt.state.zero_()
            for _ in range(3):
                out.append((float(step(x, mask)), float(crit.last_dice)))
        else:
            for _ in range(3):
                opt.zero_grad()
                loss = crit(model(x), mask)
                loss.backward()
                opt.step()
                out.append((float(loss), float(crit.last_dice)))
        losses[mode] = out
    assert losses["eager"] == losses["graph"], losses
    assert losses["eager"][2][0] != losses["eager"][0][0]      # the parameters did move
