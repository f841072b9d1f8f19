"""The drop-in epoch loops on the GPU with the hot-path model, the fused criterion and the flat optimizer, against the
reference's loop body run with torch's own loss / clip / AdamW on a second copy of the same model."""
import copy

import pytest
import torch

from conftest import seeded_model
from oracle import reference_port as O
from unet_zoo_b200 import training as T
from unet_zoo_b200.loss import BCEDiceWithLogits
from unet_zoo_b200.optim import FlatAdamW

pytestmark = pytest.mark.gpu


def _batches(n):
    out = []
    for i in range(n):
        g = torch.Generator().manual_seed(100 + i)
        out.append((torch.randn(2, 3, 64, 64, generator=g), (torch.rand(2, 1, 64, 64, generator=g) > 0.5).float(), f"p{i}"))
    return out


def test_train_and_validate_epoch_with_fused_tail_match_the_torch_tail():
    dev = torch.device("cuda", 0)
    data = _batches(3)
    a = seeded_model("unet").to(dev)
    b = seeded_model("unet").to(dev)
    opt_a = FlatAdamW(a.parameters(), lr=1e-4, weight_decay=1e-5, max_norm=1.0)
    opt_b = torch.optim.AdamW(b.parameters(), lr=1e-4, weight_decay=1e-5)
    got = T.train_one_epoch(a, data, opt_a, BCEDiceWithLogits(), dev, None, 0, "unet", None)
    crit = torch.nn.BCEWithLogitsLoss()
    b.train()
    rl = rd = 0.0
    for img, mask, _ in data:                                   # training_loop.py:108-124
        img, mask = img.to(dev), mask.to(dev)
        opt_b.zero_grad()
        out = b(img)
        loss = crit(out, mask)
        dc = O.dice_coefficient(out.detach().cpu(), mask.cpu())
        loss.backward()
        torch.nn.utils.clip_grad_norm_(b.parameters(), 1.0)
        opt_b.step()
        rl += loss.item()
        rd += float(dc)
    assert abs(got[0] - rl / 3) < 1e-5 and abs(got[1] - rd / 3) < 1e-6, (got, rl / 3, rd / 3)
    for (k, p), q in zip(a.named_parameters(), b.parameters()):
        assert float((p - q).abs().max()) <= 1e-5 * float(q.abs().max()) + 1e-8, k
    va = T.validate_one_epoch(a, data, BCEDiceWithLogits(), dev, "unet", None)
    vb = T.validate_one_epoch(b, data, crit, dev, "unet", None)           # torch criterion + torch-op Dice path
    assert abs(va[0] - vb[0]) < 1e-4 and abs(va[1] - vb[1]) < 2e-3, (va, vb)
    # the eval numbers are those of the fp32 oracle on the same weights within the bf16 bar
    sd = {k: v.detach().cpu() for k, v in a.state_dict().items()}
    ref = sum(float(O.bce_with_logits(O.FORWARDS["unet"](copy.deepcopy(sd), i, False), m)) for i, m, _ in data) / 3
    assert abs(va[0] - ref) < 2e-2 * max(1.0, abs(ref))


def test_manager_wraps_nothing_on_one_gpu_and_checkpoints_roundtrip(tmp_path):
    mgr = T.MultiGPUManager(device="cuda:0")
    m = mgr.setup_model_for_gpu(seeded_model("resunet"))
    assert next(m.parameters()).is_cuda and not hasattr(m, "module")
    path = str(tmp_path / "resunet_last.pth")
    mgr.save_model_state(m, path)
    sd = torch.load(path)
    assert len(sd) == 127 and all(v.device.type == "cpu" for v in sd.values())
    fresh = mgr.setup_model_for_gpu(seeded_model("resunet"))
    with torch.no_grad():
        for p in fresh.parameters():
            p.add_(1.0)
    mgr.load_model_state(fresh, path, torch.device("cuda", 0))
    for k, v in fresh.state_dict().items():
        assert torch.equal(v.cpu(), sd[k]), k
