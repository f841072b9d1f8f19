"""Host-side logic of the per-epoch loops and the checkpoint manager (unet_zoo_b200/training.py) on CPU with a
stand-in module: same return values as the reference's loops (training_loop.py:86-180), which read loss / Dice on the
host every batch, and the reference's checkpoint key format (multi_gpu.py:39-87)."""
import torch
import torch.nn as nn

from oracle import reference_port as O
from unet_zoo_b200 import training as T


class _Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(3, 1, 3, padding=1)
        self.bn = nn.BatchNorm2d(1)

    def forward(self, x):
        return self.bn(self.conv(x))


def _batches(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(2, 3, 16, 16, generator=g), (torch.rand(2, 1, 16, 16, generator=g) > 0.5).float(), "path")
            for _ in range(n)]


def test_dice_matches_the_oracle_including_empty_union():
    g = torch.Generator().manual_seed(3)
    z = torch.randn(2, 1, 8, 8, generator=g)
    t = (torch.rand(2, 1, 8, 8, generator=g) > 0.5).float()
    assert abs(float(T.dice_coefficient(z, t)) - float(O.dice_coefficient(z, t))) < 1e-7
    assert float(T.dice_coefficient(torch.full((1, 1, 4, 4), -3.0), torch.zeros(1, 1, 4, 4))) == 1.0


def test_epoch_loops_return_what_the_reference_loops_return():
    data = _batches(4)
    torch.manual_seed(0)
    a, b = _Tiny(), _Tiny()
    b.load_state_dict(a.state_dict())
    crit = nn.BCEWithLogitsLoss()
    opt_a = torch.optim.AdamW(a.parameters(), lr=1e-2, weight_decay=1e-5)
    opt_b = torch.optim.AdamW(b.parameters(), lr=1e-2, weight_decay=1e-5)
    got = T.train_one_epoch(a, data, opt_a, crit, torch.device("cpu"), None, 0, "tiny", None)
    # the reference's loop body, verbatim semantics (training_loop.py:108-124,143-145)
    b.train()
    rl = rd = 0.0
    for img, mask, _ in data:
        opt_b.zero_grad()
        out = b(img.float())
        loss = crit(out, mask.float())
        dc = O.dice_coefficient(out, mask)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(b.parameters(), 1.0)
        opt_b.step()
        rl += loss.item()
        rd += float(dc)
    assert abs(got[0] - rl / len(data)) < 1e-6 and abs(got[1] - rd / len(data)) < 1e-6
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-7)
    va = T.validate_one_epoch(a, data, crit, torch.device("cpu"), "tiny", None)
    b.eval()
    with torch.no_grad():
        vl = sum(crit(b(i), m).item() for i, m, _ in data) / len(data)
        vd = sum(float(O.dice_coefficient(b(i), m)) for i, m, _ in data) / len(data)
    assert abs(va[0] - vl) < 1e-6 and abs(va[1] - vd) < 1e-6
    assert not a.training                                      # validate leaves the model in eval mode, like the reference


def test_deep_supervision_lists_sum_the_head_losses():
    crit = nn.BCEWithLogitsLoss()
    g = torch.Generator().manual_seed(1)
    outs = [torch.randn(1, 1, 8, 8, generator=g) for _ in range(4)]
    t = (torch.rand(1, 1, 8, 8, generator=g) > 0.5).float()
    loss, main, dice = T._loss_and_dice(outs, t, crit)
    assert abs(float(loss) - sum(float(crit(o, t)) for o in outs)) < 1e-6
    assert main is outs[-1] and abs(float(dice) - float(O.dice_coefficient(outs[-1], t))) < 1e-7


def test_checkpoint_manager_roundtrip_and_module_prefix(tmp_path, capsys):
    mgr = T.MultiGPUManager(device="cpu")
    torch.manual_seed(0)
    m = mgr.setup_model_for_gpu(_Tiny())
    path = str(tmp_path / "last.pth")
    mgr.save_model_state(m, path)
    sd = torch.load(path)
    assert list(sd) == list(m.state_dict()) and not any(k.startswith("module.") for k in sd)
    torch.save({"module." + k: v for k, v in sd.items()}, path)            # a checkpoint written from a wrapped model
    torch.manual_seed(1)
    fresh = _Tiny()
    mgr.load_model_state(fresh, path, torch.device("cpu"))
    for k, v in fresh.state_dict().items():
        assert torch.equal(v, sd[k]), k
    assert mgr.load_model_state(fresh, str(tmp_path / "missing.pth"), "cpu") is fresh
    assert "not found" in capsys.readouterr().out
    wrapped = nn.DataParallel(fresh) if False else fresh
    assert mgr.get_model_for_inference(wrapped, "cpu").training is False
