"""Host-side logic of the per-epoch loops and the checkpoint manager (unet_zoo_b200/training.py) on CPU with a
stand-in module: same return values as the reference's loops (training_loop.py:86-180), which read loss / Dice on the
host every batch, and the reference's checkpoint key format (multi_gpu.py:39-87)."""
import torch
import torch.nn as nn
import torch.nn.functional as F

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


def test_u2net_dict_outputs_sum_seven_losses_with_weight_one():
    """training_loop.py:23-31,61-64: U2NET's dict - every map against the mask resized to the map's size (bilinear,
    align_corners=False), all weights 1.0, 'main' is the prediction the Dice score is taken from."""
    torch.manual_seed(3)
    t = (torch.rand(2, 1, 16, 16) > 0.5).float()
    keys = ["main"] + [f"side{i}" for i in range(1, 7)]
    outs = {k: torch.randn(2, 1, 16, 16, requires_grad=True) for k in keys}
    outs["side6"] = torch.randn(2, 1, 8, 8, requires_grad=True)          # a map at another size: the mask is resized
    crit = nn.BCEWithLogitsLoss()
    loss, main, dice = T._loss_and_dice(outs, t, crit)
    want = sum(crit(o, F.interpolate(t, size=o.shape[2:], mode="bilinear", align_corners=False)) for o in outs.values())
    assert abs(float(loss) - float(want)) < 1e-6
    assert main is outs["main"]
    assert abs(float(dice) - float(O.dice_coefficient(outs["main"].detach(), t))) < 1e-6


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


def test_train_model_epoch_driver_and_checkpoints(tmp_path):
    """train_model (training_loop.py:182-285): same return tuple, last checkpoint every epoch, best checkpoint on
    improvement (the reference never writes it), injected early stopping / scheduler objects with the reference's call
    shapes are honoured."""
    class Cfg:
        EPOCHS, DEVICE = 4, torch.device("cpu")

    class Stop:                                              # early_stopping.py call shape
        def __init__(self):
            self.calls, self.early_stop = [], False

        def __call__(self, score, model, epoch):
            self.calls.append((round(score, 6), epoch))
            self.early_stop = epoch >= 3

        def get_best_score(self):
            return max(c[0] for c in self.calls)

    class Sched:                                             # lr_scheduler.py call shape
        def __init__(self, opt):
            self.opt, self.calls = opt, []

        def step(self, score, epoch):
            self.calls.append(epoch)
            self.opt.param_groups[0]["lr"] *= 0.5

    torch.manual_seed(0)
    model = _Tiny()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=1e-5)
    stop, sched = Stop(), Sched(opt)
    best, last = str(tmp_path / "best.pth"), str(tmp_path / "last.pth")
    mgr = T.MultiGPUManager(device="cpu")
    out = T.train_model(model, _batches(3), _batches(2, seed=1), opt, nn.BCEWithLogitsLoss(), Cfg, "tiny", best, last,
                        logger=None, gpu_manager=mgr, early_stopping=stop, lr_scheduler=sched)
    tl, td, vl, vd, early = out
    assert early and len(tl) == len(td) == len(vl) == len(vd) == 3       # stopped by the injected object at epoch 3
    assert sched.calls == [1, 2, 3] and abs(opt.param_groups[0]["lr"] - 1e-2 / 8) < 1e-12
    assert [c[1] for c in stop.calls] == [1, 2, 3] and abs(stop.calls[-1][0] - round(vd[-1], 6)) < 1e-6
    sd_last, sd_best = torch.load(last), torch.load(best)
    assert set(sd_last) == set(model.state_dict()) == set(sd_best)
    for k, v in model.state_dict().items():
        assert torch.equal(sd_last[k], v)                                # last = the final weights
    fresh = _Tiny()
    mgr.load_model_state(fresh, best)
    assert abs(T.validate_one_epoch(fresh, _batches(2, seed=1), nn.BCEWithLogitsLoss(), torch.device("cpu"))[1] - max(vd)) < 1e-6


def test_init_distributed_single_process():
    rank, world, device = T.init_distributed()
    assert (rank, world) == (0, 1) and device.type in ("cpu", "cuda")
