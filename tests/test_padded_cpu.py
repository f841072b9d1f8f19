"""Host logic of the zero-padded operand layout of `multiresunet` (unet_zoo_b200/models/_padded.py) on CPU: parameters ->
padded stand-ins, padded gradients -> the flat gradient buffer, running statistics both ways.  Pure index arithmetic -
bit-exact."""
import torch

from conftest import seeded_model
from unet_zoo_b200 import engine as E
from unet_zoo_b200.models._padded import Lay, Shadow, ceil8


def test_layout_arithmetic():
    lay = Lay(8, 17, 26)                       # Multiresblock 1: 51 real channels in [8 | 24 | 32]
    assert (lay.real, lay.padded) == (51, 64)
    idx = lay.index().tolist()
    assert idx[:8] == list(range(8)) and idx[8:25] == list(range(8, 25)) and idx[25:] == list(range(32, 58))
    assert lay.offsets() == [(0, 8), (8, 24), (32, 32)]
    assert [ceil8(c) for c in (1, 8, 9, 853)] == [8, 8, 16, 856]
    assert Lay(142, 284, 427).padded == 864


def test_specs_cover_every_parameter_and_buffer():
    model = seeded_model("multiresunet")
    specs = model._specs()
    mods = {id(s[1]) for s in specs}
    for m in model.modules():
        if isinstance(m, (torch.nn.Conv2d, torch.nn.ConvTranspose2d, torch.nn.BatchNorm2d)):
            assert id(m) in mods
    for kind, mod, *lays in specs:
        if kind == "conv":
            assert (mod.out_channels, mod.in_channels) == (lays[1].real, lays[0].real)
        elif kind == "convt":
            assert (mod.in_channels, mod.out_channels) == (lays[0].real, lays[1].real)
        else:
            assert mod.num_features == lays[0].real and not mod.affine


def test_scatter_gather_round_trip():
    model = seeded_model("multiresunet", in_channels=1, num_classes=3)
    sh = Shadow(model, model._specs(), torch.device("cpu"))
    with torch.no_grad():
        model.multiresblock2.batch_norm1.running_mean.uniform_(-1, 1)
        model.multiresblock2.batch_norm1.running_var.uniform_(0.5, 2)
    sh.refresh(True)
    # a convolution behind a concatenation: rows follow the output layout, columns the input layout, the rest is zero
    c = model.multiresblock2.conv2d_bn_1x1.conv1
    s = sh.of[id(c)]
    im, om = Lay(8, 17, 26).index(), Lay(17, 35, 53).index()
    assert s.weight.shape == (120, 64, 1, 1) and (s.in_channels, s.out_channels) == (64, 120)
    assert torch.equal(s.weight.detach()[om][:, im], c.weight.detach())
    assert float(s.weight.detach().abs().sum()) == float(c.weight.detach().abs().sum())
    assert torch.equal(s.bias.detach()[om], c.bias.detach()) and float(s.bias.detach().abs().sum()) == float(c.bias.detach().abs().sum())
    assert s.weight.requires_grad and s.bias.requires_grad
    # ConvTranspose2d: input channels are dim 0
    u = model.upsample6
    su = sh.of[id(u)]
    assert su.weight.shape == (864, 256, 2, 2)
    assert torch.equal(su.weight.detach()[Lay(142, 284, 427).index()], u.weight.detach())
    # running statistics: padded mean 0 / variance 1
    bn = model.multiresblock2.batch_norm1
    sb = sh.of[id(bn)]
    assert sb.weight is None and sb.bias is None and sb.num_batches_tracked is bn.num_batches_tracked
    assert torch.equal(sb.running_mean[om], bn.running_mean) and torch.equal(sb.running_var[om], bn.running_var)
    pad = torch.ones(120, dtype=torch.bool)
    pad[om] = False
    assert float(sb.running_mean[pad].abs().sum()) == 0.0 and bool((sb.running_var[pad] == 1.0).all())
    sb.running_mean += 1.0
    sh.write_back_stats()
    assert torch.equal(bn.running_mean, sb.running_mean[om])
    # an optimizer step is picked up by the next refresh
    with torch.no_grad():
        c.weight.mul_(2.0)
    sh.refresh(True)
    assert torch.equal(s.weight.detach()[om][:, im], c.weight.detach())
    # gradients: every padded tensor's gradient lands in the parameter's slice of the flat buffer
    params = list(model.parameters())
    ready = []
    pg = E.ParamGrads(params, torch.device("cpu"), on_ready=lambda p: ready.append(id(p)))
    sh.begin_backward(pg)
    for (t, _off) in sh.tensors:
        pg.view(t).copy_(t.detach() * 3.0)
        pg.ready(t)                              # stand-ins do not report: the gather does, once per parameter
    assert ready == []
    sh.end_backward(pg, params)
    for p in params:
        assert torch.equal(pg.view(p), p.detach() * 3.0)
    assert ready == [id(p) for p in params]


def test_eval_refresh_is_skipped_while_nothing_changed():
    model = seeded_model("multiresunet").eval()
    sh = Shadow(model, model._specs(), torch.device("cpu"))
    sh.refresh(False)
    v0 = sh.wflat._version
    sh.refresh(False)
    assert sh.wflat._version == v0               # the executor's inference memo stays valid
    with torch.no_grad():
        model.conv_final.conv1.weight.add_(1.0)
    sh.refresh(False)
    assert sh.wflat._version > v0
    s = sh.of[id(model.conv_final.conv1)]
    assert torch.equal(s.weight.detach()[:1, Lay(8, 17, 26).index()], model.conv_final.conv1.weight.detach())
