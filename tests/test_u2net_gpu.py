"""U^2-Net (`u2net`, `u2netp`) on the GPU: the operators it adds (dilated 3x3 convolution forward /
data gradient / weight gradient, bilinear resize with align_corners=False, residual add + pooling, the zero-padded side
convolutions and the fused seven-map head) against PyTorch fp32 on the same bf16 operands, and the whole models through
create_model against the fixtures generated from the UNMODIFIED reference (oracle/make_golden.py, 2 x 3 x 128 x 128) and
the CPU oracle."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from conftest import conditioned_model, grad_errors, load_golden, parity_report, rel_err, seeded_model
from gpu_util import bf16r, get_grad, make_act, new_ctx, relmax, rel_l2, run_tape, set_grad
from oracle import reference_port as O
from unet_zoo_b200 import engine as E

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
# Fixed bars of this file.  U^2-Net is the deepest model of the zoo (the longest path crosses ~60 train-mode conv-BN-ReLU
# layers, side maps whose logits are centred on the threshold): torch's own bf16 autocast of the reference deviates from
# fp32 by 1.9 % (u2net) / 2.6 % (u2netp) on the train logits, 1.7 % / 2.5 % in eval mode, mask agreement 99.5-99.9 %
# (tools/parity_floor.py u2net,u2netp 2 128 gauss 0.8,1.2,1.0,2.0); this path measures 2.1 % / 2.7 %, 1.6 % / 2.3 %,
# 99.5-99.9 % (profiles/r02_ao_u2net_parity.jsonl).  So the plain 2e-2 / 99.9 % bars of the shallower models are
# replaced here by 3e-2 / 99 % - a documented exception like attention_unet's train-mode bar.
LOGIT_TOL = 3e-2
MASK_AGREE = 0.99


@pytest.fixture(autouse=True)
def _exact_fp32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)


# ---------------------------------------------------------------------------------------------------- operators
@pytest.mark.parametrize("N,H,W,segs,cout,d", [(2, 32, 32, (64,), 64, 2), (2, 16, 24, (32, 32), 16, 4), (3, 16, 16, (256,), 128, 8),
                                               (2, 4, 4, (64,), 64, 8), (1, 40, 24, (16,), 32, 2)])
def test_dilated_conv_bn_relu_fwd_bwd(N, H, W, segs, cout, d):
    """REBNCONV with dirate d (u2net.py:6-17) in train mode: output, running statistics, data / weight / BatchNorm gradients."""
    cin = sum(segs)
    conv = nn.Conv2d(cin, cout, 3, padding=d, dilation=d).to(DEV)
    bn = nn.BatchNorm2d(cout).to(DEV)
    with torch.no_grad():
        conv.weight.copy_(bf16r(conv.weight))
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    x = bf16r(torch.randn(N, cin, H, W, device=DEV))
    ctx = new_ctx(params=[conv.weight, conv.bias, bn.weight, bn.bias])
    xa = make_act(ctx, x)
    xs, c0 = [], 0
    for c in segs:
        xs.append(E.Act(xa.buf, c0, c))
        c0 += c
    out = E.conv_bn_act(ctx, xs, conv, bn, dilation=d)
    xr = x.clone().requires_grad_(True)
    w = conv.weight.detach().clone().requires_grad_(True)
    g = bn.weight.detach().clone().requires_grad_(True)
    be = bn.bias.detach().clone().requires_grad_(True)
    z = F.conv2d(xr, w, conv.bias.detach(), padding=d, dilation=d)
    zq = z + (bf16r(z.detach()) - z.detach())
    rm, rv = torch.zeros(cout, device=DEV), torch.ones(cout, device=DEV)
    y = F.relu(F.batch_norm(zq, rm, rv, g, be, training=True, momentum=0.1, eps=1e-5))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), y.detach()) < 1.5e-2
    assert relmax(bn.running_mean, rm) < 3e-3 and relmax(bn.running_var, rv) < 3e-3
    dy = bf16r(torch.randn_like(y) * 0.1)
    (y * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    ctx.join_side()
    torch.cuda.synchronize()
    assert rel_l2(get_grad(xa), xr.grad) < 1.5e-2
    dw = ctx.pgrads.view(conv.weight)
    assert rel_l2(dw, w.grad) < 1e-2
    for t in range(9):                       # a wrong tap shift would hide in the norm
        if float(w.grad[:, :, t // 3, t % 3].norm()) > 0:
            assert rel_l2(dw[:, :, t // 3, t % 3], w.grad[:, :, t // 3, t % 3]) < 2e-2, t
        else:                                # taps that only ever see padding (4 x 4 image, dilation 8)
            assert float(dw[:, :, t // 3, t % 3].abs().max()) == 0, t
    assert rel_l2(ctx.pgrads.view(bn.weight), g.grad) < 1e-2
    assert rel_l2(ctx.pgrads.view(bn.bias), be.grad) < 1e-2


def test_dilated_conv_eval_mode_folds_batchnorm():
    conv = nn.Conv2d(64, 32, 3, padding=4, dilation=4).to(DEV)
    bn = nn.BatchNorm2d(32).to(DEV).eval()
    with torch.no_grad():
        bn.running_mean.normal_()
        bn.running_var.uniform_(0.5, 2.0)
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.normal_()
    x = bf16r(torch.randn(2, 64, 32, 16, device=DEV))
    ctx = new_ctx(training=False, record=False)
    out = E.conv_bn_act(ctx, [make_act(ctx, x)], conv, bn, dilation=4)
    ref = F.relu(bn(conv(x)))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), ref.detach()) < 1.5e-2


@pytest.mark.parametrize("N,C,H,W,OH,OW", [(2, 64, 16, 16, 32, 32), (1, 8, 4, 4, 128, 128), (2, 16, 8, 12, 16, 24),
                                           (1, 8, 5, 7, 13, 9), (2, 32, 32, 32, 16, 16)])
def test_bilinear_resize_align_corners_false(N, C, H, W, OH, OW):
    """_upsample_like (u2net.py:19-23): F.interpolate(size=..., mode="bilinear", align_corners=False), into a channel slice."""
    x = bf16r(torch.randn(N, C, H, W, device=DEV))
    ctx = new_ctx()
    xa = make_act(ctx, x)
    wide = ctx.new_buf(N, OH, OW, 2 * C)
    wide.data.fill_(3.0)
    out = E.bilinear_resize(ctx, xa, OH, OW, out=E.Act(wide, 0, C), align_corners=False)
    xr = x.clone().requires_grad_(True)
    ref = F.interpolate(xr, size=(OH, OW), mode="bilinear", align_corners=False)
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), ref.detach()) < 1e-2
    assert bool((wide.data[..., C:] == 3.0).all())
    dy = bf16r(torch.randn_like(ref))
    (ref * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2


@pytest.mark.parametrize("N,C,H,W,pool", [(2, 64, 16, 24, True), (1, 16, 8, 8, False), (2, 8, 6, 10, True)])
def test_residual_add_with_pooling(N, C, H, W, pool):
    a = bf16r(torch.randn(N, C, H, W, device=DEV))
    b = bf16r(torch.randn(N, C, H, W, device=DEV))
    ctx = new_ctx()
    aa, ba = make_act(ctx, a), make_act(ctx, b)
    pooled = ctx.new_act(N, H // 2, W // 2, C) if pool else None
    out = E.residual_add(ctx, aa, ba, pooled=pooled)
    ar, br = a.clone().requires_grad_(True), b.clone().requires_grad_(True)
    s = ar + br
    sq = s + (bf16r(s.detach()) - s.detach())
    torch.cuda.synchronize()
    assert torch.equal(out.to_nchw(), sq.detach())
    dy = bf16r(torch.randn_like(s))
    loss = (sq * dy).sum()
    set_grad(out, dy)
    if pool:
        p = F.max_pool2d(sq, 2, 2)
        assert torch.equal(pooled.to_nchw(), p.detach())
        dp = bf16r(torch.randn_like(p))
        loss = loss + (p * dp).sum()
        set_grad(pooled, dp)
    loss.backward()
    prev = bf16r(torch.randn_like(b))
    set_grad(ba, prev)                         # b already has a gradient from another consumer (hxin -> rebnconv1)
    run_tape(ctx)
    assert relmax(get_grad(aa), ar.grad) < 1e-2
    assert relmax(get_grad(ba), br.grad + prev) < 1e-2


@pytest.mark.parametrize("K", [1, 3])
def test_side_convolutions_and_fused_head(K):
    """Two side maps at different resolutions -> 8-channel slots -> one head pass: the three logit maps [main, side1,
    side2] (K classes each) and every gradient against the reference arithmetic (u2net.py:273-285) in fp32."""
    N, H, W = 2, 32, 32
    s1 = nn.Conv2d(64, K, 3, padding=1).to(DEV)
    s2 = nn.Conv2d(32, K, 3, padding=1).to(DEV)
    outconv = nn.Conv2d(2 * K, K, 1).to(DEV)
    with torch.no_grad():
        for m in (s1, s2):
            m.weight.copy_(bf16r(m.weight))
    x1 = bf16r(torch.randn(N, 64, H, W, device=DEV))
    x2 = bf16r(torch.randn(N, 32, H // 4, W // 4, device=DEV))
    params = [s1.weight, s1.bias, s2.weight, s2.bias, outconv.weight, outconv.bias]
    ctx = new_ctx(params=params)
    a1, a2 = make_act(ctx, x1), make_act(ctx, x2)
    fuse = ctx.new_buf(N, H, W, 16)
    E.conv_padded_out(ctx, [a1], s1, E.Act(fuse, 0, 8))
    low = E.conv_padded_out(ctx, [a2], s2, ctx.new_act(N, H // 4, W // 4, 8))
    E.bilinear_resize(ctx, low, H, W, out=E.Act(fuse, 8, 8), align_corners=False)
    logits = E.fused_side_head(ctx, E.Act(fuse), outconv, K, 2)
    xr1, xr2 = x1.clone().requires_grad_(True), x2.clone().requires_grad_(True)
    ps = [p.detach().clone().requires_grad_(True) for p in params]
    d1 = F.conv2d(xr1, ps[0], ps[1], padding=1)
    d2 = F.interpolate(F.conv2d(xr2, ps[2], ps[3], padding=1), size=(H, W), mode="bilinear", align_corners=False)
    d0 = F.conv2d(torch.cat((d1, d2), 1), ps[4], ps[5])
    ref = torch.cat((d0, d1, d2), 1)
    torch.cuda.synchronize()
    assert logits.shape == ref.shape == (N, 3 * K, H, W)
    assert relmax(logits, ref.detach()) < 1e-2
    assert float(fuse.data[..., K:8].abs().max()) == 0 and float(fuse.data[..., 8 + K:].abs().max()) == 0
    g = torch.randn_like(ref)
    (ref * g).sum().backward()
    ctx.out_grads = [g]
    run_tape(ctx)
    ctx.join_side()
    torch.cuda.synchronize()
    assert rel_l2(get_grad(a1), xr1.grad) < 1.5e-2 and rel_l2(get_grad(a2), xr2.grad) < 1.5e-2
    for p, r in zip(params, ps):
        assert rel_l2(ctx.pgrads.view(p), r.grad) < 1.5e-2


# ---------------------------------------------------------------------------------------------------- whole models
def _as_list(o):
    return list(o.values()) if isinstance(o, dict) else [o]


def _step(model, x, mask, loss_fn):
    model.cuda().train()
    out = model(x.cuda())
    assert list(out) == list(O.U2NET_KEYS)
    outs = _as_list(out)
    sum(loss_fn(o, mask.cuda()) for o in outs).backward()          # training_loop.py:61-64: weight 1.0 each
    torch.cuda.synchronize()
    return [o.detach().cpu() for o in outs]


# fixed gradient bars (all conv weights concatenated, worst tensor), see tests/test_models_gpu.py
GRAD_BARS = {"u2net": (0.3, 0.6), "u2netp": (0.3, 0.6)}


@pytest.mark.parametrize("name", ["u2net", "u2netp"])
def test_train_step_against_the_reference_fixture(name):
    gold = load_golden(name)
    cond = gold["cond"]
    model, sd = conditioned_model(name)
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    outs = _step(model, x, mask, O.gaussian_projection_loss)
    assert len(outs) == len(cond["logits_train"]) == 7
    errs = [rel_err(o, g) for o, g in zip(outs, cond["logits_train"])]
    agree = [float((O.threshold_mask(o) == O.threshold_mask(g)).float().mean()) for o, g in zip(outs, cond["logits_train"])]
    bce = float(sum(O.bce_with_logits(o, mask) for o in outs))
    named = dict(model.named_parameters())
    grads = {k: p.grad.detach().cpu() for k, p in named.items()}
    _, _, ref_grads = O.train_step(name, sd, x, mask, loss_fn=O.gaussian_projection_loss)
    e_all, per = grad_errors(grads, ref_grads)
    parity_report("u2net_train/" + name, logits=errs, agree=agree, grad_all=e_all, grad_worst=per[:4], bce=[bce, cond["bce"]])
    for o, g in zip(outs, cond["logits_train"]):
        assert o.shape == g.shape and o.dtype == torch.float32
    assert max(errs) < LOGIT_TOL, errs
    assert min(agree) >= MASK_AGREE, agree
    assert abs(bce - cond["bce"]) < 2e-3 * len(outs)
    got = model.state_dict()
    for k, v in cond["bn_after"].items():
        if "running_" in k:
            assert rel_err(got[k].cpu(), v) < 2e-2, k
        else:
            assert int(got[k]) == int(v), k
    for k, p in named.items():
        assert p.grad is not None and p.grad.shape == p.shape and bool(torch.isfinite(p.grad).all()), k
    for k in ("outconv.weight", "outconv.bias", "side1.weight", "side6.weight"):
        assert float((grads[k] - ref_grads[k]).norm() / ref_grads[k].norm().clamp_min(1e-20)) < 3e-2, k
    all_bar, worst_bar = GRAD_BARS[name]
    assert e_all < all_bar and per[0][0] < worst_bar, (e_all, per[:3])


@pytest.mark.parametrize("name", ["u2net", "u2netp"])
def test_eval_forward_and_mask_agreement(name):
    gold = load_golden(name)
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    for state in ("cond", "default"):
        if state == "cond":
            model, _ = conditioned_model(name)
            _step(model, x, mask, O.gaussian_projection_loss)
            want = gold["cond"]["logits_eval"]
        else:
            model = seeded_model(name)
            _step(model, x, mask, O.bce_with_logits)
            want = gold["logits_eval"]                      # main + side1 only (fixture size)
        sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
        model.eval()
        with torch.no_grad():
            outs = [o.cpu() for o in _as_list(model(x.cuda()))]
            refs = O.FORWARDS[name](sd, x, False)
        for i, (o, r) in enumerate(zip(outs, refs)):
            agree = float((O.threshold_mask(o) == O.threshold_mask(r)).float().mean())
            parity_report(f"u2net_eval/{name}/{state}/{i}", logits=rel_err(o, r), agree=agree,
                          positive=float(O.threshold_mask(r).float().mean()))
            assert rel_err(o, r) < LOGIT_TOL, (state, i)
            assert agree >= MASK_AGREE, (state, i, agree)
        for o, g in zip(outs, want):
            assert rel_err(o, g) < 4e-2                      # the fixture used the reference's own fp32 statistics


def test_u2net_rejects_sizes_the_nesting_cannot_halve():
    model = seeded_model("u2netp").cuda()
    with pytest.raises(RuntimeError, match="multiples of 32"):
        model(torch.zeros(1, 3, 48, 64, device="cuda"))
    with pytest.raises(ValueError, match="bf16"):
        model.set_precision("fp32")


def test_u2netp_non_square_eval_and_graph_replay():
    """Eval forward at 64 x 96 (innermost maps 2 x 3), batch 3, against the oracle; the same forward replayed from one
    CUDA graph returns the same dict bit for bit."""
    from unet_zoo_b200.train_step import GraphedInference
    model = seeded_model("u2netp")
    sd = O.randomize_bn_affine({k: v.clone() for k, v in model.state_dict().items()})
    with torch.no_grad():
        for k, v in sd.items():
            if k.endswith("running_mean"):
                v.normal_(0.0, 0.2, generator=torch.Generator().manual_seed(len(k)))
            if k.endswith("running_var"):
                v.uniform_(0.5, 1.5, generator=torch.Generator().manual_seed(len(k) + 1))
    model.load_state_dict(sd)
    model.cuda().eval()
    x, _ = O.synthetic_batch(3, 3, 64, 96)
    with torch.no_grad():
        out = model(x.cuda())
        refs = O.u2netp_forward({k: v.clone() for k, v in sd.items()}, x, False)
    assert list(out) == list(O.U2NET_KEYS)
    for (k, o), r in zip(out.items(), refs):
        assert o.shape == r.shape == (3, 1, 64, 96), k
        assert rel_err(o.cpu(), r) < LOGIT_TOL, k
    g = GraphedInference(model, x.cuda())
    again = g(x.cuda())
    for k in out:
        assert torch.equal(again[k], out[k]), k


def test_u2netp_two_classes_against_the_oracle():
    """out_ch = 2 (the factory's num_classes): seven maps of two channels each, eval mode and one training step's gradients
    of the output stage against the oracle."""
    model = seeded_model("u2netp", num_classes=2)
    sd = O.randomize_bn_affine({k: v.clone() for k, v in model.state_dict().items()})
    model.load_state_dict(sd)
    x, _ = O.synthetic_batch(2, 3, 64, 64)
    model.cuda().train()
    out = model(x.cuda())
    assert all(o.shape == (2, 2, 64, 64) for o in out.values())
    sum(O.gaussian_projection_loss(o, None) for o in out.values()).backward()
    refs, _, ref_grads = O.train_step("u2netp", {k: v.clone() for k, v in sd.items()}, x, None,
                                      loss_fn=O.gaussian_projection_loss)
    for (k, o), r in zip(out.items(), refs):
        assert rel_err(o.detach().cpu(), r.detach()) < LOGIT_TOL, k
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters()}
    for k in ("outconv.weight", "outconv.bias", "side1.weight", "side1.bias", "side6.weight"):
        assert float((grads[k] - ref_grads[k]).norm() / ref_grads[k].norm().clamp_min(1e-20)) < 5e-2, k
