"""`multiresunet` on the GPU: the operators it adds against plain PyTorch fp32 math, and the whole
model (through create_model -> forward -> loss.backward) against fixtures of the unmodified reference
(tests/golden/multiresunet.pt, oracle/make_golden.py) and against the oracle on other shapes."""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from conftest import grad_errors, load_golden, parity_report, rel_err, seeded_model
from gpu_util import bf16r, get_grad, make_act, new_ctx, relmax, rel_l2, run_tape, set_grad
from oracle import reference_port as O
from unet_zoo_b200 import engine as E

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
# Fixed bars of this file.  MultiResUnet has no BatchNorm affine parameters, so the well-conditioned parity state of the
# other models (randomised gamma / beta) does not exist for it: every train-mode BatchNorm output is
# zero-mean / unit-variance, the ReLU behind it sits on the densest part of the distribution, and a bf16 rounding of the
# pre-activation flips ~0.4 * 2^-9 of the masks per layer.  torch's own bf16 autocast of the reference is the yardstick
# (tools/parity_floor.py multiresunet 2 64 gauss): train logits 6.0e-2 from fp32, masks 97.9 %, weight gradients 57 %
# median / 91 % worst; eval logits 1.0e-2, masks 100 %.  Larger inputs do not help (7.2e-2 at 2 x 128^2, 7.4e-2 at
# 4 x 128^2).  Per block the same yardstick gives 0.8e-2 on the output, 8e-2 (Multiresblock) / 4-13e-2 (Respath) on the
# gradients.  This path measures the same numbers (profiles/r02_cm_multiresunet_parity.jsonl): 6.3e-2 / 98.0 % / 64 %
# whole-model train, 0.7e-2 / 100 % eval, 8-9e-2 per block.
# Parity is therefore pinned in three ways:
#  * eval-mode forward of the whole model against the plain fp32 oracle: EVAL_TOL;
#  * every kind of block, forward AND backward, with the model's own parameters and padded operands, against the oracle
#    run with the executor's storage roundings (oracle Runner.round_storage = bf16: weights, conv inputs / outputs,
#    BatchNorm inputs / outputs rounded at the same points, everything else fp32).  Both sides' ReLU masks then see the
#    same values; what remains is accumulation order: outputs within two bf16 ulps of the largest value (measured 0 to
#    1.3e-2; Multiresblock 1 and 9 are bit-identical), gradients within BLOCK_GRAD (measured 0.2 - 3.5e-2);
#  * the whole training step against that rounding-matched oracle (MATCHED_*: measured 2.7e-2, 99.3 %, gradients 41 % -
#    one-ulp differences still flip masks and the network amplifies them), and against the reference fixture at the
#    loose LOGIT_TOL / MASK_AGREE sanity bars.
LOGIT_TOL = 2e-1
MASK_AGREE = 0.96
EVAL_TOL = 3e-2
MATCHED_TOL = 5e-2
MATCHED_AGREE = 0.985
MATCHED_GRAD = 0.6
BLOCK_OUT = 2e-2
BLOCK_GRAD = 5e-2


@pytest.fixture(autouse=True)
def _exact_fp32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(5)


# ---------------------------------------------------------------------------------------------------- operators
@pytest.mark.parametrize("N,C,H,W,pool", [(2, 64, 16, 24, True), (1, 24, 8, 8, False), (2, 120, 6, 10, True)])
def test_residual_add_relu(N, C, H, W, pool):
    """x = F.relu(x + temp) (multiresunet.py:82-83) in one pass; the gradient passes where the stored sum is positive."""
    a = bf16r(torch.randn(N, C, H, W, device=DEV))
    b = bf16r(torch.randn(N, C, H, W, device=DEV))
    ctx = new_ctx()
    aa, ba = make_act(ctx, a), make_act(ctx, b)
    pooled = ctx.new_act(N, H // 2, W // 2, C) if pool else None
    out = E.residual_add(ctx, aa, ba, pooled=pooled, relu=True)
    ar, br = a.clone().requires_grad_(True), b.clone().requires_grad_(True)
    s = ar + br
    sq = F.relu(s + (bf16r(s.detach()) - s.detach()))
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
    run_tape(ctx)
    assert relmax(get_grad(aa), ar.grad) < 1e-2
    assert relmax(get_grad(ba), br.grad) < 1e-2


@pytest.mark.parametrize("N,C,H,W,pool", [(2, 64, 16, 16, True), (2, 120, 8, 12, False), (1, 224, 8, 8, True)])
def test_batchnorm_without_affine_into_a_slice_with_pooling(N, C, H, W, pool):
    """BatchNorm2d(affine=False) of an existing activation (multiresunet.py:79,84), written into a channel slice of a
    wider buffer with MaxPool2d(2, 2) of the result as a second output."""
    bn = nn.BatchNorm2d(C, affine=False).to(DEV)
    x = bf16r(torch.randn(N, C, H, W, device=DEV) * 1.5 + 0.3)
    ctx = new_ctx()
    xa = make_act(ctx, x)
    out = E.Act(ctx.new_buf(N, H, W, C + 16), 16, C)
    pooled = ctx.new_act(N, H // 2, W // 2, C) if pool else None
    rm, rv = bn.running_mean.clone(), bn.running_var.clone()
    E.bn_act(ctx, xa, bn, relu=False, out=out, pooled=pooled)
    xr = x.clone().requires_grad_(True)
    y = F.batch_norm(xr, rm, rv, None, None, training=True, momentum=0.1, eps=1e-5)
    yq = y + (bf16r(y.detach()) - y.detach())
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), yq.detach()) < 1e-2
    assert relmax(bn.running_mean, rm) < 2e-3 and relmax(bn.running_var, rv) < 2e-3 and int(bn.num_batches_tracked) == 1
    dy = bf16r(torch.randn_like(y) * 0.1)
    loss = (yq * dy).sum()
    set_grad(out, dy)
    if pool:
        p = F.max_pool2d(yq, 2, 2)
        assert relmax(pooled.to_nchw(), p.detach()) < 1e-2
        dp = bf16r(torch.randn_like(p) * 0.1)
        loss = loss + (p * dp).sum()
        set_grad(pooled, dp)
    loss.backward()
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad) < 1.5e-2


@pytest.mark.parametrize("N,C,H,W,relu_a", [(2, 64, 16, 16, False), (2, 32, 16, 24, True), (1, 120, 8, 8, False), (3, 256, 8, 8, True)])
def test_two_batchnorms_add_relu_in_one_pass(N, C, H, W, relu_a):
    """relu([relu](BN(a)) + BN(b)) from the raw tensors (uz_dual_bn_add_relu_fwd/_bwd): a = a convolution output with a ReLU
    behind its BatchNorm (Respath) or an existing activation (the concatenation of a Multiresblock), b = the 1x1 shortcut."""
    from unet_zoo_b200 import _lib as L
    cin = 24
    convB, bnB = nn.Conv2d(cin, C, 1).to(DEV), nn.BatchNorm2d(C, affine=False).to(DEV)
    convA, bnA = nn.Conv2d(cin, C, 3, padding=1).to(DEV), nn.BatchNorm2d(C, affine=False).to(DEV)
    with torch.no_grad():
        convA.weight.copy_(bf16r(convA.weight))
        convB.weight.copy_(bf16r(convB.weight))
    x = bf16r(torch.randn(N, cin, H, W, device=DEV))
    a = bf16r(torch.randn(N, C, H, W, device=DEV) * 2.0 + 0.5)
    params = [convA.weight, convA.bias, convB.weight, convB.bias]
    ctx = new_ctx(params=params)
    xa, aa = make_act(ctx, x), make_act(ctx, a)
    A = E.bn_branch_conv(ctx, [xa], convA, bnA, relu=True) if relu_a else E.bn_branch_act(ctx, aa, bnA)
    B = E.bn_branch_conv(ctx, [xa], convB, bnB, relu=False, mode=L.CONV1X1)
    stats = ctx.f32(ctx.reduce_rows, 2, C)
    out = E.dual_bn_add_relu(ctx, A, B, stats=stats)
    ste = lambda t: t + (bf16r(t.detach()) - t.detach())      # noqa: E731
    xr, ar = x.clone().requires_grad_(True), a.clone().requires_grad_(True)
    ps = [p.detach().clone().requires_grad_(True) for p in params]
    bn = lambda t: F.batch_norm(t, None, None, None, None, training=True, eps=1e-5)      # noqa: E731
    ya = F.relu(bn(ste(F.conv2d(xr, ps[0], ps[1], padding=1)))) if relu_a else bn(ar)
    y = F.relu(ya + bn(ste(F.conv2d(xr, ps[2], ps[3]))))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), y.detach()) < 1e-2
    stored = out.to_nchw().double()      # the by-product statistics describe the tensor as stored
    assert relmax(stats[:, 0].double().sum(0), stored.sum((0, 2, 3))) < 1e-5
    assert relmax(stats[:, 1].double().sum(0), (stored * stored).sum((0, 2, 3))) < 1e-5
    dy = bf16r(torch.randn_like(y) * 0.1)
    (y * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    ctx.join_side()
    torch.cuda.synchronize()
    assert rel_l2(get_grad(xa), xr.grad) < 1.5e-2
    assert rel_l2(ctx.pgrads.view(convB.weight), ps[2].grad) < 1e-2
    if relu_a:
        assert rel_l2(ctx.pgrads.view(convA.weight), ps[0].grad) < 1e-2
        assert int(bnA.num_batches_tracked) == 1
    else:
        assert rel_l2(get_grad(aa), ar.grad) < 1.5e-2
    assert float(ctx.pgrads.view(convB.bias).abs().max()) < 2e-2 * float(dy.abs().sum() / C) + 1e-3


@pytest.mark.parametrize("C,pool", [(64, False), (64, True), (120, False)])
def test_whole_tail_with_the_following_batchnorm(C, pool):
    """BN(relu(relu(BN(conv3x3(x))) + BN(conv1x1(x)))) [+ MaxPool2d] as one operator (engine.dual_bn_add_relu_bn): without a
    pooled gradient the outer BatchNorm's backward is applied on the fly inside the fused backward (uz_bn_bwd_coef +
    bn_outer / coef_outer of uz_dual_bn_add_relu_bwd); with one, it runs on its own first."""
    from unet_zoo_b200 import _lib as L
    N, H, W, cin = 2, 16, 24, 24
    convA, bnA = nn.Conv2d(cin, C, 3, padding=1).to(DEV), nn.BatchNorm2d(C, affine=False).to(DEV)
    convB, bnB = nn.Conv2d(cin, C, 1).to(DEV), nn.BatchNorm2d(C, affine=False).to(DEV)
    bnO = nn.BatchNorm2d(C, affine=False).to(DEV)
    with torch.no_grad():
        convA.weight.copy_(bf16r(convA.weight))
        convB.weight.copy_(bf16r(convB.weight))
    x = bf16r(torch.randn(N, cin, H, W, device=DEV))
    params = [convA.weight, convA.bias, convB.weight, convB.bias]
    ctx = new_ctx(params=params)
    xa = make_act(ctx, x)
    A = E.bn_branch_conv(ctx, [xa], convA, bnA, relu=True)
    B = E.bn_branch_conv(ctx, [xa], convB, bnB, relu=False, mode=L.CONV1X1)
    out = E.Act(ctx.new_buf(N, H, W, C + 8), 8, C)
    pooled = ctx.new_act(N, H // 2, W // 2, C) if pool else None
    E.dual_bn_add_relu_bn(ctx, A, B, bnO, out=out, pooled=pooled)
    ste = lambda t: t + (bf16r(t.detach()) - t.detach())      # noqa: E731
    bn = lambda t: F.batch_norm(t, None, None, None, None, training=True, eps=1e-5)      # noqa: E731
    xr = x.clone().requires_grad_(True)
    ps = [p.detach().clone().requires_grad_(True) for p in params]
    ya = F.relu(bn(ste(F.conv2d(xr, ps[0], ps[1], padding=1))))
    s = ste(F.relu(ya + bn(ste(F.conv2d(xr, ps[2], ps[3])))))
    o = ste(bn(s))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), o.detach()) < 1e-2
    assert int(bnO.num_batches_tracked) == 1 and relmax(bnO.running_var, 0.9 + 0.1 * s.detach().var((0, 2, 3))) < 1e-2
    do = bf16r(torch.randn_like(o) * 0.1)
    loss = (o * do).sum()
    set_grad(out, do)
    if pool:
        p = F.max_pool2d(o, 2, 2)
        assert relmax(pooled.to_nchw(), p.detach()) < 1e-2
        dp = bf16r(torch.randn_like(p) * 0.1)
        loss = loss + (p * dp).sum()
        set_grad(pooled, dp)
    loss.backward()
    run_tape(ctx)
    ctx.join_side()
    torch.cuda.synchronize()
    assert rel_l2(get_grad(xa), xr.grad) < 2e-2
    assert rel_l2(ctx.pgrads.view(convA.weight), ps[0].grad) < 1.5e-2
    assert rel_l2(ctx.pgrads.view(convB.weight), ps[2].grad) < 1.5e-2


@pytest.mark.parametrize("K", [1, 3, 8])
def test_logits_extraction_and_its_gradient(K):
    N, H, W = 2, 16, 24
    y = bf16r(torch.randn(N, 8, H, W, device=DEV))
    ctx = new_ctx()
    ya = make_act(ctx, y)
    logits = E.nhwc_logits(ctx, ya, K)
    torch.cuda.synchronize()
    assert logits.dtype == torch.float32 and torch.equal(logits, y[:, :K])
    g = torch.randn(N, K, H, W, device=DEV)
    ctx.out_grads = [g]
    run_tape(ctx)
    got = get_grad(ya)
    assert torch.equal(got[:, :K], bf16r(g)) and float(got[:, K:].abs().sum()) == 0.0


# ---------------------------------------------------------------------------------------------------- blocks
# Whole-model train-mode outputs of this network are ill-conditioned (see LOGIT_TOL above), so forward AND backward
# parity is pinned block by block: every kind of block of the model, with the model's own parameters and padded
# operands, against the oracle's restatement of that block on the same bf16-rounded input and output gradient.
def _pad(x, lay):
    out = torch.zeros(x.shape[0], lay.padded, *x.shape[2:], device=x.device)
    out[:, lay.index().to(x.device)] = x
    return out


def _real(x, lay):
    return x[:, lay.index().to(x.device)]


def _pad_is_zero(x, lay):
    mask = torch.ones(lay.padded, dtype=torch.bool, device=x.device)
    mask[lay.index().to(x.device)] = False
    return float(x[:, mask].abs().sum()) == 0.0 if bool(mask.any()) else True


def _block_harness(model):
    model.cuda().train()
    model._before_forward(DEV)                     # builds / refreshes the padded operands
    params = list(model.parameters())
    ctx = new_ctx(params=params)
    model._uz_shadow.begin_backward(ctx.pgrads)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    leaves = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running_" not in k}
    return ctx, params, sd, leaves


def _finish(model, ctx, params):
    run_tape(ctx)
    ctx.join_side()
    model._uz_shadow.end_backward(ctx.pgrads, params)
    model._uz_shadow.write_back_stats()
    torch.cuda.synchronize()


def _check_param_grads(model, ctx, leaves, prefix, tag, tol):
    worst = []
    for k, p in model.named_parameters():
        if not k.startswith(prefix + ".") or p.dim() != 4:
            continue
        worst.append((rel_l2(ctx.pgrads.view(p), leaves[k].grad), k))
    assert worst
    worst.sort(reverse=True)
    parity_report("multiresunet_block/" + tag, grad_worst=worst[:3])
    assert worst[0][0] < tol, worst[:3]


def _check_stats(model, sd, prefix, twice=()):
    got = model.state_dict()
    for k, v in sd.items():
        if not k.startswith(prefix + "."):
            continue
        if "running_" in k:
            assert relmax(got[k], v) < 2e-2, k
        elif "num_batches" in k:
            assert int(got[k]) == int(v) == (2 if k[:-len(".num_batches_tracked")] in twice else 1), k


@pytest.mark.parametrize("name,N,H,W", [("multiresblock1", 2, 32, 32), ("multiresblock2", 2, 32, 16), ("multiresblock3", 2, 16, 16),
                                        ("multiresblock5", 1, 16, 16), ("multiresblock6", 1, 16, 16), ("multiresblock9", 2, 32, 32)])
def test_multiresblock_against_the_oracle(name, N, H, W, monkeypatch):
    monkeypatch.setattr(O.Runner, "round_storage", torch.bfloat16)      # same storage roundings as the executor
    model = seeded_model("multiresunet")
    ctx, params, sd, leaves = _block_harness(model)
    blk = getattr(model, name)
    specs = {id(s[1]): s for s in model._specs()}
    in_lay, out_lay = specs[id(blk.conv2d_bn_3x3.conv1)][2], specs[id(blk.conv2d_bn_1x1.conv1)][3]
    first = name == "multiresblock1"
    x = bf16r(torch.randn(N, in_lay.real, H, W, device=DEV))
    xa = make_act(ctx, _pad(x, in_lay))
    pooled = ctx.new_act(N, H // 2, W // 2, out_lay.padded)
    out = model._block(ctx, blk, [xa], pooled=pooled, need_dx=not first)
    xr = x.clone().requires_grad_(True)
    y = O._multiresblock(O.Runner(sd, True), xr, name)
    p = F.max_pool2d(y, 2, 2)
    torch.cuda.synchronize()
    e_out, e_pool = relmax(_real(out.to_nchw(), out_lay), y.detach()), relmax(_real(pooled.to_nchw(), out_lay), p.detach())
    parity_report("multiresunet_block/" + name, out=e_out, pooled=e_pool)
    assert e_out < BLOCK_OUT and e_pool < BLOCK_OUT
    assert _pad_is_zero(out.to_nchw(), out_lay) and _pad_is_zero(pooled.to_nchw(), out_lay)
    dy, dp = bf16r(torch.randn_like(y) * 0.1), bf16r(torch.randn_like(p) * 0.1)
    ((y * dy).sum() + (p * dp).sum()).backward()
    set_grad(out, _pad(dy, out_lay))
    set_grad(pooled, _pad(dp, out_lay))
    _finish(model, ctx, params)
    if not first:
        g = get_grad(xa)
        e_dx = rel_l2(_real(g, in_lay), xr.grad)
        parity_report("multiresunet_block/" + name, dx=e_dx)
        assert e_dx < BLOCK_GRAD and _pad_is_zero(g, in_lay)
    _check_param_grads(model, ctx, leaves, name, name, BLOCK_GRAD)
    _check_stats(model, sd, name, twice=(name + ".batch_norm1",))


@pytest.mark.parametrize("name,N,H,W", [("respath1", 2, 32, 32), ("respath3", 2, 16, 16), ("respath4", 1, 16, 16)])
def test_respath_against_the_oracle(name, N, H, W, monkeypatch):
    monkeypatch.setattr(O.Runner, "round_storage", torch.bfloat16)
    model = seeded_model("multiresunet")
    ctx, params, sd, leaves = _block_harness(model)
    rp = getattr(model, name)
    in_lay = {id(s[1]): s for s in model._specs()}[id(rp.conv2d_bn_3x3_initial.conv1)][2]
    x = bf16r(torch.randn(N, in_lay.real, H, W, device=DEV))
    xa = make_act(ctx, _pad(x, in_lay))
    cat = ctx.new_buf(N, H, W, 2 * rp.filters)
    out = model._respath(ctx, rp, xa, E.Act(cat, rp.filters, rp.filters))      # the skip half of the decoder's buffer
    xr = x.clone().requires_grad_(True)
    y = O._respath(O.Runner(sd, True), xr, name, rp.respath_length)
    torch.cuda.synchronize()
    e_out = relmax(out.to_nchw(), y.detach())
    parity_report("multiresunet_block/" + name, out=e_out)
    assert e_out < BLOCK_OUT
    dy = bf16r(torch.randn_like(y) * 0.1)
    (y * dy).sum().backward()
    set_grad(out, dy)
    _finish(model, ctx, params)
    g = get_grad(xa)
    e_dx = rel_l2(_real(g, in_lay), xr.grad)
    parity_report("multiresunet_block/" + name, dx=e_dx)
    assert e_dx < BLOCK_GRAD and _pad_is_zero(g, in_lay)
    _check_param_grads(model, ctx, leaves, name, name, BLOCK_GRAD)
    _check_stats(model, sd, name)


@pytest.mark.parametrize("K", [1, 3])
def test_normalised_output_stage_against_the_oracle(K):
    """conv_final = Conv2d(1x1) -> BatchNorm2d(affine=False) IS the model output (multiresunet.py:200-201,238-240)."""
    model = seeded_model("multiresunet", num_classes=K)
    ctx, params, sd, leaves = _block_harness(model)
    in_lay = {id(s[1]): s for s in model._specs()}[id(model.conv_final.conv1)][2]
    N, H, W = 2, 32, 48
    x = bf16r(torch.randn(N, in_lay.real, H, W, device=DEV))
    xa = make_act(ctx, _pad(x, in_lay))
    logits = E.nhwc_logits(ctx, model._unit(ctx, [xa], model.conv_final), K)
    xr = x.clone().requires_grad_(True)
    y = O._conv2d_bn(O.Runner(sd, True), xr, "conv_final", False, 0)
    torch.cuda.synchronize()
    assert logits.shape == y.shape and relmax(logits, y.detach()) < 1e-2
    g = torch.randn_like(y) * 0.1
    (y * g).sum().backward()
    ctx.out_grads = [g]
    _finish(model, ctx, params)
    assert rel_l2(_real(get_grad(xa), in_lay), xr.grad) < 2e-2
    _check_param_grads(model, ctx, leaves, "conv_final", f"conv_final/{K}", 2e-2)
    _check_stats(model, sd, "conv_final")


# ---------------------------------------------------------------------------------------------------- whole model
def _worst_stats_error(got, ref):
    """Largest deviation of a running statistic after ONE step from the initial (0, 1) state, in units of the statistic's
    natural scale: running_var by its largest entry; running_mean (= 0.1 * batch mean, often ~0 for normalised inputs) by
    the larger of its largest entry and 0.1 * the largest batch standard deviation (recovered from running_var)."""
    worst = (0.0, "")
    for k, v in ref.items():
        if "running_" not in k:
            continue
        scale = float(v.abs().max())
        if k.endswith("running_mean"):
            var_b = ((ref[k[:-len("mean")] + "var"].double() - 0.9) / 0.1).clamp_min(0.0)
            scale = max(scale, 0.1 * float(var_b.max().sqrt()))
        e = float((got[k].double().cpu() - v.double()).abs().max()) / max(scale, 1e-12)
        worst = max(worst, (e, k))
    return worst


def _train_step(model, x, mask, loss_fn):
    model.cuda().train()
    out = model(x.cuda())
    loss_fn(out, mask.cuda() if mask is not None else None).backward()
    torch.cuda.synchronize()
    return out.detach().cpu()


def test_train_step_against_the_reference_fixture():
    gold = load_golden("multiresunet")
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    for tag, loss_fn, want in (("bce", O.bce_with_logits, gold), ("gauss", O.gaussian_projection_loss, gold["cond"])):
        model = seeded_model("multiresunet")
        sd = copy.deepcopy(model.state_dict())
        out = _train_step(model, x, mask, loss_fn)
        ref = want["logits_train"][0]
        assert out.shape == ref.shape and out.dtype == torch.float32
        err = rel_err(out, ref)
        agree = float((O.threshold_mask(out) == O.threshold_mask(ref)).float().mean())
        named = dict(model.named_parameters())
        grads = {k: p.grad.detach().cpu() for k, p in named.items()}
        _, _, ref_grads = O.train_step("multiresunet", sd, x, mask, loss_fn=loss_fn)
        e_all, per = grad_errors(grads, ref_grads)
        parity_report("multiresunet_train/" + tag, logits=err, agree=agree, grad_all=e_all, grad_worst=per[:4])
        assert err < LOGIT_TOL, (tag, err)
        assert agree >= MASK_AGREE, (tag, agree)
        got = model.state_dict()
        assert list(got) == list(gold["fingerprint"])
        for k, v in want["bn_after"].items():
            assert got[k].shape == v.shape, k
            if "num_batches" in k:
                assert int(got[k]) == int(v), k           # batch_norm1 of every Multiresblock counts 2 per forward
        ws = _worst_stats_error(got, want["bn_after"])
        parity_report("multiresunet_train/" + tag, stats_worst=ws)
        assert ws[0] < 1e-1, ws
        for k, p in named.items():
            assert p.grad is not None and p.grad.shape == p.shape and bool(torch.isfinite(p.grad).all()), k
        assert e_all < 0.9 and per[0][0] < 1.2, (tag, e_all, per[:3])     # see LOGIT_TOL: the function's own bf16 noise


def test_train_step_against_the_oracle_with_matched_storage_rounding(monkeypatch):
    """The same training step against the oracle run with the executor's storage roundings (oracle Runner.round_storage):
    the ReLU masks of both sides then see the same bf16 values, and what is left is accumulation order."""
    gold = load_golden("multiresunet")
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    model = seeded_model("multiresunet")
    sd = copy.deepcopy(model.state_dict())
    out = _train_step(model, x, mask, O.gaussian_projection_loss)
    monkeypatch.setattr(O.Runner, "round_storage", torch.bfloat16)
    ref, _, ref_grads = O.train_step("multiresunet", sd, x, mask, loss_fn=O.gaussian_projection_loss)
    ref = ref.detach()
    err = rel_err(out, ref)
    agree = float((O.threshold_mask(out) == O.threshold_mask(ref)).float().mean())
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters()}
    e_all, per = grad_errors(grads, ref_grads)
    parity_report("multiresunet_train/matched", logits=err, agree=agree, grad_all=e_all, grad_worst=per[:4])
    assert err < MATCHED_TOL and agree >= MATCHED_AGREE
    assert e_all < MATCHED_GRAD and per[0][0] < 1.5 * MATCHED_GRAD, (e_all, per[:3])
    ws = _worst_stats_error(model.state_dict(), sd)
    parity_report("multiresunet_train/matched", stats_worst=ws)
    assert ws[0] < 5e-2, ws


def test_eval_forward_against_the_oracle_and_the_fixture():
    gold = load_golden("multiresunet")
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    model = seeded_model("multiresunet")
    _train_step(model, x, mask, O.bce_with_logits)              # running statistics of one training step
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        out = model(x.cuda()).cpu()
        again = model(x.cuda()).cpu()                           # second call: memoised folded weights
        ref = O.multiresunet_forward(sd, x, False)
    assert torch.equal(out, again)
    agree = float((O.threshold_mask(out) == O.threshold_mask(ref)).float().mean())
    parity_report("multiresunet_eval", logits=rel_err(out, ref), agree=agree, fixture=rel_err(out, gold["logits_eval"][0]))
    assert rel_err(out, ref) < EVAL_TOL
    assert agree >= 0.995
    assert rel_err(out, gold["logits_eval"][0]) < LOGIT_TOL          # the fixture used the reference's own fp32 statistics


def test_rejects_what_the_reference_rejects():
    model = seeded_model("multiresunet").cuda()
    with pytest.raises(RuntimeError, match="multiples of 16"):
        model(torch.zeros(1, 3, 40, 64, device="cuda"))
    with pytest.raises(ValueError, match="bf16"):
        model.set_precision("fp32")


def test_three_classes_non_square_and_graph_replay():
    """num_classes = 3, in_channels = 1, 48 x 80 input, batch 3: eval forward with randomised running statistics against
    the oracle, the same forward replayed from one CUDA graph, and one training step's gradients of the output stage."""
    from unet_zoo_b200.train_step import GraphedInference
    model = seeded_model("multiresunet", in_channels=1, num_classes=3)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    for k, v in sd.items():
        if k.endswith("running_mean"):
            v.normal_(0.0, 0.2, generator=torch.Generator().manual_seed(len(k)))
        if k.endswith("running_var"):
            v.uniform_(0.5, 1.5, generator=torch.Generator().manual_seed(len(k) + 1))
    model.load_state_dict(sd)
    model.cuda().eval()
    x, _ = O.synthetic_batch(3, 1, 48, 80)
    with torch.no_grad():
        out = model(x.cuda())
        ref = O.multiresunet_forward({k: v.clone() for k, v in sd.items()}, x, False)
    assert out.shape == ref.shape == (3, 3, 48, 80)
    assert rel_err(out.cpu(), ref) < EVAL_TOL
    g = GraphedInference(model, x.cuda())
    assert torch.equal(g(x.cuda()), out)
    # a checkpoint written by the reference loads unchanged and changes the result
    sd2 = {k: (v * 1.5 if k.endswith("conv_final.conv1.weight") else v) for k, v in sd.items()}
    model.load_state_dict(sd2)
    with torch.no_grad():
        out2 = model(x.cuda())
    assert not torch.equal(out2, out)
    model.load_state_dict(sd)
    out_t = _train_step(model, x, None, O.gaussian_projection_loss)
    ref_t, _, ref_grads = O.train_step("multiresunet", {k: v.clone() for k, v in sd.items()}, x, None,
                                       loss_fn=O.gaussian_projection_loss)
    assert rel_err(out_t, ref_t.detach()) < LOGIT_TOL
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters()}
    for k in ("conv_final.conv1.weight", "multiresblock9.conv2d_bn_7x7.conv1.weight", "upsample9.weight", "upsample9.bias"):
        assert float((grads[k] - ref_grads[k]).norm() / ref_grads[k].norm().clamp_min(1e-20)) < 1.0, k


def test_other_base_width_and_train_eval_transitions(monkeypatch):
    """filters = 16 (branch widths 4 / 8 / 13 ... 71 / 142 / 213: other paddings, same code), two training steps with an
    in-place parameter update between them, then eval: the padded operands follow the parameters, the inference memo is
    dropped on train() and rebuilt on the next eval forward."""
    model = seeded_model("multiresunet", filters=16)
    assert model.multiresblock1.conv2d_bn_7x7.conv1.out_channels == 13 and model.in_filters5 == 426
    x, mask = O.synthetic_batch(2, 3, 64, 64)
    sd = copy.deepcopy(model.state_dict())
    monkeypatch.setattr(O.Runner, "round_storage", torch.bfloat16)
    out1 = _train_step(model, x, mask, O.gaussian_projection_loss)
    ref1, _, _ = O.train_step("multiresunet", sd, x, mask, loss_fn=O.gaussian_projection_loss)
    assert rel_err(out1, ref1.detach()) < MATCHED_TOL
    gen = torch.Generator().manual_seed(9)
    with torch.no_grad():                                      # an "optimizer step" that bypasses the version counters
        for k, p in model.named_parameters():
            step = torch.randn(p.shape, generator=gen) * 0.3 * float(p.detach().std() if p.numel() > 1 else 1.0)
            p.data.add_(step.cuda())
            sd[k].add_(step)
            p.grad = None
    out2 = _train_step(model, x, mask, O.gaussian_projection_loss)
    ref2, _, _ = O.train_step("multiresunet", sd, x, mask, loss_fn=O.gaussian_projection_loss)
    assert rel_err(out2, ref2.detach()) < MATCHED_TOL
    monkeypatch.setattr(O.Runner, "round_storage", None)
    model.eval()
    with torch.no_grad():
        oe = model(x.cuda()).cpu()
        ref = O.multiresunet_forward({k: v.detach().cpu().clone() for k, v in model.state_dict().items()}, x, False)
    assert rel_err(oe, ref) < EVAL_TOL
    for k, v in sd.items():
        if "num_batches" in k:
            assert int(model.state_dict()[k]) == int(v)


def test_graphed_train_step_matches_eager_steps_and_learns():
    """One CUDA graph per step (GraphedTrainStep) captures the padded-operand refresh, the statistics write-back and the
    gradient gather with everything else: same losses and same parameters / buffers as three eager steps.  Then forty
    replays on one batch: the loss goes down (the optimizer's updates reach the padded operands)."""
    from unet_zoo_b200.train_step import GraphedTrainStep
    x, m = O.synthetic_batch(2, 3, 64, 64)
    x, m = x.cuda(), m.cuda()
    crit = torch.nn.BCEWithLogitsLoss()
    results = []
    for graphed in (False, True):
        model = seeded_model("multiresunet").cuda().train()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=1e-5, fused=True, capturable=True)
        losses = []
        if graphed:
            step = GraphedTrainStep(model, crit, opt, x, m, max_norm=1.0, warmup=1)     # restores the state it was given
            for _ in range(3):
                losses.append(float(step(x, m)))
        else:
            for _ in range(3):
                opt.zero_grad(set_to_none=True)
                loss = crit(model(x), m)
                loss.backward()
                torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
                opt.step()
                losses.append(float(loss))
        results.append((losses, {k: v.detach().clone() for k, v in model.state_dict().items()}))
    (l0, s0), (l1, s1) = results
    assert max(abs(a - b) for a, b in zip(l0, l1)) < 1e-6, (l0, l1)
    for k in s0:
        if s0[k].dtype.is_floating_point:
            assert torch.allclose(s0[k], s1[k], rtol=0, atol=1e-6), k
        else:
            assert int(s0[k]) == int(s1[k]), k
    more = [float(step(x, m)) for _ in range(40)]
    assert more[-1] < 0.8 * l1[0], (l1[0], more[-1])


def test_epoch_loops_with_the_fused_tail():
    """training.train_one_epoch / validate_one_epoch (drop-ins for training_loop.py:86-180) with the fused criterion and the
    flat optimizer on this model: finite numbers, the first batch's loss equals a plain forward's, parameters moved,
    running statistics and counters advanced (batch_norm1: two per forward)."""
    from unet_zoo_b200 import training as T
    from unet_zoo_b200.loss import BCEDiceWithLogits
    from unet_zoo_b200.optim import FlatAdamW
    data = []
    for i in range(3):
        g = torch.Generator().manual_seed(200 + i)
        data.append((torch.randn(2, 3, 64, 64, generator=g), (torch.rand(2, 1, 64, 64, generator=g) > 0.5).float(), f"p{i}"))
    model = seeded_model("multiresunet").to(DEV)
    probe = seeded_model("multiresunet").to(DEV).train()
    with torch.no_grad():
        first = float(F.binary_cross_entropy_with_logits(probe(data[0][0].to(DEV)), data[0][1].to(DEV)))
    w0 = model.multiresblock5.conv2d_bn_7x7.conv1.weight.detach().clone()
    opt = FlatAdamW(model.parameters(), lr=1e-3, weight_decay=1e-5, max_norm=1.0)
    crit = BCEDiceWithLogits()
    one = T.train_one_epoch(model, data[:1], opt, crit, DEV, None, 0, "multiresunet", None)
    assert abs(one[0] - first) < 1e-5, (one, first)
    got = T.train_one_epoch(model, data, opt, crit, DEV, None, 1, "multiresunet", None)
    val = T.validate_one_epoch(model, data, crit, DEV, "multiresunet", None)
    for v in tuple(got[:2]) + tuple(val[:2]):
        assert v == v and abs(v) < 10.0
    assert not torch.equal(model.multiresblock5.conv2d_bn_7x7.conv1.weight.detach(), w0)
    assert int(model.multiresblock3.batch_norm1.num_batches_tracked) == 8        # 4 training forwards x 2
    assert int(model.respath2.batch_norm_initial.num_batches_tracked) == 4
    assert float(model.conv_final.batchnorm.running_var) != 1.0


def test_baseline_resolution_512(monkeypatch):
    """2 x 3 x 512 x 512 (the resolution of every BASELINE config; multi-wave persistent scheduling, split-K plans and
    pooled / unpooled tails at their real sizes): one training step against the rounding-matched oracle, then the eval
    forward on the statistics that step produced against the plain fp32 oracle."""
    torch.set_num_threads(16)
    x, mask = O.synthetic_batch(2, 3, 512, 512)
    model = seeded_model("multiresunet")
    sd = copy.deepcopy(model.state_dict())
    out = _train_step(model, x, mask, O.gaussian_projection_loss)
    monkeypatch.setattr(O.Runner, "round_storage", torch.bfloat16)
    ref, _, ref_grads = O.train_step("multiresunet", sd, x, mask, loss_fn=O.gaussian_projection_loss)
    monkeypatch.setattr(O.Runner, "round_storage", None)
    ref = ref.detach()
    err = rel_err(out, ref)
    agree = float((O.threshold_mask(out) == O.threshold_mask(ref)).float().mean())
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters()}
    e_all, per = grad_errors(grads, ref_grads)
    ws = _worst_stats_error(model.state_dict(), sd)
    parity_report("multiresunet_512/matched", logits=err, agree=agree, grad_all=e_all, grad_worst=per[:3], stats_worst=ws)
    assert err < MATCHED_TOL and agree >= MATCHED_AGREE, (err, agree)
    assert e_all < MATCHED_GRAD, (e_all, per[:3])
    assert ws[0] < 5e-2, ws
    state = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        oe = model(x.cuda()).cpu()
        re = O.multiresunet_forward(state, x, False)
    e_eval = rel_err(oe, re)
    a_eval = float((O.threshold_mask(oe) == O.threshold_mask(re)).float().mean())
    parity_report("multiresunet_512/eval", logits=e_eval, agree=a_eval)
    assert e_eval < EVAL_TOL and a_eval >= 0.999, (e_eval, a_eval)
