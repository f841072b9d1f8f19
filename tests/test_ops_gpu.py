"""Operator-level parity on the GPU: every fused operator (through the C ABI) against plain PyTorch fp32 math on the
same bf16-rounded operands.  Tolerances: results stored in bf16 carry one rounding (2^-9 relative) -> 1e-2 of the
tensor's max; fp32 reductions (parameter gradients, statistics) -> 2e-3."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from gpu_util import bf16r, get_grad, make_act, new_ctx, relmax, rel_l2, run_tape, set_grad
from unet_zoo_b200 import engine as E
from unet_zoo_b200 import _lib as L

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _exact_fp32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(7)


def _conv_bn(cin, cout, dev):
    conv = nn.Conv2d(cin, cout, 3, padding=1).to(dev)
    bn = nn.BatchNorm2d(cout).to(dev)
    with torch.no_grad():
        conv.weight.copy_(bf16r(conv.weight))
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    return conv, bn


@pytest.mark.parametrize("N,H,W,cin,cout,ld_extra", [(2, 32, 32, 64, 64, 0), (3, 16, 16, 128, 256, 64), (2, 16, 32, 32, 32, 0)])
def test_conv_bn_relu_pool_train_fwd_bwd(N, H, W, cin, cout, ld_extra):
    dev = torch.device("cuda", 0)
    conv, bn = _conv_bn(cin, cout, dev)
    x = bf16r(torch.randn(N, cin, H, W, device=dev))
    ctx = new_ctx(params=[conv.weight, conv.bias, bn.weight, bn.bias])
    xa = make_act(ctx, x, ld=cin + 8)
    out = E.Act(ctx.new_buf(N, H, W, cout + ld_extra), ld_extra, cout)     # a slice of a wider (concat) buffer
    pooled = ctx.new_act(N, H // 2, W // 2, cout)
    rm0, rv0 = bn.running_mean.clone(), bn.running_var.clone()
    E.conv_bn_act(ctx, [xa], conv, bn, out=out, pooled=pooled)
    torch.cuda.synchronize()

    # reference with the same rounding points: z -> bf16, y -> bf16
    xr = x.clone().requires_grad_(True)
    w = conv.weight.detach().clone().requires_grad_(True)
    b = conv.bias.detach().clone().requires_grad_(True)
    g = bn.weight.detach().clone().requires_grad_(True)
    be = bn.bias.detach().clone().requires_grad_(True)
    z = F.conv2d(xr, w, b, padding=1)
    zq = z + (bf16r(z.detach()) - z.detach())                                # straight-through rounding
    rm, rv = rm0.clone(), rv0.clone()
    y = F.relu(F.batch_norm(zq, rm, rv, g, be, training=True, momentum=0.1, eps=1e-5))
    yq = y + (bf16r(y.detach()) - y.detach())
    p = F.max_pool2d(yq, 2, 2)
    assert relmax(out.to_nchw(), yq.detach()) < 1e-2
    assert relmax(pooled.to_nchw(), p.detach()) < 1e-2
    assert relmax(bn.running_mean, rm) < 2e-3 and relmax(bn.running_var, rv) < 2e-3
    assert int(bn.num_batches_tracked) == 1

    dy = bf16r(torch.randn_like(y) * 0.1)
    dp = bf16r(torch.randn_like(p) * 0.1)
    (yq * dy).sum().backward(retain_graph=True)
    (p * dp).sum().backward()
    set_grad(out, dy)
    set_grad(pooled, dp)
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2
    assert rel_l2(ctx.pgrads.view(conv.weight), w.grad) < 5e-3
    assert rel_l2(ctx.pgrads.view(bn.weight), g.grad) < 2e-3
    assert rel_l2(ctx.pgrads.view(bn.bias), be.grad) < 2e-3
    # conv bias under train-mode BN: the true gradient is identically zero (sum of dz); ours is rounding noise
    assert float(ctx.pgrads.view(conv.bias).abs().max()) < 2e-2 * float(dy.abs().sum() / cout) + 1e-3


def test_maxpool_tie_goes_to_first_maximum():
    """nn.MaxPool2d backward routes to the first maximum in row-major window order; post-ReLU zeros tie often."""
    dev = torch.device("cuda", 0)
    C = 8
    ctx = new_ctx()
    z = torch.zeros(1, C, 2, 2, device=dev)
    z[0, 0] = torch.tensor([[1.0, 1.0], [1.0, 1.0]])         # 4-way tie -> (0,0)
    z[0, 1] = torch.tensor([[0.5, 2.0], [2.0, 0.1]])         # tie between (0,1) and (1,0) -> (0,1)
    z[0, 2] = torch.tensor([[-1.0, -2.0], [-3.0, 3.0]])      # unique max at (1,1)
    za = make_act(ctx, z)
    bnbuf = torch.zeros(4, C, device=dev)
    bnbuf[1] = 1.0
    bnbuf[2] = 1.0                                           # mean 0, invstd 1, scale 1, shift 0
    dpool = torch.ones(1, C, 1, 1, device=dev)
    dpa = make_act(ctx, dpool)
    dz = ctx.new_act(1, 2, 2, C)
    dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    ctx.bn_bwd(za, None, 0, dpa.ptr, dpa.ld, bnbuf, True, dz.ptr, dz.ld, dg, db, None)
    torch.cuda.synchronize()
    assert db[0] == 1 and db[1] == 1 and db[2] == 1          # exactly one routed unit per window
    # dgamma = sum g*x^ = value of z at the routed position
    assert float(dg[0]) == 1.0 and float(dg[1]) == 2.0 and float(dg[2]) == 3.0
    assert float(db[3]) == 0.0                               # all-zero window: ReLU mask kills the routed gradient


@pytest.mark.parametrize("N,H,W,cin", [(2, 16, 16, 128), (1, 8, 24, 64)])
def test_conv_transpose_into_concat_slice(N, H, W, cin):
    dev = torch.device("cuda", 0)
    up = nn.ConvTranspose2d(cin, cin // 2, 2, 2).to(dev)
    with torch.no_grad():
        up.weight.copy_(bf16r(up.weight))
    x = bf16r(torch.randn(N, cin, H, W, device=dev))
    ctx = new_ctx(params=[up.weight, up.bias])
    xa = make_act(ctx, x)
    cat = ctx.new_buf(N, 2 * H, 2 * W, cin)
    cat.data.fill_(7.0)
    out = E.Act(cat, 0, cin // 2)
    E.conv_transpose2x2(ctx, xa, up, out)
    xr = x.clone().requires_grad_(True)
    w = up.weight.detach().clone().requires_grad_(True)
    b = up.bias.detach().clone().requires_grad_(True)
    y = F.conv_transpose2d(xr, w, b, stride=2)
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), y.detach()) < 1e-2
    assert bool((cat.data[..., cin // 2:] == 7.0).all())     # the skip half of the concat buffer is untouched
    dy = bf16r(torch.randn_like(y))
    (y * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2
    assert rel_l2(ctx.pgrads.view(up.weight), w.grad) < 5e-3
    assert rel_l2(ctx.pgrads.view(up.bias), b.grad) < 2e-3


@pytest.mark.parametrize("C,K", [(64, 1), (32, 1), (64, 3), (64, 6)])
def test_head_fwd_bwd(C, K):
    dev = torch.device("cuda", 0)
    conv = nn.Conv2d(C, K, 1).to(dev)
    y = bf16r(torch.randn(2, C, 24, 40, device=dev))
    ctx = new_ctx(params=[conv.weight, conv.bias])
    ya = make_act(ctx, y, ld=C + 16, c0=16)
    logits = E.head1x1(ctx, ya, conv)
    yr = y.clone().requires_grad_(True)
    ref = F.conv2d(yr, conv.weight, conv.bias)
    torch.cuda.synchronize()
    assert relmax(logits, ref.detach()) < 1e-5               # fp32 in, fp32 out: only summation order differs
    dl = torch.randn_like(ref)
    gw, gb = torch.autograd.grad(ref, [conv.weight, conv.bias], dl, retain_graph=True)
    (gy,) = torch.autograd.grad(ref, [yr], dl)
    ctx.out_grads = [dl]
    run_tape(ctx)
    assert rel_l2(get_grad(ya), gy) < 1e-2
    assert rel_l2(ctx.pgrads.view(conv.weight), gw) < 1e-4
    assert rel_l2(ctx.pgrads.view(conv.bias), gb) < 1e-4


def test_stem_im2col_conv_matches_conv3x3():
    dev = torch.device("cuda", 0)
    conv, bn = _conv_bn(3, 64, dev)
    x = torch.randn(2, 3, 32, 48, device=dev)
    ctx = new_ctx(params=[conv.weight, conv.bias, bn.weight, bn.bias])
    a, kind = E.stem_input(ctx, x)
    assert kind == "im2col" and a.C == 32
    out = E.conv_bn_act(ctx, [a], conv, bn, stem=kind, need_dx=False)
    z = F.conv2d(bf16r(x), conv.weight, conv.bias, padding=1)
    zq = bf16r(z)
    ref = F.relu(F.batch_norm(zq, None, None, bn.weight, bn.bias, training=True, eps=1e-5))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), ref) < 1e-2
    dy = bf16r(torch.randn_like(ref) * 0.1)
    w = conv.weight.detach().clone().requires_grad_(True)
    z2 = F.conv2d(bf16r(x), w, conv.bias, padding=1)
    y2 = F.relu(F.batch_norm(z2, None, None, bn.weight, bn.bias, training=True, eps=1e-5))
    (y2 * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    assert rel_l2(ctx.pgrads.view(conv.weight), w.grad) < 2e-2


@pytest.mark.parametrize("N,cin,H,W", [(2, 3, 32, 48), (1, 1, 16, 16), (3, 7, 40, 72), (1, 3, 9, 130), (2, 5, 64, 64)])
def test_stem_im2col_is_exact(N, cin, H, W, monkeypatch):
    """uz_stem_im2col (tiled kernel and the first-generation one) against F.unfold: pure data movement + one bf16
    rounding, so bit-exact; column k = tap * cin + ci, zero padding columns, ragged sizes (tile = 8 x 64 pixels)."""
    dev = torch.device("cuda", 0)
    x = torch.randn(N, cin, H, W, device=dev)
    cols = F.unfold(x, 3, padding=1).view(N, cin, 9, H, W).permute(0, 3, 4, 2, 1).reshape(N, H, W, 9 * cin)
    outs = []
    for gen1 in (False, True):
        if gen1:
            monkeypatch.setenv("UZ_STEM_GEN1", "1")
        ctx = new_ctx(training=False, record=False)
        a, kind = E.stem_input(ctx, x)
        torch.cuda.synchronize()
        assert kind == "im2col" and a.C in (16, 32, 64) and a.C >= 9 * cin
        got = a.buf.data.float()
        assert torch.equal(got[..., :9 * cin], cols.to(torch.bfloat16).float())
        assert bool((got[..., 9 * cin:] == 0).all())
        outs.append(got)
    assert torch.equal(outs[0], outs[1])


def test_eval_mode_folds_batchnorm():
    dev = torch.device("cuda", 0)
    conv, bn = _conv_bn(64, 128, dev)
    with torch.no_grad():
        bn.running_mean.uniform_(-0.3, 0.3)
        bn.running_var.uniform_(0.5, 2.0)
    x = bf16r(torch.randn(2, 64, 16, 16, device=dev))
    ctx = new_ctx(training=False, record=False)
    xa = make_act(ctx, x)
    pooled = ctx.new_act(2, 8, 8, 128)
    out = E.conv_bn_act(ctx, [xa], conv, bn, pooled=pooled)
    ref = F.relu(F.batch_norm(F.conv2d(x, conv.weight, conv.bias, padding=1), bn.running_mean, bn.running_var,
                              bn.weight, bn.bias, training=False, eps=1e-5))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), ref) < 1.5e-2               # folded weights are rounded to bf16 once more
    assert torch.equal(pooled.to_nchw(), F.max_pool2d(out.to_nchw(), 2, 2))   # pooling is exact on stored values


def test_library_rejects_bad_arguments():
    lib = L.load()
    assert lib.uz_bn_apply(None, 64, 1, 4, 4, 64, None, None, 0, None, 0, None, 0, None, 0, None) < 0
    assert b"bn_apply" in lib.uz_last_error()
    t = torch.zeros(1, 4, 4, 12, dtype=torch.bfloat16, device="cuda")
    assert lib.uz_colsum(t.data_ptr(), 12, 16, 12, t.data_ptr(), 0, t.data_ptr(), 0, None) < 0   # C not multiple of 8
