"""Operator-level parity at REAL layer shapes of BASELINE configs[1] (unet, 512 x 512): conv3x3 + BatchNorm + ReLU
forward, data gradient, weight gradient (every tile mode of the halo-box kernel the shapes select: single CTA N = 64,
narrow CTA pair, WIDE pair, virtual concat of two segments) and the BatchNorm backward passes, through the C ABI,
against PyTorch fp32 (cuDNN, TF32 off) on the same bf16-rounded operands with the same rounding points.

The small-shape operator tests (test_ops_gpu.py) run one wave of a few tiles; these run thousands of tiles per CTA
(persistent scheduling, accumulator ring wrap-around, split-K over > 4000 chunks)."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from gpu_util import bf16r, get_grad, make_act, new_ctx, relmax, rel_l2, run_tape, set_grad
from unet_zoo_b200 import engine as E

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _exact_fp32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(11)


@pytest.mark.parametrize("N,S,segs,cout,pool", [
    (2, 512, (64,), 64, True),          # down_convolution_1 second half: 64 -> 64 @512^2, skip + pooled outputs
    (2, 512, (64, 64), 64, False),      # up_convolution_4 first half: cat[up | skip] 128 -> 64 @512^2, two segments
    (2, 512, (128,), 64, False),        # same GEMM over ONE 128-channel operand: narrow CTA-pair weight gradient
    (32, 32, (1024,), 1024, False),     # bottle_neck second half: 1024 -> 1024 @32^2, batch 32 (WIDE pair, BN = 256)
])
def test_conv_bn_relu_train_fwd_bwd_at_layer_shape(N, S, segs, cout, pool):
    dev = torch.device("cuda", 0)
    cin = sum(segs)
    conv = nn.Conv2d(cin, cout, 3, padding=1).to(dev)
    bn = nn.BatchNorm2d(cout).to(dev)
    with torch.no_grad():
        conv.weight.copy_(bf16r(conv.weight))
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    x = bf16r(torch.randn(N, cin, S, S, device=dev))
    ctx = new_ctx(params=[conv.weight, conv.bias, bn.weight, bn.bias])
    # the segments are channel slices of ONE wide buffer, like the decoder's concat buffer
    xa = make_act(ctx, x)
    xs, c0 = [], 0
    for c in segs:
        xs.append(E.Act(xa.buf, c0, c))
        c0 += c
    out = ctx.new_act(N, S, S, cout)
    pooled = ctx.new_act(N, S // 2, S // 2, cout) if pool else None
    E.conv_bn_act(ctx, xs, conv, bn, out=out, pooled=pooled)
    torch.cuda.synchronize()

    xr = x.clone().requires_grad_(True)
    w = conv.weight.detach().clone().requires_grad_(True)
    g = bn.weight.detach().clone().requires_grad_(True)
    be = bn.bias.detach().clone().requires_grad_(True)
    z = F.conv2d(xr, w, conv.bias.detach(), padding=1)
    zq = z + (bf16r(z.detach()) - z.detach())                      # straight-through rounding of the stored z
    rm, rv = torch.zeros(cout, device=dev), torch.ones(cout, device=dev)
    y = F.relu(F.batch_norm(zq, rm, rv, g, be, training=True, momentum=0.1, eps=1e-5))
    yq = y + (bf16r(y.detach()) - y.detach())
    assert relmax(out.to_nchw(), yq.detach()) < 1e-2
    assert relmax(bn.running_mean, rm) < 2e-3 and relmax(bn.running_var, rv) < 2e-3
    dy = bf16r(torch.randn_like(y) * 0.1)
    (yq * dy).sum().backward(retain_graph=pool)
    set_grad(out, dy)
    if pool:
        p = F.max_pool2d(yq, 2, 2)
        assert relmax(pooled.to_nchw(), p.detach()) < 1e-2
        dp = bf16r(torch.randn_like(p) * 0.1)
        (p * dp).sum().backward()
        set_grad(pooled, dp)
    run_tape(ctx)
    ctx.join_side()
    torch.cuda.synchronize()
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2                     # data gradient (bf16 output)
    dw = ctx.pgrads.view(conv.weight)
    assert rel_l2(dw, w.grad) < 5e-3                                # split-K fp32 weight gradient
    for t in range(9):                                              # per tap: a wrong tap pairing hides in the norm
        assert rel_l2(dw[:, :, t // 3, t % 3], w.grad[:, :, t // 3, t % 3]) < 1e-2, t
    assert rel_l2(ctx.pgrads.view(bn.weight), g.grad) < 3e-3
    assert rel_l2(ctx.pgrads.view(bn.bias), be.grad) < 3e-3


def test_conv_transpose_at_layer_shape():
    """up_convolution_4.up: ConvTranspose2d 128 -> 64, 256^2 -> 512^2 into the first half of the level's concat buffer,
    forward, data gradient and weight gradient."""
    dev = torch.device("cuda", 0)
    N, H, cin = 2, 256, 128
    up = nn.ConvTranspose2d(cin, cin // 2, 2, 2).to(dev)
    with torch.no_grad():
        up.weight.copy_(bf16r(up.weight))
    x = bf16r(torch.randn(N, cin, H, H, device=dev))
    ctx = new_ctx(params=[up.weight, up.bias])
    xa = make_act(ctx, x)
    cat = ctx.new_buf(N, 2 * H, 2 * H, cin)
    cat.data.fill_(7.0)
    out = E.Act(cat, 0, cin // 2)
    E.conv_transpose2x2(ctx, xa, up, out)
    xr = x.clone().requires_grad_(True)
    w = up.weight.detach().clone().requires_grad_(True)
    b = up.bias.detach().clone().requires_grad_(True)
    y = F.conv_transpose2d(xr, w, b, stride=2)
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), y.detach()) < 1e-2
    assert bool((cat.data[..., cin // 2:] == 7.0).all())            # the skip half is untouched
    dy = bf16r(torch.randn_like(y) * 0.1)
    (y * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    ctx.join_side()
    torch.cuda.synchronize()
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2
    assert rel_l2(ctx.pgrads.view(up.weight), w.grad) < 5e-3
    assert rel_l2(ctx.pgrads.view(up.bias), b.grad) < 5e-3
