"""Operators behind `multiresunet` against plain PyTorch fp32 math on the same bf16-rounded
operands: convolutions whose channel counts are multiples of 8 but of nothing larger (the zero-padded widths of the
Multiresblock branches: 8 / 24 / 32 / 40 / 56 / 72 / 112 / 120 / 144 / 216 / 224 / 288 / 432 / 864), BatchNorm2d(affine=False),
the residual add with ReLU, and the NHWC bf16 -> NCHW fp32 extraction of the normalised logits."""
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
    torch.manual_seed(11)


@pytest.mark.parametrize("N,H,W,cin,cout,k,relu", [
    (2, 32, 32, 8, 8, 3, True),          # Multiresblock 1: padded image -> first branch
    (2, 32, 32, 8, 24, 3, True),
    (2, 32, 32, 24, 32, 3, True),
    (2, 32, 32, 8, 64, 1, False),        # the 1x1 shortcut of block 1
    (2, 32, 32, 64, 24, 3, True),
    (2, 16, 16, 120, 40, 3, True),
    (2, 16, 16, 40, 56, 3, True),
    (2, 16, 16, 120, 120, 1, False),
    (1, 16, 16, 224, 72, 3, True),
    (1, 16, 16, 72, 112, 3, True),
    (1, 16, 16, 432, 144, 3, True),
    (1, 16, 16, 144, 288, 3, True),
    (1, 16, 16, 288, 432, 3, True),
    (1, 16, 16, 512, 864, 1, False),
    (2, 32, 32, 64, 8, 1, False),        # conv_final with its single class padded to 8
])
def test_conv_bn_with_channel_counts_that_are_only_multiples_of_8(N, H, W, cin, cout, k, relu):
    dev = torch.device("cuda", 0)
    conv = nn.Conv2d(cin, cout, k, padding=k // 2).to(dev)
    bn = nn.BatchNorm2d(cout, affine=False).to(dev)
    with torch.no_grad():
        conv.weight.copy_(bf16r(conv.weight))
    x = bf16r(torch.randn(N, cin, H, W, device=dev))
    ctx = new_ctx(params=[conv.weight, conv.bias])
    xa = make_act(ctx, x)
    rm0, rv0 = bn.running_mean.clone(), bn.running_var.clone()
    out = E.conv_bn_act(ctx, [xa], conv, bn, relu=relu, mode=L.CONV3X3 if k == 3 else L.CONV1X1)
    torch.cuda.synchronize()

    xr = x.clone().requires_grad_(True)
    w = conv.weight.detach().clone().requires_grad_(True)
    z = F.conv2d(xr, w, conv.bias.detach(), padding=k // 2)
    zq = z + (bf16r(z.detach()) - z.detach())
    rm, rv = rm0.clone(), rv0.clone()
    y = F.batch_norm(zq, rm, rv, None, None, training=True, momentum=0.1, eps=1e-5)
    if relu:
        y = F.relu(y)
    assert relmax(out.to_nchw(), y.detach()) < 1e-2
    assert relmax(bn.running_mean, rm) < 2e-3 and relmax(bn.running_var, rv) < 2e-3
    dy = bf16r(torch.randn_like(y) * 0.1)
    (y * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2
    assert rel_l2(ctx.pgrads.view(conv.weight), w.grad) < 5e-3


@pytest.mark.parametrize("N,H,W,cin,cout", [(1, 8, 8, 864, 256), (2, 8, 16, 432, 128), (2, 16, 16, 224, 64), (2, 16, 16, 120, 32)])
def test_conv_transpose_from_padded_widths(N, H, W, cin, cout):
    dev = torch.device("cuda", 0)
    up = nn.ConvTranspose2d(cin, cout, 2, 2).to(dev)
    with torch.no_grad():
        up.weight.copy_(bf16r(up.weight))
    x = bf16r(torch.randn(N, cin, H, W, device=dev))
    ctx = new_ctx(params=[up.weight, up.bias])
    xa = make_act(ctx, x)
    cat = ctx.new_buf(N, 2 * H, 2 * W, 2 * cout)
    out = E.conv_transpose2x2(ctx, xa, up, E.Act(cat, 0, cout))
    torch.cuda.synchronize()
    xr = x.clone().requires_grad_(True)
    w = up.weight.detach().clone().requires_grad_(True)
    b = up.bias.detach().clone().requires_grad_(True)
    y = F.conv_transpose2d(xr, w, b, stride=2)
    assert relmax(out.to_nchw(), y.detach()) < 1e-2
    dy = bf16r(torch.randn_like(y) * 0.1)
    (y * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2
    assert rel_l2(ctx.pgrads.view(up.weight), w.grad) < 5e-3
    assert rel_l2(ctx.pgrads.view(up.bias), b.grad) < 2e-3
