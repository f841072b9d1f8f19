"""The bandwidth operators around unet_transformer's attention blocks (csrc/transformer_ops.cu) through the engine, forward
and backward, against PyTorch fp32 on the same bf16 operands: positional-encoding add (+ MaxPool2d of the sum), adaptive
average pooling (down, identity and up), bilinear resize with align_corners=True at assorted ratios."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import bf16r, get_grad, make_act, new_ctx, relmax, rel_l2, run_tape, set_grad
from unet_zoo_b200 import engine as E

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(5)


@pytest.mark.parametrize("N,C,H,W", [(2, 64, 32, 24), (1, 256, 9, 7), (3, 8, 16, 16)])
def test_add_posenc_plain_and_pooled(N, C, H, W):
    x = bf16r(torch.randn(N, C, H, W, device=DEV))
    pe = torch.randn(H, W, C, device=DEV)                       # any fixed table
    pe_nchw = pe.permute(2, 0, 1)[None]
    # plain: y = x + pe
    ctx = new_ctx()
    xa = make_act(ctx, x)
    y, _ = E.add_posenc(ctx, xa, pe)
    ref = bf16r(x + pe_nchw)
    torch.cuda.synchronize()
    assert relmax(y.to_nchw(), ref) < 1e-6
    dy = bf16r(torch.randn_like(ref))
    set_grad(y, dy)
    run_tape(ctx)
    assert torch.equal(get_grad(xa), dy)
    # pooled only: MaxPool2d(2)(x + pe), gradient routed to the first maximum
    ctx = new_ctx()
    xa = make_act(ctx, x)
    _, p = E.add_posenc(ctx, xa, pe, pooled=ctx.new_act(N, H // 2, W // 2, C))
    xr = x.clone().requires_grad_(True)
    s = xr + pe_nchw
    sq = s + (bf16r(s.detach()) - s.detach())
    pr = F.max_pool2d(sq, 2)
    torch.cuda.synchronize()
    assert relmax(p.to_nchw(), pr.detach()) < 1e-6
    dp = bf16r(torch.randn_like(pr))
    (pr * dp).sum().backward()
    set_grad(p, dp)
    run_tape(ctx)
    assert relmax(get_grad(xa), xr.grad) < 1e-6


@pytest.mark.parametrize("N,C,H,W,OH,OW", [(2, 64, 128, 128, 64, 64), (1, 128, 64, 64, 64, 64), (2, 64, 8, 8, 64, 64),
                                           (1, 64, 100, 70, 64, 64), (2, 8, 5, 9, 3, 4)])
def test_adaptive_avgpool_fwd_bwd(N, C, H, W, OH, OW):
    x = bf16r(torch.randn(N, C, H, W, device=DEV))
    ctx = new_ctx()
    xa = make_act(ctx, x)
    out = E.adaptive_avgpool(ctx, xa, OH, OW)
    xr = x.clone().requires_grad_(True)
    ref = F.adaptive_avg_pool2d(xr, (OH, OW))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), ref.detach()) < 1e-2
    dy = bf16r(torch.randn_like(ref))
    (ref * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad) < 1e-2


@pytest.mark.parametrize("N,C,H,W,OH,OW", [(2, 64, 64, 64, 128, 128), (1, 64, 64, 64, 512, 512), (2, 128, 64, 64, 16, 16),
                                           (1, 8, 5, 7, 13, 9), (2, 64, 64, 64, 64, 64), (1, 8, 1, 1, 4, 4)])
def test_bilinear_resize_fwd_bwd(N, C, H, W, OH, OW):
    x = bf16r(torch.randn(N, C, H, W, device=DEV))
    ctx = new_ctx()
    xa = make_act(ctx, x)
    out = E.bilinear_resize(ctx, xa, OH, OW)
    xr = x.clone().requires_grad_(True)
    ref = F.interpolate(xr, size=(OH, OW), mode="bilinear", align_corners=True)
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), ref.detach()) < 1e-2
    dy = bf16r(torch.randn_like(ref))
    (ref * dy).sum().backward()
    set_grad(out, dy)
    prev = bf16r(torch.randn_like(x))
    set_grad(xa, prev)                                          # accumulate onto an existing gradient
    run_tape(ctx)
    assert rel_l2(get_grad(xa), xr.grad + prev) < 1e-2
