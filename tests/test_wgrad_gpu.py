"""3x3 weight gradient through uz_conv_wgrad on the halo-box kernel in all of its tile modes - single CTA / CTA pair,
N = 64 / WIDE (128 output channels per work unit, two passes over the tap pairs) - against torch's conv2d weight
gradient on the same bf16-rounded operands, including ragged channel counts and image sizes off the 8x16 chunk grid."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import bf16r, make_act, new_ctx, rel_l2
from unet_zoo_b200 import _lib as L

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,H,W,cin,cout", [
    (2, 32, 32, 64, 64),      # single CTA, N = 64
    (2, 32, 32, 64, 128),     # single CTA, WIDE
    (2, 32, 16, 64, 192),     # single CTA, WIDE, second block half empty
    (3, 16, 24, 128, 128),    # CTA pair, WIDE
    (2, 32, 32, 192, 320),    # pair with an odd number of input-channel blocks, ragged output block
    (1, 48, 40, 256, 64),     # CTA pair, N = 64 (narrow pair, SWIZZLE_64B halves)
    (5, 16, 8, 128, 1024),    # many output blocks, one chunk per image
])
def test_conv3x3_weight_gradient_all_tile_modes(N, H, W, cin, cout):
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(N * 1000 + cin + cout)
    x = bf16r(torch.randn(N, cin, H, W, generator=g)).to(dev)
    dz = bf16r(torch.randn(N, cout, H, W, generator=g) * 0.1).to(dev)
    ctx = new_ctx(record=False)
    xa = make_act(ctx, x, ld=cin + 8)
    dza = make_act(ctx, dz, ld=cout + 16, c0=8)
    dw = torch.full((cout, cin, 3, 3), float("nan"), device=dev)
    ctx.wgrad(L.CONV3X3, xa.ptr, N, H, W, cin, xa.ld, dza.ptr, cout, dza.ld, dw, cin, 0)
    torch.cuda.synchronize()
    w = torch.zeros(cout, cin, 3, 3, device=dev, requires_grad=True)
    (F.conv2d(x, w, padding=1) * dz).sum().backward()
    assert bool(torch.isfinite(dw).all())
    assert rel_l2(dw, w.grad) < 2e-3, rel_l2(dw, w.grad)
    # per-tap check: a wrong tap pairing would hide in the global norm of a large tensor
    for t in range(9):
        a, b = dw[:, :, t // 3, t % 3], w.grad[:, :, t // 3, t % 3]
        assert rel_l2(a, b) < 2e-3, (t, rel_l2(a, b))
    # accumulate into a slice of a wider gradient (concat input): dW[:, c_off : c_off + cin]
    wide = torch.ones(cout, cin + 64, 3, 3, device=dev)
    ctx.wgrad(L.CONV3X3, xa.ptr, N, H, W, cin, xa.ld, dza.ptr, cout, dza.ld, wide, cin + 64, 64, accumulate=True)
    torch.cuda.synchronize()
    assert rel_l2(wide[:, 64:] - 1.0, w.grad) < 2e-3 and bool((wide[:, :64] == 1.0).all())
