import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch, torch.nn as nn, torch.nn.functional as F
from gpu_util import bf16r, get_grad, make_act, new_ctx, relmax, rel_l2, run_tape, set_grad
from unet_zoo_b200.models import _graphs as G
from unet_zoo_b200.models.blocks import ResidualConv
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
DEV = "cuda:0"
for (N, H, W, cin, cout, stride) in [(2, 16, 16, 64, 128, 2), (2, 16, 16, 128, 64, 1), (1, 32, 16, 64, 64, 2), (4, 32, 32, 64, 64, 1)]:
    torch.manual_seed(11)
    blk = ResidualConv(cin, cout, stride, 1).to(DEV)
    with torch.no_grad():
        for m in blk.modules():
            if isinstance(m, nn.Conv2d): m.weight.copy_(bf16r(m.weight))
            if isinstance(m, nn.BatchNorm2d): m.weight.uniform_(0.5, 1.5); m.bias.uniform_(-0.3, 0.3)
    x = bf16r(torch.randn(N, cin, H, W, device=DEV))
    ctx = new_ctx(params=list(blk.parameters()))
    xa = make_act(ctx, x)
    out = G._residual_conv(ctx, xa, blk)
    xr = x.clone().requires_grad_(True)
    bn = lambda t, m: F.batch_norm(t, None, None, m.weight, m.bias, training=True, eps=m.eps)
    cb, cs = blk.conv_block, blk.conv_skip
    h0 = F.relu(bn(xr, cb[0])); h0.retain_grad()
    h = F.conv2d(h0, cb[2].weight, None, stride=stride, padding=1)
    h = F.relu(bn(h, cb[3]))
    h = F.conv2d(h, cb[5].weight, None, padding=1)
    ref = h + bn(F.conv2d(xr, cs[0].weight, None, stride=stride), cs[1])
    dy = bf16r(torch.randn_like(ref) * 0.1)
    (ref * dy).sum().backward()
    set_grad(out, dy)
    run_tape(ctx)
    print((N, H, W, cin, cout, stride), "fwd", relmax(out.to_nchw(), ref.detach()), "dx", rel_l2(get_grad(xa), xr.grad))
    for name, p in blk.named_parameters():
        print("   ", name, round(rel_l2(ctx.pgrads.view(p), p.grad), 4))
