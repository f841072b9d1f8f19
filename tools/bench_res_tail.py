"""Times the fused inference tail of ResidualConv (engine.residual_conv_tail_fused: 3x3 over h + 1x1 over x in one GEMM)
against the two-launch form at one layer shape.  SHAPE="N,H,W,ch,cx,cout,s2" (default: up_residual_conv2 of resunet)."""
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unet_zoo_b200 import _lib as L  # noqa: E402
from unet_zoo_b200 import engine as E  # noqa: E402


def main():
    N, H, W, ch, cx, cout, s2 = [int(v) for v in os.environ.get("SHAPE", "32,256,256,128,256,128,0").split(",")]
    dev = torch.device("cuda", 0)
    conv2 = nn.Conv2d(ch, cout, 3, padding=1, bias=False).to(dev)
    conv_s = nn.Conv2d(cx, cout, 1, stride=2 if s2 else 1, bias=False).to(dev)
    bn_s = nn.BatchNorm2d(cout).to(dev).eval()
    ctx = E.Ctx(dev, training=False, record=False)
    ctx.eval_cache = {}
    h = ctx.new_act(N, H, W, ch)
    x = ctx.new_act(N, H * (2 if s2 else 1), W * (2 if s2 else 1), cx)
    h.buf.data.normal_()
    x.buf.data.normal_()
    out = ctx.new_act(N, H, W, cout)

    def fused():
        E.residual_conv_tail_fused(ctx, h, x, conv2, conv_s, bn_s, out, bool(s2))

    def plain():
        E.conv_bn_act(ctx, [x], conv_s, bn_s, out=out, relu=False, mode=L.CONV1X1_S2 if s2 else L.CONV1X1)
        E.conv_plain(ctx, [h], conv2, out=out, accumulate=True)

    for name, fn in (("fused", fused), ("two launches", plain)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{name}: {e0.elapsed_time(e1) / 20:.3f} ms")


main()
