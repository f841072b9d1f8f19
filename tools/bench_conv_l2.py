"""conv3x3 CIN -> COUT (default 64 -> 64) at 512 x 512 for growing batch: per-tile time when the operands fit in L2
(batch 1: 67 MB in + out) against the DRAM-streaming regime (batch 32) - separates the per-tile pipeline cost from HBM."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unet_zoo_b200 import engine as E  # noqa: E402
from unet_zoo_b200 import _lib as L  # noqa: E402

cin, cout = int(os.environ.get("CIN", "64")), int(os.environ.get("COUT", "64"))
mode = {"c3": L.CONV3X3, "c1": L.CONV1X1}[os.environ.get("KIND", "c3")]
taps = 9 if mode == L.CONV3X3 else 1
dev = torch.device("cuda", 0)
ctx = E.Ctx(dev, training=True, record=False)
for N in (1, 2, 4, 8, 32):
    H = W = 512
    x = E.Act(ctx.new_buf(N, H, W, cin))
    x.buf.data.normal_()
    w = torch.randn(cout, taps, cin, device=dev).to(torch.bfloat16)
    out = E.Act(ctx.new_buf(N, H, W, cout))
    bias = torch.randn(cout, device=dev)
    times = []
    for i in range(12):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.conv(mode, [x], w, cout, bias, out.ptr, out.ld, H, W, None, 0)
        e1.record()
        torch.cuda.synchronize()
        if i >= 4:
            times.append(e0.elapsed_time(e1))
    ms = min(times)
    tiles = N * H * W / 128
    print(f"{os.environ.get('KIND', 'c3')} {cin}->{cout} N={N}: {ms * 1e3:.1f} us, {ms * 1e3 / (tiles / 148):.3f} us per tile per SM, "
          f"{2.0 * N * H * W * (cin + cout) / ms / 1e6:.0f} GB/s")
