"""ConvTranspose2d k2s2 forward (four 1x1 GEMMs through strided output views) at the unet decoder shapes, batch 32."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unet_zoo_b200 import engine as E  # noqa: E402
from unet_zoo_b200 import _lib as L  # noqa: E402

dev = torch.device("cuda", 0)
ctx = E.Ctx(dev, training=True, record=False)
N = 32
for (H, cin, cout) in ((32, 1024, 512), (64, 512, 256), (128, 256, 128), (256, 128, 64)):
    x = E.Act(ctx.new_buf(N, H, H, cin))
    x.buf.data.normal_()
    w = torch.randn(cout, 4, cin, device=dev).to(torch.bfloat16)
    out = E.Act(ctx.new_buf(N, 2 * H, 2 * H, 2 * cout), 0, cout)
    bias = torch.randn(cout, device=dev)
    times = []
    for i in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.conv(L.CONVT2X2, [x], w, cout, bias, out.ptr, out.ld, 2 * H, 2 * H, None, 0)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            times.append(e0.elapsed_time(e1))
    ms = min(times)
    fl = 2.0 * N * H * H * cin * cout * 4
    gb = 2.0 * N * H * H * (cin + 4 * cout) / 1e9
    print(f"convT {cin}->{cout} @{H}->{2 * H}: {ms:.3f} ms  {fl / ms / 1e9:.0f} TFLOP/s  {gb / ms * 1e3:.0f} GB/s")
