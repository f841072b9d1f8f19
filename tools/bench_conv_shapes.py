"""Same conv3x3 (CIN -> COUT, default 64 -> 64) over the same number of pixels laid out as different (N, H, W): probes how
the halo-box kernel's throughput depends on the image geometry (row pitch of the boxes in DRAM) rather than on the work."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unet_zoo_b200 import engine as E  # noqa: E402
from unet_zoo_b200 import _lib as L  # noqa: E402

cin, cout = int(os.environ.get("CIN", "64")), int(os.environ.get("COUT", "64"))
dev = torch.device("cuda", 0)
ctx = E.Ctx(dev, training=True, record=False)
for (N, H, W) in ((32, 512, 512), (32, 4096, 64), (32, 64, 4096), (128, 256, 256), (512, 128, 128), (2048, 64, 64), (8192, 32, 32),
                  (32, 2048, 128), (32, 128, 2048)):
    x = E.Act(ctx.new_buf(N, H, W, cin))
    x.buf.data.normal_()
    w = torch.randn(cout, 9, cin, device=dev).to(torch.bfloat16)
    out = E.Act(ctx.new_buf(N, H, W, cout))
    st = ctx.f32(ctx.stats_rows, 2, cout)
    bias = torch.randn(cout, device=dev)
    times = []
    for i in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.conv(L.CONV3X3, [x], w, cout, bias, out.ptr, out.ld, H, W, st, L.FLAG_STATS)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            times.append(e0.elapsed_time(e1))
    ms = sum(times) / len(times)
    gb = 2.0 * N * H * W * (cin + cout) / 1e9
    print(f"{cin}->{cout} N={N} H={H} W={W}: {ms:.3f} ms  {gb / ms:.0f} GB/s algorithmic  {2.0 * N * H * W * cin * cout * 9 / ms / 1e9:.0f} TFLOP/s")
