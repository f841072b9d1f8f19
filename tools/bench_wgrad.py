"""Per-layer timing of the conv3x3 weight-gradient kernel (+ split-K reduce) at the unet batch-32 / 512x512 shapes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unet_zoo_b200 import engine as E  # noqa: E402
from unet_zoo_b200 import _lib as L  # noqa: E402

N = int(os.environ.get("BATCH", "32"))
dev = torch.device("cuda", 0)
ctx = E.Ctx(dev, training=True, record=False)
LAYERS = [(512, 64, 64), (512, 128, 64), (256, 64, 128), (256, 128, 128), (256, 256, 128), (128, 128, 256),
          (128, 256, 256), (128, 512, 256), (64, 256, 512), (64, 512, 512), (64, 1024, 512), (32, 512, 1024),
          (32, 1024, 1024)]
MODE, K = (L.CONV1X1, 1) if os.environ.get("MODE") == "1x1" else (L.CONV3X3, 3)      # MODE=1x1: first-generation kernel
only = os.environ.get("ONLY")        # e.g. ONLY=128,256,256 (H, cin, cout)
if only:
    LAYERS = [tuple(int(v) for v in only.split(","))]
tot = 0.0
for H, cin, cout in LAYERS:
    x = E.Act(ctx.new_buf(N, H, H, cin))
    dz = E.Act(ctx.new_buf(N, H, H, cout))
    x.buf.data.normal_()
    dz.buf.data.normal_()
    dw = torch.empty(cout, cin, K, K, device=dev)
    times = []
    for i in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.wgrad(MODE, x.ptr, N, H, H, cin, x.ld, dz.ptr, cout, dz.ld, dw, cin, 0)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            times.append(e0.elapsed_time(e1))
    ms = sum(times) / len(times)
    fl = 2.0 * N * H * H * cin * cout * K * K
    tot += ms
    print(f"wgrad {cin:5d}->{cout:5d} @{H:3d}  {ms:8.3f} ms  {fl / ms / 1e9:8.1f} TFLOP/s")
    del x, dz, dw
print("total", tot)
