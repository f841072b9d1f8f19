"""Per-layer timing of the conv kernels at the unet batch-32 / 512x512 shapes (CUDA events, 5 reps after 2 warm-ups)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unet_zoo_b200 import engine as E  # noqa: E402
from unet_zoo_b200 import _lib as L  # noqa: E402

N = int(os.environ.get("BATCH", "32"))
dev = torch.device("cuda", 0)
ctx = E.Ctx(dev, training=True, record=False)
# (name, mode, H, cin, cout, stats)
LAYERS = [
    ("stem 1x1 32->64 @512", L.CONV1X1, 512, 32, 64, True),
    ("c3 64->64 @512", L.CONV3X3, 512, 64, 64, True),
    ("c3 128->64 @512", L.CONV3X3, 512, 128, 64, True),
    ("dg 64->128 @512", L.CONV3X3_DGRAD, 512, 64, 128, False),
    ("c3 64->128 @256", L.CONV3X3, 256, 64, 128, True),
    ("c3 128->128 @256", L.CONV3X3, 256, 128, 128, True),
    ("c3 256->128 @256", L.CONV3X3, 256, 256, 128, True),
    ("dg 128->256 @256", L.CONV3X3_DGRAD, 256, 128, 256, False),
    ("c3 128->256 @128", L.CONV3X3, 128, 128, 256, True),
    ("c3 256->256 @128", L.CONV3X3, 128, 256, 256, True),
    ("c3 512->256 @128", L.CONV3X3, 128, 512, 256, True),
    ("c3 256->512 @64", L.CONV3X3, 64, 256, 512, True),
    ("c3 512->512 @64", L.CONV3X3, 64, 512, 512, True),
    ("c3 1024->512 @64", L.CONV3X3, 64, 1024, 512, True),
    ("c3 512->1024 @32", L.CONV3X3, 32, 512, 1024, True),
    ("c3 1024->1024 @32", L.CONV3X3, 32, 1024, 1024, True),
]
spec = os.environ.get("SPEC")        # e.g. SPEC=c3,512,32,32 (kind c3|dg|c1, H, cin, cout) with BATCH=8: one custom layer
if spec:
    kind, hh, ci, co = spec.split(",")
    LAYERS = [(f"{kind} {ci}->{co} @{hh}", {"c3": L.CONV3X3, "dg": L.CONV3X3_DGRAD, "c1": L.CONV1X1}[kind], int(hh), int(ci),
               int(co), kind != "dg")]
only = os.environ.get("ONLY")
tot = 0.0
for name, mode, H, cin, cout, stats in LAYERS:
    if only and only not in name:
        continue
    taps = 1 if mode == L.CONV1X1 else 9
    x = E.Act(ctx.new_buf(N, H, H, cin))
    x.buf.data.normal_()
    w = torch.randn(cout, taps, cin, device=dev).to(torch.bfloat16)
    out = E.Act(ctx.new_buf(N, H, H, cout))
    st = ctx.f32(ctx.stats_rows, 2, cout) if stats else None
    bias = torch.randn(cout, device=dev)
    times = []
    for i in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ctx.conv(mode, [x], w, cout, bias, out.ptr, out.ld, H, H, st, L.FLAG_STATS if stats else 0)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            times.append(e0.elapsed_time(e1))
    ms = sum(times) / len(times)
    fl = 2.0 * N * H * H * cin * cout * taps
    tot += ms
    print(f"{name:24s} {ms:8.3f} ms  {fl / ms / 1e9:8.1f} TFLOP/s   io {(N*H*H*(cin+cout)*2)/ms/1e6:7.0f} GB/s")
    del x, out, w
print("total", tot)
