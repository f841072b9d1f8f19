"""uz_stem_im2col at the bench shape (32 x 3 x 512 x 512 -> 32-wide im2col operand): tiled kernel against the
first-generation one-thread-per-pixel kernel (UZ_STEM_GEN1=1).  Algorithmic bytes: 4*cin in + 2*kp out per pixel."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from unet_zoo_b200 import _lib as L  # noqa: E402
from unet_zoo_b200 import engine as E  # noqa: E402


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ctx = E.Ctx(torch.device("cuda", 0), training=True, record=False)
    print("| N x cin x H x W | kp | GB | gen1 ms | tiled ms | GB/s |")
    print("|---|---|---|---|---|---|")
    for (N, cin, H, W) in [(32, 3, 512, 512), (16, 3, 512, 512), (8, 3, 512, 512), (32, 1, 512, 512), (4, 3, 1024, 1024)]:
        kp = 16 if 9 * cin <= 16 else 32 if 9 * cin <= 32 else 64
        x = torch.randn(N, cin, H, W, device="cuda")
        out = torch.empty(N, H, W, kp, dtype=torch.bfloat16, device="cuda")
        gb = N * H * W * (4 * cin + 2 * kp) / 1e9
        res = {}
        for gen1 in (True, False):
            if gen1:
                os.environ["UZ_STEM_GEN1"] = "1"
            else:
                os.environ.pop("UZ_STEM_GEN1", None)
            res[gen1] = timed(lambda: L.check(ctx.lib.uz_stem_im2col(x.data_ptr(), N, cin, H, W, out.data_ptr(), kp, ctx.stream), "stem"))
        print(f"| {N} x {cin} x {H} x {W} | {kp} | {gb:.2f} | {res[True]:.3f} | {res[False]:.3f} | {gb / res[False] * 1e3:.0f} |")


if __name__ == "__main__":
    main()
