"""uz_bilinear2x_fwd / _bwd at the level shapes of nested_unet / transatt_unet: achieved HBM GB/s of the
row-per-block kernels against the first-generation element-per-thread kernels (UZ_BILINEAR_GEN1=1).

Algorithmic bytes: forward reads the low-res tensor once and writes the high-res one (C*2*(1/4 + 1) per high-res
pixel); backward reads the high-res gradient and writes the low-res one (same bytes)."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from unet_zoo_b200 import _lib as L  # noqa: E402
from unet_zoo_b200 import engine as E  # noqa: E402

SHAPES = [(32, 256, 256, 64), (32, 128, 128, 128), (32, 64, 64, 256), (32, 32, 32, 512), (8, 256, 256, 64), (8, 256, 256, 32)]


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
    lib, st = ctx.lib, ctx.stream
    print("| N x H x W x C (low-res) | GB | fwd gen1 ms | fwd rows ms | GB/s | bwd gen1 ms | bwd rows ms | GB/s |")
    print("|---|---|---|---|---|---|---|---|")
    for (N, H, W, C) in SHAPES:
        lo = ctx.new_act(N, H, W, C)
        hi = E.Act(ctx.new_buf(N, 2 * H, 2 * W, 2 * C), C, C)      # a slice of a concat buffer, as in the models
        lo.buf.data.normal_()
        hi.buf.data.normal_()
        dlo = torch.empty_like(lo.buf.data)
        gb = N * H * W * C * 2 * 5 / 1e9
        res = {}
        for gen1 in (True, False):
            if gen1:
                os.environ["UZ_BILINEAR_GEN1"] = "1"
            else:
                os.environ.pop("UZ_BILINEAR_GEN1", None)
            f = timed(lambda: L.check(lib.uz_bilinear2x_fwd(lo.ptr, lo.ld, N, H, W, C, hi.ptr, hi.ld, st), "fwd"))
            b = timed(lambda: L.check(lib.uz_bilinear2x_bwd(hi.ptr, hi.ld, N, H, W, C, dlo.data_ptr(), C, 0, st), "bwd"))
            res[gen1] = (f, b)
        print(f"| {N} x {H} x {W} x {C} | {gb:.2f} | {res[True][0]:.3f} | {res[False][0]:.3f} | {gb / res[False][0] * 1e3:.0f} "
              f"| {res[True][1]:.3f} | {res[False][1]:.3f} | {gb / res[False][1] * 1e3:.0f} |")


if __name__ == "__main__":
    main()
