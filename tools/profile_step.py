"""Kernel-time breakdown of one eager training step (torch.profiler / CUPTI): which kernels the step spends its time
in, the library's own and torch's (the stock-torch bottleneck island of transatt_unet).  Not a bench: numbers taken
under the profiler only rank kernels.

    python tools/profile_step.py --model transatt_unet --batch 32 --size 512
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import unet_zoo_b200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="unet")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--rows", type=int, default=30)
    ap.add_argument("--eval", action="store_true", help="profile the eval-mode forward instead of a training step")
    a = ap.parse_args()
    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(a.model).cuda().train()
    x = torch.randn(a.batch, 3, a.size, a.size, device="cuda")
    m = (torch.rand(a.batch, 1, a.size, a.size, device="cuda") > 0.5).float()
    crit = torch.nn.BCEWithLogitsLoss()

    if a.eval:
        model.eval()

    def step():
        if a.eval:
            with torch.no_grad():
                model(x)
            return
        model.zero_grad(set_to_none=True)
        out = model(x)
        outs = list(out.values()) if isinstance(out, dict) else (out if isinstance(out, (list, tuple)) else [out])
        sum(crit(o, m) for o in outs).backward()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
    total = sum(e.device_time_total for e in rows)
    print(f"{a.model} batch {a.batch} {a.size}^2: {total / 1000.0:.2f} ms of kernel time in one "
          f"{'eval forward' if a.eval else 'training step'} (serialised sum)")
    for e in rows[:a.rows]:
        print(f"{e.device_time_total / 1000.0:9.3f} ms {100.0 * e.device_time_total / total:5.1f} %  x{e.count:<4d} {e.key[:110]}")


if __name__ == "__main__":
    main()
