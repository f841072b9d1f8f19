"""Host-side cost of ENQUEUEING one eager training step (cProfile over a few steps): which Python / ctypes calls the
~300 launches of a step spend their host time in.  The GPU runs ahead asynchronously; only the enqueue time matters."""
import argparse
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import unet_zoo_b200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="unet")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--rows", type=int, default=35)
    a = ap.parse_args()
    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(a.model).cuda().train()
    x = torch.randn(a.batch, 3, a.size, a.size, device="cuda")
    m = (torch.rand(a.batch, 1, a.size, a.size, device="cuda") > 0.5).float()
    crit = torch.nn.BCEWithLogitsLoss()

    def step():
        model.zero_grad(set_to_none=True)
        crit(model(x), m).backward()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        step()
    host = (time.perf_counter() - t0) / 5
    torch.cuda.synchronize()
    total = (time.perf_counter() - t0) / 5
    print(f"{a.model} batch {a.batch} {a.size}^2: host enqueue {host * 1e3:.2f} ms per step, wall {total * 1e3:.2f} ms per step")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(3):
        step()
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(a.rows)


if __name__ == "__main__":
    main()
