"""BASELINE configs[0]: `unet` in=3 classes=1, batch 1, 256x256 fp32 eval forward (the reference's README example) -
the fp32 parity mode (csrc/f32.cu) on one B200 next to the oracle port on the box's host cores.  One JSON line.

    python tools/bench_fp32.py [--model unet] [--batch 1] [--size 256] [--steps 20]
"""
import argparse
import copy
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import unet_zoo_b200  # noqa: E402
from oracle import reference_port as O  # noqa: E402  (CPU baseline leg + parity check only)

FWD_GFLOP_512 = {"unet": 385.34, "attention_unet": 531.40, "nested_unet": 275.88, "resunet": 311.08}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="unet")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(a.model, precision="fp32")
    sd = copy.deepcopy(model.state_dict())
    x = torch.randn(a.batch, 3, a.size, a.size, generator=torch.Generator().manual_seed(1234))
    # CPU: the oracle port (same ATen kernels as the reference modules), best of 5 after 2 warm-ups
    cpu_t = []
    with torch.no_grad():
        for i in range(7):
            t0 = time.perf_counter()
            ref = O.FORWARDS[a.model](sd, x, False)
            cpu_t.append(time.perf_counter() - t0)
    cpu_ms = min(cpu_t[2:]) * 1e3
    model.cuda().eval()
    xg = x.cuda()
    with torch.no_grad():
        for _ in range(a.warmup):
            out = model(xg)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            out = model(xg)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    outs = out if isinstance(out, (list, tuple)) else [out]
    refs = ref if isinstance(ref, (list, tuple)) else [ref]
    err = max(float((o.cpu().double() - r.double()).abs().max() / r.double().abs().max()) for o, r in zip(outs, refs))
    gf = FWD_GFLOP_512[a.model] * a.batch * (a.size / 512.0) ** 2
    print(json.dumps({"workload": f"{a.model} batch {a.batch} {a.size}x{a.size} fp32 eval forward (BASELINE configs[0])",
                      "gpu_ms": ms, "gpu_images_per_sec": a.batch / ms * 1e3, "gpu_fp32_tflops": gf / ms,
                      "launches": model._last_launches, "rel_err_vs_oracle": err,
                      "cpu_ms": cpu_ms, "cpu_images_per_sec": a.batch / cpu_ms * 1e3,
                      "cpu_threads": torch.get_num_threads()}))


if __name__ == "__main__":
    main()
