"""Headline benchmark: images/sec of the `unet` 3->1, 512x512, bf16-compute training step, batch 32 per GPU.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference's CPU path (oracle port of it) on the host cores

A step is the reference's training-loop body (unet_zoo/utils/training_loop.py:109-121): zero_grad -> forward ->
BCEWithLogitsLoss -> backward (+ gradient all-reduce when N > 1) -> clip_grad_norm_(1.0) -> AdamW(lr 1e-4, wd 1e-5).
Prints ONE JSON line on rank 0.  `value` = device-resident inputs; `e2e` = same step through the public API with the
batch copied from pinned host memory and the loss read back every step.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# conv + transposed-conv FLOPs (2*MAC) of one fwd+bwd per image at 512x512, SURVEY.md section 6 / 8(d)
FWD_BWD_GFLOP_PER_IMG_512 = {"unet": 1155.11, "attention_unet": 1593.29, "nested_unet": 827.20, "resunet": 931.44,
                              "transatt_unet": 967.47}
UNET_FWD_BWD_GFLOP_PER_IMG_512 = FWD_BWD_GFLOP_PER_IMG_512["unet"]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="images per GPU")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--model", default="unet")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tail", default="fused", choices=["fused", "torch"],
                    help="loss + clip + AdamW: the library's device-side tail (uz_bce_dice / uz_adamw_flat) or torch's")
    ap.add_argument("--no-graph", action="store_true", help="issue the step eagerly instead of replaying one CUDA graph")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"tflops": float(p["bf16_tflops_sustained"]), "hbm": float(p["hbm_gbs"]), "src": "measured"}
    except Exception:
        return {"tflops": 1400.0, "hbm": 6650.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_train_step_rate(model_name, size, steps, warmup, threads=None):
    """img/s of the oracle port of the reference's CPU training step (fp32, batch 1) on this host's cores."""
    import copy

    import torch

    import unet_zoo_b200
    from oracle import reference_port as O
    if not threads:
        # all host cores: torchrun exports OMP_NUM_THREADS=1 to its workers, which would time a single thread
        threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    sd = copy.deepcopy(unet_zoo_b200.create_model(model_name).state_dict())
    leaves = {k: v.clone().requires_grad_(True) for k, v in sd.items()
              if v.dtype.is_floating_point and "running_" not in k}
    work = dict(sd)
    work.update(leaves)
    params = list(leaves.values())
    opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=1e-5)
    x, m = O.synthetic_batch(1, 3, size, size)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad()
        loss = O.bce_with_logits(O.FORWARDS[model_name](work, x, True), m)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        float(loss)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return 1.0 / (sum(times) / len(times)), torch.get_num_threads()


def metric_name(model):
    return "images/sec UNet 512^2 bf16 train step" if model == "unet" else f"images/sec {model} bf16 train step"


def workload_name(model, B, S):
    cfg = {"unet": "BASELINE.json configs[1]", "attention_unet": "BASELINE.json configs[2]",
           "nested_unet": "BASELINE.json configs[3]"}.get(model, "not a BASELINE.json config")
    return f"{model} in=3 classes=1, batch {B}/GPU, {S}x{S}, fwd+bwd+clip+AdamW training step ({cfg})"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rate, cores = cpu_train_step_rate(args.model, args.size, args.steps, max(1, min(args.warmup, 2)))
    sample = f"batch 1 of the {args.size}x{args.size} fp32 training step per step (oracle port of the reference path, torch CPU)"
    line = {
        "impl": "reference", "metric": metric_name(args.model), "value": rate, "unit": "images/sec",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 / rate,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.model, args.batch, args.size),
                   "sample": "each step = batch 1 of that workload on the host cores, fp32 (the reference has no bf16 path)"},
        "cpu_baseline": {"value": rate, "unit": "images/sec", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "images/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    import unet_zoo_b200
    from unet_zoo_b200 import engine as E
    from unet_zoo_b200 import parallel as P

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the single JSON line (NCCL prints its version there)
        dist.init_process_group("nccl", device_id=dev)

    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(args.model, in_channels=3, num_classes=1).to(dev).train()
    if world > 1:
        model = P.DataParallel(model, bucket_mb=float(os.environ.get("UZ_BUCKET_MB", "25")))   # bucketed NCCL all-reduce overlapped with backward
    params = [p for p in model.parameters() if p.requires_grad]
    # N > 1: capturing the bucketed NCCL all-reduces of the communication stream into the step graph was tried
    # (capture_error_mode="thread_local") and hung at N = 2.  Default there: two graphs (fwd+bwd | clip+AdamW) around ONE
    # eager all-reduce of the flat gradient buffer (DataParallel.defer_sync); UZ_DP_GRAPH=0 or --no-graph issue the step
    # eagerly with the bucketed all-reduces overlapped with backward (host enqueue 24-36 ms of a 43-47 ms step)
    use_graph = not args.no_graph and (world == 1 or os.environ.get("UZ_DP_GRAPH", "1") != "0")
    if world > 1 and use_graph:
        model.defer_sync = True
    if args.tail == "fused":
        # BCE-with-logits + Dice + loss gradient in one pass, clip_grad_norm_(1.0) + AdamW as one flat update
        # (training_loop.py:115-121 incl. the per-batch Dice of :81 that the torch tail below leaves out)
        from unet_zoo_b200.loss import BCEDiceWithLogits
        from unet_zoo_b200.optim import FlatAdamW
        opt = FlatAdamW(params, lr=1e-4, weight_decay=1e-5, max_norm=1.0)
        crit = BCEDiceWithLogits()
        clip, tail_launches = None, 5
    else:
        opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=1e-5, fused=True, capturable=use_graph)
        crit = torch.nn.BCEWithLogitsLoss()
        clip, tail_launches = 1.0, 0

    B, S = args.batch, args.size
    gx = torch.Generator().manual_seed(1234 + rank)
    gm = torch.Generator().manual_seed(4321 + rank)
    x_h = torch.randn(B, 3, S, S, generator=gx).pin_memory()
    m_h = (torch.rand(B, 1, S, S, generator=gm) > 0.5).float().pin_memory()
    x_d, m_d = x_h.to(dev), m_h.to(dev)

    def eager_step(x, m):
        opt.zero_grad(set_to_none=True)
        out = model(x)
        loss = crit(out, m)
        loss.backward()
        if getattr(model, "defer_sync", False):
            model.sync_gradients()
        if clip is not None:
            torch.nn.utils.clip_grad_norm_(params, clip)
        opt.step()
        return loss

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for _ in range(k):
            fn()
        e1.record()
        timed.host_ms = (time.perf_counter() - t0) * 1000.0 / k   # host time to ENQUEUE one step (launch-bound check)
        sync()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    step, graph_note = eager_step, "eager"
    if use_graph:
        from unet_zoo_b200.train_step import GraphedTrainStep
        try:
            if world > 1:
                step = GraphedTrainStep(model, crit, opt, x_d, m_d, max_norm=clip, params=params,
                                        grad_sync=model.sync_gradients)
                graph_note = ("two CUDA graphs per step (fwd+bwd | clip+AdamW) around one eager NCCL all-reduce of the "
                              "flat gradient buffer (unet_zoo_b200.train_step.GraphedTrainStep, grad_sync)")
            else:
                step = GraphedTrainStep(model, crit, opt, x_d, m_d, max_norm=clip, params=params)
                graph_note = "one CUDA graph per step (unet_zoo_b200.train_step.GraphedTrainStep)"
        except Exception as e:   # report, never hide: the JSON line says which mode was measured
            graph_note = f"eager (graph capture failed: {type(e).__name__}: {e})"[:300]
            if world > 1:
                model.defer_sync = False
            torch.cuda.synchronize()
    for _ in range(max(args.warmup, 3)):
        step(x_d, m_d)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ms = timed(lambda: step(x_d, m_d), args.steps)
    host_ms = timed.host_ms
    base = model.module if hasattr(model, "module") else model
    launches = (base._last_launches + base._last_launches_bwd + tail_launches) * args.steps

    # end to end: every step copies ITS batch from pinned host memory (on a copy stream, double-buffered so the copy of
    # batch i+1 overlaps the compute of batch i - what a prefetching input pipeline does) and reads the loss back
    copy_stream = torch.cuda.Stream(device=dev)
    bufs = [(torch.empty_like(x_d), torch.empty_like(m_d)) for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    state = {"i": 0}

    def prefetch(i):
        with torch.cuda.stream(copy_stream):
            bufs[i % 2][0].copy_(x_h, non_blocking=True)
            bufs[i % 2][1].copy_(m_h, non_blocking=True)
            ready[i % 2].record(copy_stream)

    def e2e_step():
        i = state["i"]
        torch.cuda.current_stream().wait_event(ready[i % 2])
        prefetch(i + 1)            # the other buffer is free: the previous step ended with a host read of its loss
        v = float(step(*bufs[i % 2]).item())
        state["i"] = i + 1
        return v
    prefetch(0)
    e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    clocks = sampler.stop() if sampler else None

    # roofline leg: per-launch CUDA-event timing of the tensor-core kernels during three more steps (averaged: a single
    # eager step right after the e2e leg read 0.82-0.96 of peak for the same kernels from run to run)
    # (kernels timed one at a time: the side-stream overlap of the weight gradients is switched off for these steps)
    overlap_env = os.environ.get("UZ_OVERLAP_WGRAD")
    os.environ["UZ_OVERLAP_WGRAD"] = "0"
    eager_step(x_d, m_d)                  # untimed: first eager launch sequence after the graph replays
    torch.cuda.synchronize()
    PROFILE_STEPS = 3
    E.PROFILE = []
    for _ in range(PROFILE_STEPS):
        eager_step(x_d, m_d)
    torch.cuda.synchronize()
    if overlap_env is None:
        del os.environ["UZ_OVERLAP_WGRAD"]
    else:
        os.environ["UZ_OVERLAP_WGRAD"] = overlap_env
    agg = {}
    for kind, flops, a, b, _tag in E.PROFILE:
        t, f, n = agg.get(kind, (0.0, 0.0, 0))
        agg[kind] = (t + a.elapsed_time(b), f + flops, n + 1)
    E.PROFILE = None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    imgs = B * world * args.steps
    value = imgs / (ms / 1000.0)
    top = max((k for k in agg if k.startswith("igemm")), key=lambda k: agg[k][0]) if agg else None
    roof = None
    traffic = None
    try:   # DRAM bytes per launch of the family, from the committed ncu capture (profiles/r01_traffic.json)
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            traffic = json.load(f)[top]["dram_bytes_per_launch"] if args.model == "unet" and B == 32 and S == 512 else None
    except Exception:
        traffic = None
    if top:
        t, f, n = agg[top]
        ach = f / (t / 1000.0) / 1e12
        roof = {"kernel": top, "bound": "tensor", "achieved": ach, "peak": pk["tflops"], "unit": "TFLOP/s",
                "frac": ach / pk["tflops"], "traffic": traffic, "peak_source": pk["src"] + " (sustained bf16)",
                "launches": n // PROFILE_STEPS, "avg_launch_ms": t / n, "steps_averaged": PROFILE_STEPS,
                "families": {k: ({"ms": v[0] / PROFILE_STEPS, "tflops": v[1] / (v[0] / 1000.0) / 1e12,
                                  "launches": v[2] // PROFILE_STEPS}
                                 if k.startswith("igemm") else
                                 {"ms": v[0] / PROFILE_STEPS, "gbs": v[1] / (v[0] / 1000.0) / 1e9,
                                  "frac_hbm": v[1] / (v[0] / 1000.0) / 1e9 / pk["hbm"], "launches": v[2] // PROFILE_STEPS})
                             for k, v in agg.items()},
                "step_tflops": value / world * FWD_BWD_GFLOP_PER_IMG_512[args.model] * (S / 512.0) ** 2 / 1000.0}
    line = {
        "metric": metric_name(args.model), "value": value, "unit": "images/sec", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(args.model, B, S), "global_batch": B * world,
                   "parallelism": f"dp{world}", "launch": graph_note,
                   "tail": ("uz_bce_dice + uz_adamw_flat (loss, Dice, clip, AdamW on the device)" if args.tail == "fused"
                            else "torch BCEWithLogitsLoss + clip_grad_norm_ + fused AdamW"), "l2": "inputs larger than L2 (multi-GB activations per step)"},
        "e2e": {"value": imgs / (ms_e2e / 1000.0), "unit": "images/sec",
                "h2d_bytes_per_step": x_h.numel() * 4 + m_h.numel() * 4, "d2h_bytes_per_step": 4},
        "gpu_launches": launches, "host_enqueue_ms_per_step": host_ms, "roofline": roof, "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        rate, cores = cpu_train_step_rate(args.model, S, 3, 1)
        line["cpu_baseline"] = {"value": rate, "unit": "images/sec", "cores": cores, "kind": "port",
                                "sample": f"3 timed fp32 training steps at batch 1, {S}x{S} (oracle port, torch CPU)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
