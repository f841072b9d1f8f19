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
                              "transatt_unet": 967.47, "unet_transformer": 1830.0,   # last: 3 x the 610 GFLOP forward
                              "u2net": 900.07, "u2netp": 303.45,       # 3 x the conv FLOPs of one forward (hooks on the reference)
                              "multiresunet": 375.72}                  # ditto, on the reference's (unpadded) channel counts
FWD_GFLOP_PER_IMG_512 = {"unet": 385.34, "attention_unet": 531.40, "nested_unet": 275.88, "resunet": 311.08}
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
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the secondary legs (BASELINE configs[2..4] and the stock torch + cuDNN comparator)")
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


def _families(profile, pk, steps):
    """Per kernel family of one profiled step: ms, TFLOP/s and fraction of the tensor roof (GEMM kinds), GB/s of
    ALGORITHMIC bytes and fraction of the HBM roof (all kinds)."""
    agg = {}
    for kind, flops, a, b, _tag, nbytes in profile:
        t, f, by, n = agg.get(kind, (0.0, 0.0, 0.0, 0))
        agg[kind] = (t + a.elapsed_time(b), f + flops, by + nbytes, n + 1)
    out = {}
    for k, (t, f, by, n) in agg.items():
        fam = {"ms": t / steps, "launches": n // steps, "gbs": by / (t / 1000.0) / 1e9}
        fam["frac_hbm"] = fam["gbs"] / pk["hbm"]
        if k.startswith("igemm"):
            fam["tflops"] = f / (t / 1000.0) / 1e12
            fam["frac_tensor"] = fam["tflops"] / pk["tflops"]
        out[k] = fam
    return out, agg


def train_leg(args, model_name, B, S, steps, warmup, dev, world, rank, e2e=True, profile_steps=3, sampler=None):
    """Times `steps` training steps of `model_name` at batch B per GPU (device-resident inputs), optionally the same
    through the public API with host batches, and a per-launch profile of `profile_steps` extra eager steps."""
    import gc

    import torch
    import torch.distributed as dist

    import unet_zoo_b200
    from unet_zoo_b200 import engine as E
    from unet_zoo_b200 import parallel as P

    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(model_name, in_channels=3, num_classes=1).to(dev).train()
    # UZ_DP_MODE=replicas: N independent single-GPU steps with NO gradient exchange (not a training mode - it measures what
    # the GPUs of the box do on their own, so that max-over-ranks power-cap spread can be told from exposed communication)
    replicas = world > 1 and os.environ.get("UZ_DP_MODE", "") == "replicas"
    if world > 1 and not replicas:
        model = P.DataParallel(model, bucket_mb=float(os.environ.get("UZ_BUCKET_MB", "25")))
    params = [p for p in model.parameters() if p.requires_grad]
    # N > 1 launch modes (UZ_DP_MODE): "overlap" (default) = ONE CUDA graph per step with the bucketed NCCL all-reduces
    # captured on the communication stream (they overlap the rest of backward); "split" = two graphs around one eager
    # all-reduce of the flat gradient buffer (round 1's default); "eager" = eager launches + bucketed all-reduces
    dp_mode = os.environ.get("UZ_DP_MODE", "overlap") if world > 1 else "single"
    if replicas:
        dp_mode = "single"
    if args.no_graph:
        dp_mode = "eager" if world > 1 else "single"
    use_graph = not args.no_graph
    if dp_mode == "split":
        model.defer_sync = True
    if args.tail == "fused":
        from unet_zoo_b200.loss import BCEDiceWithLogits
        from unet_zoo_b200.optim import FlatAdamW
        opt = FlatAdamW(params, lr=1e-4, weight_decay=1e-5, max_norm=1.0)
        crit = BCEDiceWithLogits()
        clip, tail_launches = None, 5
    else:
        opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=1e-5, fused=True, capturable=use_graph)
        crit = torch.nn.BCEWithLogitsLoss()
        clip, tail_launches = 1.0, 0

    gx = torch.Generator().manual_seed(1234 + rank)
    gm = torch.Generator().manual_seed(4321 + rank)
    x_h = torch.randn(B, 3, S, S, generator=gx).pin_memory()
    m_h = (torch.rand(B, 1, S, S, generator=gm) > 0.5).float().pin_memory()
    x_d, m_d = x_h.to(dev), m_h.to(dev)

    def eager_step(x, m):
        opt.zero_grad(set_to_none=True)
        out = model(x)
        if isinstance(out, dict):      # U2NET: seven maps, weight 1.0 each (training_loop.py:23-31, :61-64)
            loss = sum(crit(o, m) for o in out.values())
        else:
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
        mine = torch.tensor([e0.elapsed_time(e1)], device=dev)
        timed.per_rank = [float(mine)]
        if world > 1:
            allr = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allr, mine)
            timed.per_rank = [float(t) for t in allr]
        return max(timed.per_rank)

    step, graph_note = eager_step, "eager launches" + (", bucketed NCCL all-reduce overlapped with backward" if world > 1 else "")
    if use_graph and dp_mode != "eager":
        from unet_zoo_b200.train_step import GraphedTrainStep
        try:
            if dp_mode == "split":
                step = GraphedTrainStep(model, crit, opt, x_d, m_d, max_norm=clip, params=params,
                                        grad_sync=model.sync_gradients)
                graph_note = ("two CUDA graphs per step (fwd+bwd | clip+AdamW) around one eager NCCL all-reduce of the "
                              "flat gradient buffer (GraphedTrainStep, grad_sync)")
            elif dp_mode == "overlap":
                step = GraphedTrainStep(model, crit, opt, x_d, m_d, max_norm=clip, params=params, capture_nccl=True)
                graph_note = ("one CUDA graph per step; the bucketed NCCL all-reduces (%s MB buckets) are captured on the "
                              "communication stream and overlap the rest of backward" % os.environ.get("UZ_BUCKET_MB", "25"))
            else:
                step = GraphedTrainStep(model, crit, opt, x_d, m_d, max_norm=clip, params=params)
                graph_note = "one CUDA graph per step (unet_zoo_b200.train_step.GraphedTrainStep)"
                if replicas:
                    graph_note += "; UZ_DP_MODE=replicas: independent replicas, NO gradient exchange (diagnostic)"
        except Exception as e:   # report, never hide: the JSON line says which mode was measured
            graph_note = f"eager (graph capture failed: {type(e).__name__}: {e})"[:300]
            if world > 1:
                model.defer_sync = False
            torch.cuda.synchronize()
    for _ in range(max(warmup, 3)):
        step(x_d, m_d)
    if sampler:
        sampler.start()
    ms = timed(lambda: step(x_d, m_d), steps)
    res = {"ms_per_step": ms / steps, "host_ms": timed.host_ms, "rank_ms_per_step": [t / steps for t in timed.per_rank],
           "launch": graph_note, "B": B, "S": S}
    base = model.module if hasattr(model, "module") else model
    res["launches_per_step"] = base._last_launches + base._last_launches_bwd + tail_launches

    if e2e:
        # end to end: every step copies ITS batch from pinned host memory (on a copy stream, double-buffered so the copy
        # of batch i+1 overlaps the compute of batch i - what a prefetching input pipeline does) and reads the loss back
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
            prefetch(i + 1)        # the other buffer is free: the previous step ended with a host read of its loss
            v = float(step(*bufs[i % 2]).item())
            state["i"] = i + 1
            return v
        prefetch(0)
        e2e_step()
        res["e2e_ms_per_step"] = timed(e2e_step, steps) / steps
        res["h2d_bytes"] = x_h.numel() * 4 + m_h.numel() * 4
    if sampler:
        res["clocks"] = sampler.stop()

    # roofline leg: per-launch CUDA-event timing of the kernels during extra eager steps (kernels timed one at a time:
    # the side-stream overlap of the weight gradients is switched off for these steps)
    if profile_steps:
        overlap_env = os.environ.get("UZ_OVERLAP_WGRAD")
        os.environ["UZ_OVERLAP_WGRAD"] = "0"
        eager_step(x_d, m_d)              # untimed: first eager launch sequence after the graph replays
        torch.cuda.synchronize()
        E.PROFILE = []
        for _ in range(profile_steps):
            eager_step(x_d, m_d)
        torch.cuda.synchronize()
        if overlap_env is None:
            del os.environ["UZ_OVERLAP_WGRAD"]
        else:
            os.environ["UZ_OVERLAP_WGRAD"] = overlap_env
        res["profile"] = E.PROFILE
        E.PROFILE = None
    del step, model, opt, crit, params, x_d, m_d
    gc.collect()
    torch.cuda.empty_cache()
    return res


def infer_leg(model_name, B, S, iters, dev, pk, world=1):
    """Eval-mode forward (BatchNorm folded) of `model_name` replayed as one CUDA graph; per-family roofline of one eager
    forward.  N > 1: independent replicas, max over ranks."""
    import gc

    import torch
    import torch.distributed as dist

    import unet_zoo_b200
    from unet_zoo_b200 import engine as E
    from unet_zoo_b200.train_step import GraphedInference
    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(model_name).to(dev).eval()
    x = torch.randn(B, 3, S, S, device=dev)
    g = GraphedInference(model, x)
    for _ in range(3):
        g(x)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g(x)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
    with torch.no_grad():
        model(x)
        E.PROFILE = []
        model(x)
        torch.cuda.synchronize()
    fam, _ = _families(E.PROFILE, pk, 1)
    E.PROFILE = None
    tf = B * FWD_GFLOP_PER_IMG_512[model_name] * (S / 512.0) ** 2 / ms
    out = {"workload": f"{model_name} inference (eval, BatchNorm folded), batch {B}/GPU, {S}x{S}, one CUDA graph per forward "
                       "(BASELINE.json configs[4] corner)", "n_gpus": world, "ms": ms,
           "images_per_sec": world * B * 1000.0 / ms, "conv_tflops_per_gpu": tf, "frac_tensor": tf / pk["tflops"],
           "launches": model._last_launches, "families": fam}
    del g, model, x
    gc.collect()
    torch.cuda.empty_cache()
    return out


def library_leg(B, S, dev, steps=3, warmup=2):
    """The GPU comparator that matters: the same UNet written with stock torch.nn modules, run by stock PyTorch + cuDNN on
    this GPU (autocast bf16, channels_last, fused AdamW, clip) - what the reference gives a B200 user without this
    library.  tools/bench_stock_torch.py holds the module; none of the library's kernels run here."""
    import gc

    import torch
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from bench_stock_torch import StockUNet
    torch.manual_seed(0)
    torch.backends.cudnn.benchmark = True
    model = StockUNet().to(dev).to(memory_format=torch.channels_last).train()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-5, fused=True)
    crit = torch.nn.BCEWithLogitsLoss()
    x = torch.randn(B, 3, S, S, device=dev).contiguous(memory_format=torch.channels_last)
    y = (torch.rand(B, 1, S, S, device=dev) > 0.5).float()

    def step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
        loss = crit(out.float(), y)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    out = {"impl": "stock torch.nn + cuDNN %s, autocast bf16, channels_last, fused AdamW (tools/bench_stock_torch.py)"
                   % torch.backends.cudnn.version(), "workload": f"unet batch {B}, {S}x{S} training step",
           "ms_per_step": ms, "value": B * 1000.0 / ms, "unit": "images/sec", "steps": steps}
    del model, opt, x, y
    gc.collect()
    torch.cuda.empty_cache()
    return out


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the single JSON line (NCCL prints its version there)
        # NCCL calls are captured into the step graph: the process group's watchdog must not poll events meanwhile
        os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")
        dist.init_process_group("nccl", device_id=dev)

    pk = peaks()
    B, S = args.batch, args.size
    sampler = ClockSampler(local) if rank == 0 else None
    r = train_leg(args, args.model, B, S, args.steps, args.warmup, dev, world, rank, e2e=True, profile_steps=3,
                  sampler=sampler)
    secondary = None
    if not args.no_secondary and args.model == "unet":
        secondary = {}
        if world == 1:
            # BASELINE configs[2], [3], then the neighbouring models of SURVEY 8(f) rank 4 (not BASELINE configs: batch 32)
            for name, b in (("attention_unet", 16), ("nested_unet", 8), ("transatt_unet", 32), ("unet_transformer", 32),
                            ("u2net", 32), ("u2netp", 32), ("multiresunet", 32)):
                try:
                    q = train_leg(args, name, b, S, 5, 3, dev, world, rank, e2e=False, profile_steps=1)
                    fam, _ = _families(q["profile"], pk, 1)
                    v = b * 1000.0 / q["ms_per_step"]
                    tf = v * FWD_BWD_GFLOP_PER_IMG_512[name] * (S / 512.0) ** 2 / 1000.0
                    if name == "attention_unet":
                        # its four nearest-x2 convs run in phase form (16 instead of 36 tap-pixels): the families' TFLOP/s
                        # count EXECUTED FLOPs, step_tflops the reference's algorithmic FLOPs (what img/s is worth)
                        fam["note"] = "families: executed FLOPs (up-convs in phase form); step_tflops: reference FLOPs"
                    secondary[name] = {"workload": workload_name(name, b, S), "value": v, "unit": "images/sec",
                                       "ms_per_step": q["ms_per_step"], "steps": 5, "step_tflops": tf,
                                       "frac_tensor": tf / pk["tflops"], "launch": q["launch"], "families": fam}
                except Exception as e:
                    secondary[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
            corners = ((64, 256), (32, 512), (8, 1024))
        else:
            corners = ((32, 512),)
        inf = []
        for b, s in corners:
            try:
                inf.append(infer_leg("resunet", b, s, 10, dev, pk, world))
            except Exception as e:
                inf.append({"workload": f"resunet inference batch {b} {s}x{s}", "error": f"{type(e).__name__}: {e}"[:300]})
        secondary["resunet_inference"] = inf

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    steps = args.steps
    ms = r["ms_per_step"]
    value = B * world * 1000.0 / ms
    fam, agg = _families(r["profile"], pk, 3)
    top = max((k for k in agg if k.startswith("igemm")), key=lambda k: agg[k][0]) if agg else None
    roof = None
    traffic = None
    try:   # DRAM bytes per launch of the family, from the committed ncu capture (profiles/r01_traffic.json)
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            traffic = json.load(f)[top]["dram_bytes_per_launch"] if args.model == "unet" and B == 32 and S == 512 else None
    except Exception:
        traffic = None
    if top:
        t, f, _by, n = agg[top]
        ach = f / (t / 1000.0) / 1e12
        roof = {"kernel": top, "bound": "tensor", "achieved": ach, "peak": pk["tflops"], "unit": "TFLOP/s",
                "frac": ach / pk["tflops"], "traffic": traffic, "peak_source": pk["src"] + " (sustained bf16)",
                "launches": n // 3, "avg_launch_ms": t / n, "steps_averaged": 3, "families": fam,
                "step_tflops": value / world * FWD_BWD_GFLOP_PER_IMG_512[args.model] * (S / 512.0) ** 2 / 1000.0}
    rk = sorted(r["rank_ms_per_step"])
    line = {
        "metric": metric_name(args.model), "value": value, "unit": "images/sec", "n_gpus": world,
        "steps": steps, "warmup": max(args.warmup, 3), "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(args.model, B, S), "global_batch": B * world,
                   "parallelism": f"dp{world}", "launch": r["launch"],
                   "tail": ("uz_bce_dice + uz_adamw_flat (loss, Dice, clip, AdamW on the device)" if args.tail == "fused"
                            else "torch BCEWithLogitsLoss + clip_grad_norm_ + fused AdamW"), "l2": "inputs larger than L2 (multi-GB activations per step)"},
        "e2e": {"value": B * world * 1000.0 / r["e2e_ms_per_step"], "unit": "images/sec",
                "h2d_bytes_per_step": r["h2d_bytes"], "d2h_bytes_per_step": 4},
        "gpu_launches": r["launches_per_step"] * steps, "host_enqueue_ms_per_step": r["host_ms"],
        # each rank's own device-timed step: the spread separates power-cap differences between GPUs from exposed comms
        "rank_ms_per_step": {"min": rk[0], "median": rk[len(rk) // 2], "max": rk[-1]},
        "roofline": roof, "clocks": r.get("clocks"),
    }
    if world == 1 and not args.no_cpu_baseline:
        rate, cores = cpu_train_step_rate(args.model, S, 3, 1)
        line["cpu_baseline"] = {"value": rate, "unit": "images/sec", "cores": cores, "kind": "port",
                                "sample": f"3 timed fp32 training steps at batch 1, {S}x{S} (oracle port, torch CPU)"}
    if secondary is not None:
        line["secondary"] = secondary
        if world == 1:
            try:
                line["gpu_library_baseline"] = library_leg(B, S, dev)
                line["gpu_library_baseline"]["speedup_of_this_path"] = value / line["gpu_library_baseline"]["value"]
            except Exception as e:
                line["gpu_library_baseline"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
