"""BASELINE config #5: `resunet` (or any hot-path model) inference sweep over batch x resolution on one GPU.

    python tools/bench_infer.py [--model resunet] [--sizes 256,384,512,768,1024] [--batches 1,2,4,8,16,32,64]

Eval-mode forward (BatchNorm folded / running statistics), inputs resident on the device, CUDA-event timing of `iters`
forwards after 3 warm-ups.  Prints one JSON line per (batch, size): images/s, ms, achieved conv TFLOP/s (algorithmic
forward FLOPs of SURVEY.md section 6, scaled by H*W/512^2) and its fraction of the measured sustained bf16 peak.
--graph replays the forward as one CUDA graph (unet_zoo_b200.train_step.GraphedInference) and reports both.
--layers BATCH,SIZE prints the per-layer table of the convolution launches (CUDA events around every launch) as markdown.
Under torchrun (N > 1) every rank runs the same sweep on its own GPU with `batch` images PER GPU (the batch is sharded,
inference needs no communication - "replicas only"); times are the max over ranks and rank 0 prints whole-job img/s:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_infer.py --graph
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FWD_GFLOP_512 = {"unet": 385.34, "attention_unet": 531.40, "nested_unet": 275.88, "resunet": 311.08, "multiresunet": 125.24}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resunet")
    ap.add_argument("--sizes", default="256,384,512,768,1024")
    ap.add_argument("--batches", default="1,2,4,8,16,32,64")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--layers", default="")
    ap.add_argument("--max-pixels", type=float, default=64 * 512 * 512, help="skip (batch, size) above this many pixels")
    a = ap.parse_args()
    import torch
    import unet_zoo_b200
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))

    def over_ranks(ms):
        """Step time of the job = slowest replica (barrier before, max over ranks after)."""
        if dist is None:
            return ms
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def sync():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]
    except Exception:
        peak = 1400.0
    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(a.model).cuda().eval()
    if a.layers:
        from unet_zoo_b200 import engine as E
        b, sz = [int(v) for v in a.layers.split(",")]
        x = torch.randn(b, 3, sz, sz, device="cuda")
        with torch.no_grad():
            for _ in range(3):
                model(x)
            E.PROFILE = []
            model(x)
            torch.cuda.synchronize()
        rows = [(kind, tag or kind, flops if kind.startswith("igemm") else 0.0, e0.elapsed_time(e1), nb)
                for kind, flops, e0, e1, tag, nb in E.PROFILE]
        E.PROFILE = None
        try:
            hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
        except Exception:
            hbm = 6650.0
        tot_ms = sum(r[3] for r in rows)
        print(f"# {a.model} inference, batch {b}, {sz}x{sz}: every profiled launch of one eval-mode forward (BatchNorm folded), "
              "CUDA events around each launch\n")
        print(f"sum {tot_ms:.3f} ms, {sum(r[2] for r in rows) / tot_ms / 1e9:.0f} TFLOP/s; roofs: sustained bf16 peak "
              f"{peak:.0f} TFLOP/s, HBM copy {hbm:.0f} GB/s (MEASURED_PEAKS.json).  A layer is bound by the roof it is "
              "closer to: 3x3 layers with >= 128 channels by the tensor pipe, 1x1 / transposed / 64-channel layers by HBM.\n")
        print("| launch | GFLOP | algorithmic MB | ms | TFLOP/s | of tensor roof | GB/s | of HBM roof |\n|---|---|---|---|---|---|---|---|")
        for kind, tag, flops, ms, nb in rows:
            tf = flops / ms / 1e9
            gbs = nb / ms / 1e6
            print(f"| {tag} | {flops / 1e9:.1f} | {nb / 1e6:.0f} | {ms:.3f} | {tf:.0f} | {tf / peak:.2f} | {gbs:.0f} | {gbs / hbm:.2f} |")
        return
    for s in [int(v) for v in a.sizes.split(",")]:
        for b in [int(v) for v in a.batches.split(",")]:
            if b * s * s > a.max_pixels:
                continue
            x = torch.randn(b, 3, s, s, device="cuda")
            with torch.no_grad():
                for _ in range(3):
                    model(x)
                sync()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(a.iters):
                    model(x)
                e1.record()
                torch.cuda.synchronize()
            ms = over_ranks(e0.elapsed_time(e1) / a.iters)
            ms_graph = None
            if a.graph:
                from unet_zoo_b200.train_step import GraphedInference
                g = GraphedInference(model, x)
                for _ in range(3):
                    g(x)
                sync()
                e0.record()
                for _ in range(a.iters):
                    g(x)
                e1.record()
                torch.cuda.synchronize()
                ms_graph = over_ranks(e0.elapsed_time(e1) / a.iters)
                del g
            tf = b * FWD_GFLOP_512[a.model] * (s / 512.0) ** 2 / ms      # GFLOP / ms = TFLOP/s
            if rank == 0:
                line = {"model": a.model, "mode": "inference", "n_gpus": world, "batch_per_gpu": b, "size": s,
                        "ms": round(ms, 3), "images_per_sec": round(world * b / ms * 1000.0, 1),
                        "conv_tflops_per_gpu": round(tf, 1), "frac_of_sustained_bf16_peak": round(tf / peak, 3),
                        "launches": model._last_launches}
                if ms_graph:
                    line.update(ms_graph=round(ms_graph, 3),
                                images_per_sec_graph=round(world * b / ms_graph * 1000.0, 1))
                print(json.dumps(line), flush=True)
            del x
            torch.cuda.empty_cache()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
