"""One eager training step (or eval forward) for an ncu launch-list capture, plus a tag file that names every GEMM / BatchNorm
launch of that step (layer shape, algorithmic FLOPs and bytes) in issue order - tools/per_launch_table.py joins the two.

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 6000 --csv \
        --log-file gpurun_out/launches.csv python tools/ncu_step.py --model unet --batch 32 --tags gpurun_out/tags.json
"""
import argparse
import json
import os
import sys

os.environ["UZ_OVERLAP_WGRAD"] = "0"          # ncu serialises kernels: one stream, issue order = execution order
import torch  # noqa: E402

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import unet_zoo_b200  # noqa: E402
from unet_zoo_b200 import engine as E  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="unet")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--eval", action="store_true")
    ap.add_argument("--tags", default="gpurun_out/tags.json")
    a = ap.parse_args()
    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(a.model).cuda()
    model.eval() if a.eval else model.train()
    x = torch.randn(a.batch, 3, a.size, a.size, device="cuda")
    m = (torch.rand(a.batch, 1, a.size, a.size, device="cuda") > 0.5).float()
    crit = torch.nn.BCEWithLogitsLoss()

    def step():
        if a.eval:
            with torch.no_grad():
                model(x)
            return
        model.zero_grad(set_to_none=True)
        out = model(x)
        outs = list(out.values()) if isinstance(out, dict) else (out if isinstance(out, (list, tuple)) else [out])
        sum(crit(o, m) for o in outs).backward()

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    E.PROFILE = []
    step()
    torch.cuda.synchronize()
    tags = [{"kind": k, "tag": t, "flops": f if k.startswith("igemm") else 0.0, "bytes": nb,
             "ms_events": e0.elapsed_time(e1)} for k, f, e0, e1, t, nb in E.PROFILE]
    E.PROFILE = None
    with open(a.tags, "w") as f:
        json.dump({"model": a.model, "batch": a.batch, "size": a.size, "eval": a.eval, "tags": tags}, f)


if __name__ == "__main__":
    main()
