"""Join an ncu launch list (gpu__time_duration.sum, dram__bytes_read.sum, dram__bytes_write.sum per launch) of
tools/ncu_step.py with its tag file: per GEMM / BatchNorm launch of the LAST step the layer, time, TFLOP/s and fraction
of the tensor roof, algorithmic bytes, measured DRAM bytes, their ratio and the achieved fraction of the HBM roof.

    python tools/per_launch_table.py gpurun_out/launches.csv gpurun_out/tags.json > profiles/<name>.md
"""
import csv
import json
import os
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3,
        "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# kernel-name classes per tag kind: how many launches of which kernels one tagged call issues, in order
CLASS = {"igemm_fwd": ("igemm_fwd",), "igemm_wgrad": ("igemm_wgrad",), "bn_apply": ("bn_apply", "bn_head_fwd", "dual_bn_fwd"),
         "bn_bwd": ("bn_bwd_reduce", "bn_head_bwd_reduce")}      # "dual_bn_bwd_reduce" contains "bn_bwd_reduce"


def main():
    path, tagpath = sys.argv[1], sys.argv[2]
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        tens, hbm = pk["bf16_tflops_sustained"], pk["hbm_gbs"]
    except Exception:
        tens, hbm = 1400.0, 6650.0
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    per, order = {}, []
    for r in csv.DictReader(lines):
        i = r["ID"]
        if i not in per:
            per[i] = {"name": r["Kernel Name"], "ms": 0.0, "rd": 0.0, "wr": 0.0}
            order.append(i)
        v = float(r["Metric Value"].replace(",", "")) * UNIT.get(r.get("Metric Unit", ""), 1.0)
        mn = r.get("Metric Name")
        if mn == "gpu__time_duration.sum":
            per[i]["ms"] = v
        elif mn == "dram__bytes_read.sum":
            per[i]["rd"] = v
        elif mn == "dram__bytes_write.sum":
            per[i]["wr"] = v
    rows = [per[i] for i in order]
    meta = json.load(open(tagpath))
    marker = "stem_im2col"
    idx = [i for i, r in enumerate(rows) if marker in r["name"]]
    step = rows[idx[-1]:] if idx else rows
    if not idx:      # multiresunet: no im2col stem; uz_nchw_to_nhwc opens the forward and (output gradient) the backward
        idx = [i for i, r in enumerate(rows) if "nchw_to_nhwc_kernel" in r["name"]]
        step = rows[idx[-2]:] if len(idx) >= 2 else rows
    total = sum(r["ms"] for r in step)
    print(f"# {meta['model']} batch {meta['batch']} {meta['size']}^2 {'eval forward' if meta['eval'] else 'training step'}: "
          f"per-launch table (ncu, serialised, cold cache: read ratios and shares, not absolute times)\n")
    print(f"step under ncu: {total:.2f} ms in {len(step)} launches; DRAM read {sum(r['rd'] for r in step) / 1e9:.1f} GB, "
          f"written {sum(r['wr'] for r in step) / 1e9:.1f} GB; roofs: {tens:.0f} TFLOP/s (sustained bf16), {hbm:.0f} GB/s\n")
    cursor = {k: 0 for k in CLASS}
    pos = {k: [i for i, r in enumerate(step) if any(c in r["name"] for c in CLASS[k])] for k in CLASS}
    print("| kind | layer | ms | TFLOP/s | of tensor roof | algorithmic MB | DRAM MB (rd + wr) | DRAM / alg | alg GB/s | of HBM roof |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    fam = {}
    for t in meta["tags"]:
        k = t["kind"]
        if k not in pos or cursor[k] >= len(pos[k]):
            continue
        i = pos[k][cursor[k]]
        cursor[k] += 1
        r = step[i]
        ms, rd, wr = r["ms"], r["rd"], r["wr"]
        if k in ("igemm_wgrad", "bn_bwd"):     # multi-kernel calls: add the launches up to the next call of any class
            j = i + 1
            while j < len(step) and not any(any(c in step[j]["name"] for c in cl) for cl in CLASS.values()) and \
                    any(s in step[j]["name"] for s in ("reduce", "finalize", "bn_bwd_apply", "bn_head_bwd_apply")):
                ms, rd, wr = ms + step[j]["ms"], rd + step[j]["rd"], wr + step[j]["wr"]
                j += 1
        tf = t["flops"] / ms / 1e9 if t["flops"] else 0.0
        gbs = t["bytes"] / ms / 1e6
        f = fam.setdefault(k, [0.0, 0.0, 0.0, 0.0])
        f[0] += ms; f[1] += t["flops"]; f[2] += t["bytes"]; f[3] += rd + wr
        print(f"| {k} | {t['tag'] or '-'} | {ms:.3f} | {tf:.0f} | {tf / tens:.2f} | {t['bytes'] / 1e6:.0f} | "
              f"{rd / 1e6:.0f} + {wr / 1e6:.0f} | {(rd + wr) / max(t['bytes'], 1):.2f} | {gbs:.0f} | {gbs / hbm:.2f} |")
    print("\n| family | ms | share of step | TFLOP/s | algorithmic GB | DRAM GB | DRAM / alg |\n|---|---|---|---|---|---|---|")
    for k, (ms, fl, by, dr) in fam.items():
        print(f"| {k} | {ms:.2f} | {100 * ms / total:.1f}% | {fl / ms / 1e9:.0f} | {by / 1e9:.2f} | {dr / 1e9:.2f} | {dr / max(by, 1):.2f} |")


if __name__ == "__main__":
    main()
