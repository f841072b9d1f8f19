"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table of ONE step.

    python tools/summarize_launches.py gpurun_out/launches.csv [marker-substring] > profiles/<name>.md

A step is the run of launches between the last two occurrences of the marker kernel (default: stem_im2col, the first
kernel of every forward).  Per-launch times under ncu are serialised and cold-cache: read the SHARES.
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name):
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("void ", "")
    return name[:90]


def main():
    path = sys.argv[1]
    marker = sys.argv[2] if len(sys.argv) > 2 else "stem_im2col"
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((r["Kernel Name"], float(r["Metric Value"]) / 1e6, r["Grid Size"], r["Block Size"]))
    idx = [i for i, r in enumerate(rows) if marker in r[0]]
    if len(idx) >= 2:
        step = rows[idx[-2]:idx[-1]]
    else:
        step = rows
    agg = OrderedDict()
    for name, ms, _, _ in step:
        k = short(name)
        t, n = agg.get(k, (0.0, 0))
        agg[k] = (t + ms, n + 1)
    total = sum(t for t, _ in agg.values())
    print(f"step total under ncu: {total:.2f} ms, {len(step)} launches\n")
    print("| ms | share | launches | kernel |\n|---|---|---|---|")
    for k, (t, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"| {t:.3f} | {100 * t / total:.1f}% | {n} | `{k}` |")


if __name__ == "__main__":
    main()
