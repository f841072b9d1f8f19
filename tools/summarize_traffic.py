"""DRAM bytes per launch of the GEMM kernel families of ONE step, from an ncu CSV that holds
dram__bytes_read.sum / dram__bytes_write.sum per launch:

    python tools/summarize_traffic.py gpurun_out/launches.csv "<source note>" > profiles/r01_traffic.json
"""
import csv
import json
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    path, note = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    per = {}   # launch id -> [name, bytes]
    order = []
    for r in csv.DictReader(lines):
        if r.get("Metric Name") in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            v = float(r["Metric Value"].replace(",", "")) * UNIT.get(r.get("Metric Unit", "byte"), 1.0)
            if r["ID"] not in per:
                per[r["ID"]] = [r["Kernel Name"], 0.0]
                order.append(r["ID"])
            per[r["ID"]][1] += v
    rows = [per[i] for i in order]
    idx = [i for i, r in enumerate(rows) if "stem_im2col" in r[0]]
    step = rows[idx[-2]:idx[-1]] if len(idx) >= 2 else rows
    fam = {"igemm_fwd": lambda n: "igemm_fwd" in n, "igemm_wgrad": lambda n: "igemm_wgrad" in n}
    out = {}
    for k, pred in fam.items():
        sel = [b for n, b in step if pred(n)]
        if sel:
            out[k] = {"launches": len(sel), "dram_bytes_per_launch": sum(sel) / len(sel)}
    out["source"] = note
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
