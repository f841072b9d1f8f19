"""PC samples of an `ncu --import-source on` capture aggregated per CUDA source line.
    python tools/ncu_hot_lines.py prof.ncu-rep [N]"""
import collections
import csv
import io
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
n = int(sys.argv[2]) if len(sys.argv) > 2 else 24
cur_file, hdr, cur_line, si = None, None, None, None
agg, srcs, stalls = collections.Counter(), {}, {}
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        si = [i for i, h in enumerate(hdr) if h == "# Samples"][0]
        sc = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None:
        continue
    if r[0].isdigit():
        cur_line = (cur_file, int(r[0]))
        srcs[cur_line] = r[1]
        continue
    if r[0] == "" and len(r) > si and cur_line:
        try:
            v = int(r[si] or 0)
        except ValueError:
            continue
        agg[cur_line] += v
        st = stalls.setdefault(cur_line, collections.Counter())
        for i in sc:
            try:
                st[hdr[i]] += int(r[i] or 0)
            except ValueError:
                pass
tot = sum(agg.values())
print("total samples", tot)
for (f, ln), s in agg.most_common(n):
    top = ", ".join(f"{k[6:]}={v}" for k, v in stalls[(f, ln)].most_common(2))
    print(f"{100 * s / max(tot, 1):5.1f}%  {f}:{ln}  {srcs[(f, ln)].strip()[:90]}   [{top}]")
