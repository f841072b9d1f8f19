"""Top stalled SASS instructions of the first kernel in an `ncu --page source --csv` dump.
    ncu -i prof.ncu-rep --page source --csv > src.csv ; python tools/ncu_hot.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
start = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
k = int(sys.argv[3]) if len(sys.argv) > 3 else 0
lo = start[k]; hi = start[k + 1] if k + 1 < len(start) else len(rows)
print(rows[lo][1])
hdr = rows[lo + 1]; body = rows[lo + 2:hi]
ci = {h: i for i, h in enumerate(hdr)}
samp = ci["# Samples"]; src = ci["Source"]
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[samp] or 0) for r in body)
print("total samples", tot)
for r in sorted(body, key=lambda r: -int(r[samp] or 0))[:n]:
    st = sorted(((int(r[i] or 0), hdr[i]) for i in stall_cols), reverse=True)[:2]
    print(f"{int(r[samp]):7d} {100*int(r[samp])/tot:5.1f}%  {r[src][:70]:70s} {st}")
