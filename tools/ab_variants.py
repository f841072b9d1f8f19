"""Same-box A/B of library variants (tools/build_variants.sh): per-layer conv timings and the unet bench for each
UZ_LIB_PATH, interleaved and repeated.   python tools/ab_variants.py k0 k1 cur [--bench]"""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
names = [a for a in sys.argv[1:] if not a.startswith("--")]
do_bench = "--bench" in sys.argv
model = next((a.split("=")[1] for a in sys.argv if a.startswith("--model=")), "unet")
batch = {"unet": 32, "attention_unet": 16, "nested_unet": 8}.get(model, 32)


def env_for(n):
    e = dict(os.environ)
    e.pop("UZ_LIB_PATH", None)
    if n != "cur":
        e["UZ_LIB_PATH"] = os.path.join(ROOT, "build", f"libuz_{n}.so")
    return e


rows = {}
for rep in range(2):
    for n in names:
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "bench_conv.py")], env=env_for(n),
                             capture_output=True, text=True).stdout
        for ln in out.splitlines():
            m = re.match(r"(.+?)\s+([\d.]+) ms", ln)
            if m:
                rows.setdefault(m.group(1).strip(), {}).setdefault(n, []).append(float(m.group(2)))
        m = re.search(r"total ([\d.]+)", out)
        rows.setdefault("TOTAL", {}).setdefault(n, []).append(float(m.group(1)) if m else float("nan"))
print("layer".ljust(26) + "".join(n.rjust(16) for n in names))
for k, v in rows.items():
    print(k.ljust(26) + "".join(("/".join(f"{t:.3f}" for t in v.get(n, []))).rjust(16) for n in names))
if do_bench:
    for rep in range(2):
        for n in names:
            out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--no-secondary", "--no-cpu-baseline",
                                  "--model", model, "--batch", str(batch)], env=env_for(n), capture_output=True, text=True).stdout
            line = json.loads(out.strip().splitlines()[-1])
            print(f"{n} {model}: {line['value']:.1f} img/s {line['ms_per_step']:.3f} ms  sm {line['clocks']['sm_mhz']} MHz")
