"""bench.py's reference arm on CPU (the one leg that needs no GPU): the JSON line carries the keys the driver reads
(`impl`, metric / unit / config of the CUDA arm, `cpu_baseline`, an `e2e` object without copies), and the CUDA arm
refuses to run without a device instead of falling back to anything."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def _run(*args, timeout=600):
    env = dict(os.environ, OMP_NUM_THREADS="4")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          cwd=ROOT, env=env, timeout=timeout)


def test_reference_arm_json_line():
    out = _run("--impl", "reference", "--size", "32", "--steps", "1", "--warmup", "1")
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "images/sec" and d["higher_is_better"] is True
    assert d["metric"] == "images/sec UNet 512^2 bf16 train step" and d["n_gpus"] == 1 and d["value"] > 0
    assert "unet in=3 classes=1" in d["config"]["workload"] and "configs[1]" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "images/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_cuda_arm_needs_a_device():
    import torch
    if torch.cuda.is_available():
        return                                               # covered by the GPU bench itself
    out = _run("--steps", "1", "--warmup", "1", "--no-cpu-baseline", timeout=300)
    assert out.returncode != 0
    assert not [ln for ln in out.stdout.splitlines() if ln.startswith("{")]     # no number without the CUDA path
