"""Micro-benchmark of the bandwidth kernels on the layer shapes of the unet config (batch 32): achieved GB/s against
the algorithmic bytes.   python tools/bench_bn.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from unet_zoo_b200 import engine as E

def timeit(fn, iters=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

ctx = E.Ctx(torch.device("cuda", 0), training=True, record=False)
N = 32
for C, S in ((64, 512), (128, 256), (256, 128), (512, 64)):
    z = ctx.new_act(N, S, S, C); z.buf.data.normal_()
    dy = ctx.new_act(N, S, S, C); dy.buf.data.normal_()
    y = ctx.new_act(N, S, S, C)
    dz = ctx.new_act(N, S, S, C)
    pooled = ctx.new_act(N, S // 2, S // 2, C); pooled.buf.data.normal_()
    bn = torch.zeros(4, C, device="cuda"); bn[1] = 1; bn[2] = 1
    dg, db = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    gb = N * S * S * C * 2 / 1e9
    t = timeit(lambda: ctx.bn_apply(z, bn[2].data_ptr(), bn[3].data_ptr(), True, y))
    t2 = timeit(lambda: ctx.bn_apply(z, bn[2].data_ptr(), bn[3].data_ptr(), True, y, pooled))
    t3 = timeit(lambda: ctx.bn_bwd(z, dy.ptr, dy.ld, None, 0, bn, True, dz.ptr, dz.ld, dg, db, None))
    t4 = timeit(lambda: ctx.bn_bwd(z, dy.ptr, dy.ld, pooled.ptr, pooled.ld, bn, True, dz.ptr, dz.ld, dg, db, None))
    print(f"C={C} S={S} ({gb:.2f} GB/pass): apply {t:.3f} ms {2*gb/t:.0f} GB/s | apply+pool {t2:.3f} ms {2.25*gb/t2:.0f} GB/s | "
          f"bwd {t3:.3f} ms {5*gb/t3:.0f} GB/s | bwd+pool {t4:.3f} ms {5.5*gb/t4:.0f} GB/s", flush=True)
    del z, dy, y, dz, pooled
    torch.cuda.empty_cache()
