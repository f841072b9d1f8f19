"""Secondary reference point (SURVEY 8d / BASELINE.md 3): the same UNet architecture (DoubleConv x2 per level,
MaxPool2d, ConvTranspose2d k2s2, cat [up, skip], 1x1 head; widths 64..1024) written with stock torch.nn modules and run
by stock PyTorch / cuDNN on the same GPU: autocast(bfloat16) + channels_last, AdamW(fused), clip_grad_norm_(1.0).
This is what a user of the reference gets on a B200 without this library.  Prints one JSON line.

    python tools/bench_stock_torch.py [--batch 32] [--size 512] [--steps 5] [--warmup 2] [--graph]
"""
import argparse
import json

import torch
import torch.nn as nn


def block(cin, cout):
    return nn.Sequential(nn.Conv2d(cin, cout, 3, padding=1), nn.BatchNorm2d(cout), nn.ReLU(inplace=True),
                         nn.Conv2d(cout, cout, 3, padding=1), nn.BatchNorm2d(cout), nn.ReLU(inplace=True))


class StockUNet(nn.Module):
    def __init__(self, cin=3, classes=1, widths=(64, 128, 256, 512, 1024)):
        super().__init__()
        self.enc = nn.ModuleList([block(c0, c1) for c0, c1 in zip((cin,) + widths[:-2], widths[:-1])])
        self.pool = nn.MaxPool2d(2, 2)
        self.mid = block(widths[-2], widths[-1])
        self.up = nn.ModuleList([nn.ConvTranspose2d(w, w // 2, 2, 2) for w in reversed(widths[1:])])
        self.dec = nn.ModuleList([block(w, w // 2) for w in reversed(widths[1:])])
        self.head = nn.Conv2d(widths[0], classes, 1)

    def forward(self, x):
        skips = []
        for e in self.enc:
            x = e(x)
            skips.append(x)
            x = self.pool(x)
        x = self.mid(x)
        for u, d, s in zip(self.up, self.dec, reversed(skips)):
            x = d(torch.cat([u(x), s], 1))
        return self.head(x)


# ---- the MultiResUnet architecture (three chained conv-BN-ReLU branches concatenated, 1x1 shortcut, BatchNorm applied
# twice; Respath skips; every BatchNorm without affine parameters; widths int(int(f * 1.67) * {0.167, 0.333, 0.5})) in stock
# torch.nn modules, for the same-GPU library number of that model
def _cb(cin, cout, k, relu):
    layers = [nn.Conv2d(cin, cout, k, padding=k // 2), nn.BatchNorm2d(cout, affine=False)]
    return nn.Sequential(*(layers + ([nn.ReLU(inplace=True)] if relu else [])))


def _mr_widths(f):
    w = int(f * 1.67)
    return int(w * 0.167), int(w * 0.333), int(w * 0.5)


class StockMRB(nn.Module):
    def __init__(self, cin, f):
        super().__init__()
        a, b, c = _mr_widths(f)
        self.out = a + b + c
        self.short = _cb(cin, self.out, 1, False)
        self.c3, self.c5, self.c7 = _cb(cin, a, 3, True), _cb(a, b, 3, True), _cb(b, c, 3, True)
        self.bn = nn.BatchNorm2d(self.out, affine=False)

    def forward(self, x):
        t = self.short(x)
        a = self.c3(x)
        b = self.c5(a)
        c = self.c7(b)
        return self.bn(torch.relu(self.bn(torch.cat([a, b, c], 1)) + t))


class StockRespath(nn.Module):
    def __init__(self, cin, f, n):
        super().__init__()
        self.steps = nn.ModuleList([nn.ModuleList([_cb(cin if i == 0 else f, f, 1, False), _cb(cin if i == 0 else f, f, 3, True),
                                                   nn.BatchNorm2d(f, affine=False)]) for i in range(n)])

    def forward(self, x):
        for short, conv, bn in self.steps:
            x = bn(torch.relu(conv(x) + short(x)))
        return x


class StockMultiResUnet(nn.Module):
    def __init__(self, cin=3, classes=1, f=32):
        super().__init__()
        self.enc, self.paths = nn.ModuleList(), nn.ModuleList()
        c = cin
        for i in range(4):
            blk = StockMRB(c, f << i)
            self.enc.append(blk)
            self.paths.append(StockRespath(blk.out, f << i, 4 - i))
            c = blk.out
        self.mid = StockMRB(c, f << 4)
        c = self.mid.out
        self.up, self.dec = nn.ModuleList(), nn.ModuleList()
        for i in (3, 2, 1, 0):
            self.up.append(nn.ConvTranspose2d(c, f << i, 2, 2))
            blk = StockMRB(2 * (f << i), f << i)
            self.dec.append(blk)
            c = blk.out
        self.final = _cb(c, classes, 1, False)
        self.pool = nn.MaxPool2d(2, 2)

    def forward(self, x):
        skips = []
        for blk, path in zip(self.enc, self.paths):
            x = blk(x)
            skips.append(path(x))
            x = self.pool(x)
        x = self.mid(x)
        for u, d, s in zip(self.up, self.dec, reversed(skips)):
            x = d(torch.cat([u(x), s], 1))
        return self.final(x)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--model", default="unet", choices=["unet", "multiresunet"])
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    torch.backends.cudnn.benchmark = True
    model = (StockUNet() if a.model == "unet" else StockMultiResUnet()).to(dev).to(memory_format=torch.channels_last).train()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-5, fused=True)
    crit = nn.BCEWithLogitsLoss()
    x = torch.randn(a.batch, 3, a.size, a.size, device=dev).contiguous(memory_format=torch.channels_last)
    y = (torch.rand(a.batch, 1, a.size, a.size, device=dev) > 0.5).float()

    def step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
        loss = crit(out.float(), y)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        return loss

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    print(json.dumps({"impl": "stock torch.nn + cuDNN, autocast bf16, channels_last", "model": a.model, "torch": torch.__version__,
                      "cudnn": torch.backends.cudnn.version(), "batch": a.batch, "size": a.size, "ms_per_step": ms,
                      "images_per_sec": a.batch * 1000.0 / ms, "loss": float(loss),
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30}))


if __name__ == "__main__":
    main()
