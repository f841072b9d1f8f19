import sys, torch, torch.nn as nn, torch.nn.functional as F
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from gpu_util import bf16r, make_act, new_ctx
from unet_zoo_b200 import engine as E, _lib as L
torch.manual_seed(0)
dev = torch.device("cuda", 0)
N, H, W, cin, cout = 1, 16, 8, 64, 64
w = bf16r(torch.randn(cout, cin, 3, 3, device=dev) * 0.1)
ctx = new_ctx(record=False)
# dgrad alone: dz high-res -> dx low-res
dz = bf16r(torch.randn(N, cout, 2 * H, 2 * W, device=dev))
dza = make_act(ctx, dz)
for phase_form in (1, 0):
    out = ctx.new_act(N, H, W, cin)
    wpd = ctx.pack_up2x(w, True) if phase_form else ctx.pack(w, "conv_dgrad")
    if not phase_form:
        import os; os.environ["UZ_UP2X_GEN1"] = "1"
    x = torch.zeros(N, cin, H, W, device=dev, dtype=torch.double, requires_grad=True)
    z = F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), w.double(), padding=1)
    (z * dz.double()).sum().backward()
    if phase_form:
        ctx.conv(L.UP2X_CONV3X3_DGRAD, [dza], wpd, cin, None, out.ptr, out.ld, H, W)
        torch.cuda.synchronize()
        got = out.to_nchw().double()
        err = (got - x.grad).abs()
        print("phase form rel_l2", float((got - x.grad).norm() / x.grad.norm()))
        print("err map (max over channels):")
        print((err.amax(1)[0] / x.grad.abs().mean()).cpu().numpy().round(2))
        print("err per channel (max over pixels), first 16:", (err.amax((2, 3))[0][:16] / x.grad.abs().mean()).cpu().numpy().round(2))
