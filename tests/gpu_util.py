"""Helpers for the GPU parity tests: move NCHW fp32 tensors in and out of the executor's NHWC bf16 views."""
import torch

from unet_zoo_b200 import engine as E


def bf16r(t):
    """Round to bf16 and back: the values the CUDA path actually sees."""
    return t.to(torch.bfloat16).float()


def make_act(ctx, x_nchw, ld=None, c0=0):
    """NCHW fp32 (already bf16-representable or not) -> Act view [c0, c0+C) of a fresh buffer with `ld` channels."""
    N, C, H, W = x_nchw.shape
    ld = ld or C
    buf = ctx.new_buf(N, H, W, ld)
    buf.data.zero_()
    buf.data[..., c0:c0 + C] = x_nchw.permute(0, 2, 3, 1).to(torch.bfloat16)
    return E.Act(buf, c0, C)


def set_grad(act, g_nchw):
    """Install g as the gradient of `act` (as if a consumer had written it)."""
    gb = act.buf.grad_buf()
    gb[..., act.c0:act.c0 + act.C] = g_nchw.permute(0, 2, 3, 1).to(torch.bfloat16)
    act.buf.written.append((act.c0, act.c0 + act.C))


def get_grad(act):
    return act.buf.grad_buf()[..., act.c0:act.c0 + act.C].permute(0, 3, 1, 2).float()


def new_ctx(training=True, record=True, params=()):
    ctx = E.Ctx(torch.device("cuda", 0), training=training, record=record)
    ctx.pgrads = E.ParamGrads(list(params), ctx.device)
    return ctx


def run_tape(ctx):
    for step in reversed(ctx.tape):
        step(ctx)
    ctx.tape.clear()
    torch.cuda.synchronize()


def relmax(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
