"""`transatt_unet` host logic on CPU: the stock-torch bottleneck modules against the oracle's
restatement of the reference (transatt_unet.py:26-107,143-154), and the tape mechanics of engine.torch_island on host
tensors (no kernel runs: a duck-typed context stands in for the CUDA executor's)."""
import copy

import torch

from conftest import rel_err, seeded_model
from oracle import reference_port as O
from unet_zoo_b200 import engine as E


def test_bottleneck_modules_match_the_oracle_forward_and_backward():
    model = seeded_model("transatt_unet", attn_dropout=0.0)
    with torch.no_grad():
        model.pam.gamma.fill_(0.7)                              # zero at init: make the PAM branch count
    sd = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in model.state_dict().items()}
    x = torch.randn(2, 512, 6, 4, generator=torch.Generator().manual_seed(3))
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    for training in (False, True):
        model.train(training)
        got = model._bottleneck(xa)
        x5 = xb + O._position_embedding(sd, xb)
        ref = O._sdpa(x5, 512.0, 0.0, training) + O._pam(O.Runner(sd, training), x5)
        assert got.shape == ref.shape == x.shape
        assert rel_err(got.detach(), ref.detach()) < 1e-5
    w = torch.randn(x.shape, generator=torch.Generator().manual_seed(4))
    (got * w).sum().backward()
    (ref * w).sum().backward()
    assert rel_err(xa.grad, xb.grad) < 1e-4
    for k, p in model.named_parameters():
        if k.startswith(("pos.", "pam.")):
            assert rel_err(p.grad, sd[k].grad) < 1e-4, k


def test_position_table_limit_is_reported():
    model = seeded_model("transatt_unet")
    try:
        model.pos(torch.zeros(1, 512, 33, 8))
    except RuntimeError as e:
        assert "position table" in str(e)
    else:
        raise AssertionError("a 33-row bottleneck map must be rejected (32-entry embedding, transatt_unet.py:59)")


def test_oversized_input_is_rejected_before_any_kernel_runs():
    model = seeded_model("transatt_unet")
    for bad in (torch.zeros(1, 3, 528, 64), torch.zeros(1, 64, 528, 3, dtype=torch.uint8)):
        try:
            model(bad)
        except RuntimeError as e:
            assert "up to 512x512" in str(e), str(e)
        else:
            raise AssertionError("a 528-pixel side must be rejected (33 > 32 position entries)")
    try:
        model(torch.zeros(1, 3, 512, 512))                     # size is fine: the next check is the device
    except RuntimeError as e:
        assert "no CPU path" in str(e)


class _HostCtx:
    """The members of engine.Ctx that torch_island touches, on host memory."""

    def __init__(self, params, record=True):
        self.record, self.tape, self.device, self.fp32 = record, [], torch.device("cpu"), False
        self.pgrads = E.ParamGrads(list(params), self.device)

    def new_act(self, N, H, W, C):
        return E.Act(E.Buf(N, H, W, C, self.device))

    def pgrad(self, p):
        return self.pgrads.view(p) if p.requires_grad else None

    def pready(self, *ps):
        for p in ps:
            self.pgrads.ready(p)


def test_torch_island_tape_on_host_tensors():
    torch.manual_seed(5)
    conv = torch.nn.Conv2d(16, 24, 1)
    frozen = torch.nn.Parameter(torch.ones(1), requires_grad=False)
    ctx = _HostCtx([conv.weight, conv.bias, frozen])
    buf = E.Buf(2, 4, 4, 32, ctx.device)
    buf.data.copy_(torch.randn(2, 4, 4, 32))
    x = E.Act(buf, 8, 16)                                       # a channel slice of a wider buffer
    out = E.torch_island(ctx, x, lambda t: torch.tanh(conv(t)) * frozen, [conv.weight, conv.bias, frozen])
    xr = x.to_nchw().clone().requires_grad_(True)
    ref = torch.tanh(conv(xr))
    assert (out.N, out.H, out.W, out.C) == (2, 4, 4, 24)
    assert rel_err(out.to_nchw(), ref.detach()) < 5e-3            # stored as bf16
    # gradient arrives as bf16 in the output's gradient buffer; the input slice already holds a partial gradient in its
    # first 8 channels (accumulate there, overwrite the rest)
    dy = torch.randn(2, 24, 4, 4).to(torch.bfloat16)
    out.buf.grad_buf().copy_(dy.permute(0, 2, 3, 1))
    out.buf.written.append((0, 24))
    prior = torch.randn(2, 4, 4, 8).to(torch.bfloat16)
    buf.grad_buf().zero_()
    buf.grad[..., 8:16] = prior
    buf.written.append((8, 16))
    ctx.tape[-1](ctx)
    ref.backward(dy.float())
    want = xr.grad.permute(0, 2, 3, 1).clone()
    want[..., :8] = (want[..., :8].to(torch.bfloat16) + prior).float()         # bf16 accumulate, as add_ does
    got = buf.grad[..., 8:24].float()
    assert rel_err(got[..., 8:], want[..., 8:]) < 5e-3
    assert rel_err(got[..., :8], want[..., :8]) < 2e-2
    assert torch.allclose(ctx.pgrads.view(conv.weight), conv.weight.grad, rtol=1e-4, atol=1e-5)
    assert torch.allclose(ctx.pgrads.view(conv.bias), conv.bias.grad, rtol=1e-4, atol=1e-5)
    assert ctx.pgrads.done == {id(conv.weight), id(conv.bias)}


def test_torch_island_inference_keeps_no_graph():
    conv = torch.nn.Conv2d(8, 8, 1)
    ctx = _HostCtx([conv.weight, conv.bias], record=False)
    x = ctx.new_act(1, 2, 2, 8)
    x.buf.data.copy_(torch.randn(1, 2, 2, 8))
    with torch.no_grad():
        out = E.torch_island(ctx, x, conv, list(conv.parameters()))
    assert not ctx.tape and out.C == 8
    try:
        E.torch_island(ctx, x, lambda t: t[:, :, :1], [])
    except RuntimeError as e:
        assert "torch island returned" in str(e)
    else:
        raise AssertionError("a spatial-size change inside the island must be rejected")


def test_island_math_follows_the_tf32_api_the_process_uses():
    """torch >= 2.9 raises when `allow_tf32` is read after `fp32_precision` was written (and vice versa): the island's
    precision guard must work, and restore the previous setting, on either front end.  Run in a subprocess: the
    choice of front end is sticky for the whole process."""
    import subprocess
    import sys
    code = r'''
import torch
from unet_zoo_b200 import engine as E
mm, cd = torch.backends.cuda.matmul, torch.backends.cudnn
mm.allow_tf32, cd.allow_tf32 = False, True                    # legacy front end
with E._island_math(True):
    assert mm.allow_tf32 and cd.allow_tf32
assert (mm.allow_tf32, cd.allow_tf32) == (False, True)
with E._island_math(False):
    assert not mm.allow_tf32 and not cd.allow_tf32
assert (mm.allow_tf32, cd.allow_tf32) == (False, True)
if hasattr(mm, "fp32_precision"):
    mm.fp32_precision = "tf32"                                 # the application switches to the new front end:
    cd.conv.fp32_precision = "tf32"                            # reading mm.allow_tf32 now raises inside torch
    try:
        mm.allow_tf32
        mixed = False
    except RuntimeError:
        mixed = True
    with E._island_math(False):
        assert mm.fp32_precision != "tf32" and cd.conv.fp32_precision != "tf32"
    assert (mm.fp32_precision, cd.conv.fp32_precision) == ("tf32", "tf32")
    with E._island_math(True):
        assert mm.fp32_precision == "tf32" and cd.conv.fp32_precision == "tf32"
    assert (mm.fp32_precision, cd.conv.fp32_precision) == ("tf32", "tf32")
    print("mixed" if mixed else "consistent")
print("ok")
'''
    import os
    env = dict(os.environ, PYTHONPATH=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]
