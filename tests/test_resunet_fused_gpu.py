"""Inference form of ResidualConv (common_layers.py:182-199): the 1x1 skip branch riding along in the last 3x3 GEMM
(UZ_FLAG_RES1X1 / _S2, engine.residual_conv_tail_fused) and ResUnet's stem pair (resunet.py:29-33,56) against the
reference arithmetic in fp32 on the same bf16 operands."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from conftest import seeded_model, rel_err
from gpu_util import bf16r, make_act, new_ctx, relmax
from oracle import reference_port as O
from unet_zoo_b200 import engine as E

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


@pytest.fixture(autouse=True)
def _exact_fp32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(21)


@pytest.mark.parametrize("N,H,W,ch,cx,cout,s2", [(2, 32, 32, 64, 128, 64, False), (2, 16, 24, 128, 64, 128, True),
                                                 (1, 48, 40, 256, 512, 256, False), (3, 16, 8, 64, 32, 64, True)])
def test_residual_tail_fused(N, H, W, ch, cx, cout, s2):
    conv2 = nn.Conv2d(ch, cout, 3, padding=1, bias=False).to(DEV)
    conv_s = nn.Conv2d(cx, cout, 1, stride=2 if s2 else 1, bias=False).to(DEV)
    bn_s = nn.BatchNorm2d(cout).to(DEV).eval()
    with torch.no_grad():
        bn_s.running_mean.normal_()
        bn_s.running_var.uniform_(0.5, 2.0)
        bn_s.weight.uniform_(0.5, 1.5)
        bn_s.bias.normal_()
    h = bf16r(torch.randn(N, ch, H, W, device=DEV))
    x = bf16r(torch.randn(N, cx, H * (2 if s2 else 1), W * (2 if s2 else 1), device=DEV))
    ctx = new_ctx(training=False, record=False)
    wide = ctx.new_buf(N, H, W, 2 * cout)
    wide.data.fill_(5.0)
    out = E.residual_conv_tail_fused(ctx, make_act(ctx, h), make_act(ctx, x), conv2, conv_s, bn_s, E.Act(wide, cout, cout), s2)
    assert out is not None
    with torch.no_grad():
        ref = conv2(h) + bn_s(conv_s(x))
    torch.cuda.synchronize()
    assert relmax(out.to_nchw(), ref) < 1.5e-2
    assert bool((wide.data[..., :cout] == 5.0).all())


def test_resunet_eval_fused_equals_two_launch_form(monkeypatch):
    """Whole model, eval mode: the fused graph (default) against the two-launch form (UZ_RESUNET_UNFUSED) and the oracle."""
    model = seeded_model("resunet").cuda().eval()
    x, _ = O.synthetic_batch(2, 3, 128, 128)
    with torch.no_grad():
        a = model(x.cuda()).cpu()
        n_fused = model._last_launches
    monkeypatch.setenv("UZ_RESUNET_UNFUSED", "1")
    model._eval_cache = None
    with torch.no_grad():
        b = model(x.cuda()).cpu()
        n_plain = model._last_launches
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        ref = O.resunet_forward(sd, x, False)
    assert n_fused < n_plain
    assert rel_err(a, ref) < 2e-2 and rel_err(b, ref) < 2e-2 and rel_err(a, b) < 2e-2
