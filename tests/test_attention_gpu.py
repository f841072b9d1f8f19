"""The attention-bottleneck kernels (uz_bmm_tn on the tcgen05 1x1 kernel with a per-image B operand, its CUDA-core
fallback for shapes below one tile, uz_softmax_rows / _bwd, uz_transpose_bf16) and their autograd wrappers
(unet_zoo_b200/attention.py) against PyTorch fp32 on the same bf16 operands."""
import pytest
import torch

from gpu_util import bf16r, rel_l2, relmax
from unet_zoo_b200 import attention as A

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


@pytest.fixture(autouse=True)
def _exact_fp32():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)


@pytest.mark.parametrize("N,M,K,J", [
    (3, 1024, 64, 1024),     # PAM energy at 512^2 input: q k^T, one 64-channel K block, four N tiles of 256
    (2, 1024, 1024, 512),    # attention applied to V: K = 1024 tokens (16 K blocks), BN = 256 x 2
    (2, 512, 1024, 512),     # channel attention x x^T over 1024 pixels
    (1, 4096, 512, 64),      # a 1x1 projection of all tokens with ONE weight matrix
    (2, 136, 72, 40),        # ragged: 17 pixel rows, K and J not multiples of 64
    (5, 128, 8, 8),          # smallest fast-path shape
    (2, 16, 512, 16),        # far below one 128-row tile (4x4 bottleneck of a 64 x 64 image): TMA zero fill / clipping
    (8, 64, 4096, 64),       # per-image weight gradient of a 64-channel projection: half a tile, long K
    (3, 100, 36, 52),        # nothing aligned: CUDA-core kernel
])
def test_bmm_tn_and_its_gradients(N, M, K, J):
    a = bf16r(torch.randn(N, M, K, device=DEV) * 0.5).to(torch.bfloat16).requires_grad_(True)
    b = bf16r(torch.randn(N, J, K, device=DEV) * 0.5).to(torch.bfloat16).requires_grad_(True)
    c = A.bmm_tn(a, b)
    ar, br = a.detach().float().requires_grad_(True), b.detach().float().requires_grad_(True)
    ref = ar @ br.transpose(1, 2)
    assert c.shape == ref.shape and c.dtype == torch.bfloat16
    assert relmax(c.float(), ref.detach()) < 1e-2
    g = bf16r(torch.randn_like(ref) * 0.1)
    c.backward(g.to(torch.bfloat16))
    ref.backward(g)
    assert rel_l2(a.grad.float(), ar.grad) < 1e-2
    assert rel_l2(b.grad.float(), br.grad) < 1e-2


@pytest.mark.parametrize("N,R,C,scale", [(2, 1024, 1024, 1.0), (3, 512, 512, 512 ** -0.5), (1, 16, 16, 1.0), (2, 100, 4096, 0.25),
                                         (1, 7, 33, 1.0)])
def test_softmax_rows_fwd_bwd(N, R, C, scale):
    x = bf16r(torch.randn(N, R, C, device=DEV) * 3).to(torch.bfloat16).requires_grad_(True)
    y = A.softmax_rows(x, scale)
    xr = x.detach().float().requires_grad_(True)
    ref = torch.softmax(scale * xr, dim=-1)
    assert relmax(y.float(), ref.detach()) < 1e-2
    assert float((y.float().sum(-1) - 1).abs().max()) < 2e-2
    g = bf16r(torch.randn_like(ref))
    y.backward(g.to(torch.bfloat16))
    # reference backward through the SAME (bf16-rounded) probabilities
    yq = y.detach().float()
    want = scale * yq * (g - (g * yq).sum(-1, keepdim=True))
    assert rel_l2(x.grad.float(), want) < 1e-2


@pytest.mark.parametrize("N,R,C,scale", [(2, 4096, 4096, 1 / 16.0), (3, 64, 200, 1.0), (1, 7, 33, 0.5), (2, 1024, 64, 1.0)])
def test_softmax_cols_fwd_bwd(N, R, C, scale):
    """nn.Softmax(dim=1) of a [b, rows, cols] matrix (unet_transformer's attention, normalised over the queries)."""
    x = bf16r(torch.randn(N, R, C, device=DEV) * 3).to(torch.bfloat16).requires_grad_(True)
    y = A.softmax_cols(x, scale)
    ref = torch.softmax(scale * x.detach().float(), dim=1)
    assert relmax(y.float(), ref) < 1e-2
    assert float((y.float().sum(1) - 1).abs().max()) < 2e-2
    g = bf16r(torch.randn_like(ref))
    y.backward(g.to(torch.bfloat16))
    yq = y.detach().float()
    want = scale * yq * (g - (g * yq).sum(1, keepdim=True))
    assert rel_l2(x.grad.float(), want) < 1e-2


@pytest.mark.parametrize("N,R,C", [(2, 1024, 512), (3, 17, 40), (1, 1, 8), (4, 64, 64), (2, 4096, 4096), (3, 72, 200), (2, 64, 8), (1, 4096, 64)])
def test_transpose(N, R, C):
    x = torch.randn(N, R, C, device=DEV).to(torch.bfloat16)
    assert torch.equal(A.transpose(x), x.transpose(1, 2).contiguous())


def test_linear_tokens_matches_conv1x1():
    conv = torch.nn.Conv2d(512, 64, 1).to(DEV)
    t = bf16r(torch.randn(2, 32 * 32, 512, device=DEV)).to(torch.bfloat16)
    y = A.linear_tokens(t, conv.weight, conv.bias)
    ref = t.float() @ bf16r(conv.weight.detach().reshape(64, 512)).t() + conv.bias.detach()
    assert relmax(y.float(), ref) < 1e-2


@pytest.mark.parametrize("N,T,K,J", [(4, 1024, 512, 64), (2, 1024, 512, 512), (2, 16, 512, 64), (3, 256, 64, 256)])
def test_linear_tokens_gradients(N, T, K, J):
    """Input and weight gradient (per-image orientation with the larger of K, J as rows) against torch fp32."""
    w = torch.nn.Parameter(bf16r(torch.randn(J, K, 1, 1, device=DEV) * 0.05))
    bias = torch.nn.Parameter(torch.randn(J, device=DEV) * 0.1)
    t = bf16r(torch.randn(N, T, K, device=DEV)).to(torch.bfloat16).requires_grad_(True)
    y = A.linear_tokens(t, w, bias)
    tr, wr, br = t.detach().float().requires_grad_(True), w.detach().clone().requires_grad_(True), bias.detach().clone().requires_grad_(True)
    ref = tr @ wr.reshape(J, K).t() + br
    g = bf16r(torch.randn_like(ref) * 0.1)
    y.backward(g.to(torch.bfloat16))
    ref.backward(g)
    assert relmax(y.float(), ref.detach()) < 1e-2
    assert rel_l2(t.grad.float(), tr.grad) < 1e-2
    assert rel_l2(w.grad, wr.grad) < 1e-2 and w.grad.shape == w.shape
    assert rel_l2(bias.grad, br.grad) < 1e-2


def test_cpu_tensors_are_rejected():
    with pytest.raises(RuntimeError, match="no CPU path"):
        A.bmm_tn(torch.zeros(1, 128, 8, dtype=torch.bfloat16), torch.zeros(1, 8, 8, dtype=torch.bfloat16))
