"""Attention bottlenecks on the kernel library: the batched matrix products and the softmax of
transatt_unet's PAM / channel self-attention (transatt_unet.py:26-107) and unet_transformer's MHSA / MHCA
(unet_transformer.py:115-229) as ``torch.autograd.Function`` s over ``uz_bmm_tn`` (tcgen05 1x1 kernel with a per-image
B operand), ``uz_softmax_rows`` and ``uz_transpose_bf16``.  Tensors are bf16 token matrices ``[N, tokens, channels]``
(the NHWC activation seen flat), products accumulate in fp32.  Only the glue between these calls (bias / residual adds,
the learned scalar, dtype casts) is left to torch's elementwise kernels.
"""
import torch

from . import _lib as L


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _check(t, name):
    if not t.is_cuda or t.dtype != torch.bfloat16 or t.dim() != 3:
        raise RuntimeError(f"unet_zoo_b200.attention: {name} must be a CUDA bf16 tensor [N, rows, cols] "
                           f"(got {tuple(t.shape)} {t.dtype} on {t.device}); there is no CPU path")
    return t if t.is_contiguous() else t.contiguous()


def transpose_raw(x: torch.Tensor) -> torch.Tensor:
    """[N, R, C] -> [N, C, R] (contiguous), uz_transpose_bf16."""
    x = _check(x, "x")
    n, r, c = x.shape
    out = torch.empty((n, c, r), dtype=torch.bfloat16, device=x.device)
    with torch.cuda.device(x.device):
        L.check(L.load().uz_transpose_bf16(x.data_ptr(), c, r * c, out.data_ptr(), r, r * c, n, r, c, _stream(x)),
                "uz_transpose_bf16")
    return out


def bmm_tn_raw(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """C[n, m, j] = sum_k A[n, m, k] * B[n, j, k]; a [N, M, K], b [N, J, K] (or [1, J, K] with N == 1) -> [N, M, J]."""
    a, b = _check(a, "a"), _check(b, "b")
    n, m, k = a.shape
    if b.shape[0] != n or b.shape[2] != k:
        raise RuntimeError(f"unet_zoo_b200.attention.bmm_tn: {tuple(a.shape)} x {tuple(b.shape)}")
    j = b.shape[1]
    out = torch.empty((n, m, j), dtype=torch.bfloat16, device=a.device)
    with torch.cuda.device(a.device):
        L.check(L.load().uz_bmm_tn(a.data_ptr(), k, m * k, b.data_ptr(), j * k, out.data_ptr(), j, m * j, n, m, k, j,
                                   _stream(a)), "uz_bmm_tn")
    return out


class _Transpose(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return transpose_raw(x)

    @staticmethod
    def backward(ctx, g):
        return transpose_raw(g)


class _BmmTN(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return bmm_tn_raw(a, b)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        g = _check(g, "grad")
        da = db = None
        if ctx.needs_input_grad[0]:            # dA[n, m, k] = sum_j g[n, m, j] * B[n, j, k]
            da = bmm_tn_raw(g, transpose_raw(b))
        if ctx.needs_input_grad[1]:            # dB[n, j, k] = sum_m g[n, m, j] * A[n, m, k]
            db = bmm_tn_raw(transpose_raw(g), transpose_raw(a))
        return da, db


class _Softmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale):
        x = _check(x, "x")
        n, r, c = x.shape
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            L.check(L.load().uz_softmax_rows(x.data_ptr(), c, y.data_ptr(), c, n * r, c, float(scale), _stream(x)),
                    "uz_softmax_rows")
        ctx.save_for_backward(y)
        ctx.scale = float(scale)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        g = _check(g, "grad")
        n, r, c = y.shape
        dx = torch.empty_like(y)
        with torch.cuda.device(y.device):
            L.check(L.load().uz_softmax_rows_bwd(y.data_ptr(), c, g.data_ptr(), c, dx.data_ptr(), c, n * r, c, ctx.scale,
                                                 _stream(y)), "uz_softmax_rows_bwd")
        return dx, None


class _SoftmaxCols(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale):
        x = _check(x, "x")
        n, r, c = x.shape
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            L.check(L.load().uz_softmax_cols(x.data_ptr(), y.data_ptr(), n, r, c, float(scale), _stream(x)), "uz_softmax_cols")
        ctx.save_for_backward(y)
        ctx.scale = float(scale)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        g = _check(g, "grad")
        n, r, c = y.shape
        dx = torch.empty_like(y)
        with torch.cuda.device(y.device):
            L.check(L.load().uz_softmax_cols_bwd(y.data_ptr(), g.data_ptr(), dx.data_ptr(), n, r, c, ctx.scale, _stream(y)),
                    "uz_softmax_cols_bwd")
        return dx, None


def softmax_cols(x, scale: float = 1.0):
    """Differentiable ``softmax(scale * x, dim=1)`` on bf16 [N, R, C] (normalised over the rows of every column)."""
    return _SoftmaxCols.apply(x, scale)


def transpose(x):
    """[N, R, C] -> [N, C, R], differentiable."""
    return _Transpose.apply(x)


def bmm_tn(a, b):
    """Differentiable ``a @ b.transpose(1, 2)`` on bf16 [N, M, K] x [N, J, K]."""
    return _BmmTN.apply(a, b)


def softmax_rows(x, scale: float = 1.0):
    """Differentiable ``softmax(scale * x, dim=-1)`` on bf16 [N, R, C]."""
    return _Softmax.apply(x, scale)


class _LinearTokens(torch.autograd.Function):
    """y[n, t, j] = sum_k x[n, t, k] * w[j, k] with ONE weight matrix for all tokens.  The weight gradient is taken per
    image in the orientation whose row count is the larger of (K, J) - dW^T[n] = x[n]^T g[n] as a [K, J] product over the
    image's tokens - so that it runs on the tensor-core path with N x (K / 128) tiles (a single [J, K] product over all
    N * T tokens has J = 64 rows for the PAM query / key projections: below one tile, and no parallelism over K)."""

    @staticmethod
    def forward(ctx, x, w):
        n, tok, k = x.shape
        ctx.save_for_backward(x, w)
        return bmm_tn_raw(x.reshape(1, n * tok, k), w[None]).reshape(n, tok, -1)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        n, tok, k = x.shape
        g = _check(g, "grad")
        j = g.shape[2]
        dx = dw = None
        if ctx.needs_input_grad[0]:            # dx[t, k] = sum_j g[t, j] * w[j, k]
            dx = bmm_tn_raw(g.reshape(1, n * tok, j), transpose_raw(w[None])).reshape(n, tok, k)
        if ctx.needs_input_grad[1]:
            xt, gt = transpose_raw(x), transpose_raw(g)          # [N, K, T], [N, J, T]
            if k >= j:
                dw = bmm_tn_raw(xt, gt).float().sum(0).t()        # [N, K, J] -> [J, K]
            else:
                dw = bmm_tn_raw(gt, xt).float().sum(0)
            dw = dw.to(w.dtype)
        return dx, dw


def linear_tokens(t, weight, bias=None):
    """``t @ weight.T + bias`` for tokens t [N, T, K] (bf16) and an fp32 parameter weight [J, K] (a 1x1 Conv2d's weight
    viewed as a matrix): one product over all N * T tokens with the weight as the single B matrix."""
    w = weight.reshape(weight.shape[0], -1).to(torch.bfloat16)
    y = _LinearTokens.apply(_check(t, "t"), w)
    if bias is not None:
        y = (y.float() + bias).to(torch.bfloat16)
    return y
