"""fp32 parity mode of the graph executor (forward only).

BASELINE.json asks for "fp32 mode within 1e-4 relative on logits" and names `unet`, batch 1, 256x256 fp32 forward as
configs[0] (the reference's README example, README.md:83-90).  The bf16 tensor-core operators of :mod:`engine` cannot
meet 1e-4, so every operator has a second implementation here on NHWC **fp32** activations, launching the CUDA-core
kernels of ``csrc/f32.cu`` through the same C ABI.  The model graphs (``models/unet.py``, ``models/_graphs.py``) are
shared: :class:`engine.Ctx` carries ``fp32=True`` and each operator of :mod:`engine` dispatches here.

Same fusion structure as the bf16 path where it costs nothing (channel-slice outputs instead of ``torch.cat``, nearest
x2 folded into the convolution's gather, max-pool as a second output of the BatchNorm pass, inference-mode BatchNorm
folded into the packed weights); BatchNorm statistics are a separate pass with fp64 accumulators.  There is no tape:
the mode exists to pin the forward arithmetic against the fp32 reference, training runs in bf16.
"""
import ctypes
from typing import Optional, Sequence, Tuple

import torch

from . import _lib as L
from . import engine as E

_c_void_p = ctypes.c_void_p
_MAX_SEGS = 4


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _coutp(cout: int) -> int:
    return (cout + 3) // 4 * 4


def pack(ctx, w: torch.Tensor, kind: str, cin_total: Optional[int] = None,
         row_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 parameter -> [T][cin_total][coutp] GEMM operand (uz_f32_pack_weight)."""
    w = w.detach()
    if not w.is_contiguous():
        w = w.contiguous()
    T = w.shape[2] * w.shape[3]
    if kind == "conv":       # OIHW
        cout, cin, s_co, s_ci = w.shape[0], w.shape[1], w.shape[1] * T, T
    elif kind == "convt":    # IOHW
        cin, cout, s_co, s_ci = w.shape[0], w.shape[1], T, w.shape[1] * T
    else:
        raise ValueError(kind)
    cin_total = cin if cin_total is None else cin_total
    out = torch.empty((T, cin_total, _coutp(cout)), dtype=torch.float32, device=w.device)
    rc = ctx.lib.uz_f32_pack_weight(w.data_ptr(), out.data_ptr(), T, cin, cin_total, cout, _coutp(cout), s_co, s_ci,
                                    _ptr(row_scale), ctx.stream)
    L.check(rc, "uz_f32_pack_weight")
    ctx.launches += 1
    return out


def conv(ctx, mode: int, segs: Sequence["E.Act"], wp: torch.Tensor, cout: int, bias: Optional[torch.Tensor],
         out_ptr: int, out_ld: int, flags: int = 0):
    """uz_f32_conv over a virtual concat of `segs`."""
    n = len(segs)
    if n > _MAX_SEGS:
        raise RuntimeError(f"unet_zoo_b200: fp32 convolution takes at most {_MAX_SEGS} concat segments (got {n})")
    ptrs = (_c_void_p * n)(*[s.ptr for s in segs])
    cs = (ctypes.c_int * n)(*[s.C for s in segs])
    lds = (ctypes.c_int * n)(*[s.ld for s in segs])
    a = segs[0]
    rc = ctx.lib.uz_f32_conv(mode, a.N, a.H, a.W, n, ptrs, cs, lds, wp.data_ptr(), cout, _coutp(cout), _ptr(bias),
                             out_ptr, out_ld, flags, ctx.stream)
    L.check(rc, "uz_f32_conv")
    ctx.launches += 1


def affine_act(ctx, z: "E.Act", scale_ptr, shift_ptr, relu: bool, y: Optional["E.Act"],
               pooled: Optional["E.Act"] = None, residual: Optional["E.Act"] = None):
    rc = ctx.lib.uz_f32_affine_act(z.ptr, z.ld, z.N, z.H, z.W, z.C, scale_ptr, shift_ptr, int(relu),
                                   residual.ptr if residual else None, residual.ld if residual else 0,
                                   y.ptr if y else None, y.ld if y else 0,
                                   pooled.ptr if pooled else None, pooled.ld if pooled else 0, ctx.stream)
    L.check(rc, "uz_f32_affine_act")
    ctx.launches += 1


def batch_stats(ctx, x: "E.Act", bn) -> torch.Tensor:
    """Training-mode statistics of `x`: bn[4][C] = {mean, invstd, scale, shift}; running statistics updated."""
    rows = ctx.reduce_rows
    stats = ctx.f32(rows, 2, x.C)
    L.check(ctx.lib.uz_f32_bn_stats(x.ptr, x.ld, x.npix, x.C, stats.data_ptr(), ctx.stream), "uz_f32_bn_stats")
    ctx.launches += 1
    bnbuf = ctx.f32(4, x.C)
    ctx.bn_finalize(stats, rows, x.C, x.npix, bn, bnbuf)
    return bnbuf


def _out_hw(mode: int, H: int, W: int) -> Tuple[int, int]:
    if mode in (L.CONV3X3_S2, L.CONV1X1_S2):
        return (H - 1) // 2 + 1, (W - 1) // 2 + 1
    if mode == L.UP2X_CONV3X3:
        return 2 * H, 2 * W
    return H, W


# ====================================================================================================== operators
def stem_input(ctx, x: torch.Tensor):
    """NCHW fp32 images -> NHWC fp32 with channels zero-padded to a multiple of 4 (16-byte pixel rows)."""
    N, cin, H, W = x.shape
    cpad = (cin + 3) // 4 * 4
    a = ctx.new_act(N, H, W, cpad)
    L.check(ctx.lib.uz_f32_nchw_to_nhwc(x.data_ptr(), N, cin, H, W, a.ptr, cpad, ctx.stream), "uz_f32_nchw_to_nhwc")
    ctx.launches += 1
    return a, "nhwc32"


def conv_bn_act(ctx, xs, conv_m, bn, out=None, pooled=None, relu=True, mode=L.CONV3X3, stem=None, need_dx=True):
    """fp32 counterpart of engine.conv_bn_act (DoubleConv / ConvBlock / VGGBlock halves, UpConvBlock, DownSample)."""
    a0 = xs[0]
    cout = conv_m.out_channels
    Ho, Wo = _out_hw(mode, a0.H, a0.W)
    if out is None:
        out = ctx.new_act(a0.N, Ho, Wo, cout)
    cin_total = sum(a.C for a in xs)
    if not ctx.training:
        scale, bias = E._fold_bn(conv_m.bias, bn) if bn is not None else (None, conv_m.bias)
        wp = pack(ctx, conv_m.weight, "conv", cin_total, row_scale=scale)
        conv(ctx, mode, xs, wp, cout, bias, out.ptr, out.ld, L.FLAG_RELU if relu else 0)
        if pooled is not None:
            affine_act(ctx, out, None, None, False, None, pooled)
        return out
    wp = pack(ctx, conv_m.weight, "conv", cin_total)
    z = ctx.new_act(a0.N, Ho, Wo, cout)
    conv(ctx, mode, xs, wp, cout, conv_m.bias, z.ptr, z.ld)
    bnbuf = batch_stats(ctx, z, bn)
    affine_act(ctx, z, bnbuf[2].data_ptr(), bnbuf[3].data_ptr(), relu, out, pooled)
    return out


def conv_transpose2x2(ctx, x, up, out):
    """ConvTranspose2d(k=2, s=2) into a channel slice of the concat buffer (common_layers.py:104,108,204)."""
    wp = pack(ctx, up.weight, "convt")
    conv(ctx, L.CONVT2X2, [x], wp, up.out_channels, up.bias, out.ptr, out.ld)
    return out


def head1x1(ctx, y, conv_m) -> torch.Tensor:
    """1x1 classifier head -> fp32 NCHW logits (OutConv, common_layers.py:125-128)."""
    K = conv_m.out_channels
    wp = pack(ctx, conv_m.weight, "conv")
    nhwc = torch.empty((y.N, y.H, y.W, K), dtype=torch.float32, device=ctx.device)
    conv(ctx, L.CONV1X1, [y], wp, K, conv_m.bias, nhwc.data_ptr(), K)
    if K == 1:
        logits = nhwc.view(y.N, 1, y.H, y.W)          # one class: NHWC and NCHW are the same memory
    else:
        logits = torch.empty((y.N, K, y.H, y.W), dtype=torch.float32, device=ctx.device)
        L.check(ctx.lib.uz_f32_nhwc_to_nchw(nhwc.data_ptr(), K, y.N, K, y.H, y.W, logits.data_ptr(), ctx.stream),
                "uz_f32_nhwc_to_nchw")
        ctx.launches += 1
    ctx.outputs.append(logits)
    return logits


def conv_bn_act_head(ctx, xs, conv_m, bn, head, relu=True) -> torch.Tensor:
    return head1x1(ctx, conv_bn_act(ctx, xs, conv_m, bn, relu=relu), head)


def bn_act(ctx, x, bn, relu=True):
    """[ReLU](BatchNorm2d(x)) - the pre-activation half of ResidualConv (common_layers.py:185-186,189-190)."""
    out = ctx.new_act(x.N, x.H, x.W, x.C)
    aff = batch_stats(ctx, x, bn) if ctx.training else E._running_affine(bn, ctx.device)
    affine_act(ctx, x, aff[2].data_ptr(), aff[3].data_ptr(), relu, out)
    return out


def conv_plain(ctx, xs, conv_m, out=None, mode=L.CONV3X3, accumulate=False, stem=None, need_dx=True):
    """Conv2d without normalisation, optionally out += conv(x) (common_layers.py:191,198; resunet.py:29-33,56)."""
    a0 = xs[0]
    cout = conv_m.out_channels
    Ho, Wo = _out_hw(mode, a0.H, a0.W)
    if out is None:
        if accumulate:
            raise ValueError("accumulate needs an existing output")
        out = ctx.new_act(a0.N, Ho, Wo, cout)
    wp = pack(ctx, conv_m.weight, "conv", sum(a.C for a in xs))
    conv(ctx, mode, xs, wp, cout, conv_m.bias, out.ptr, out.ld, L.FLAG_ACCUMULATE if accumulate else 0)
    return out


def bilinear_up2x(ctx, x, out):
    """nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True) into a concat slice (nested_unet.py:33)."""
    L.check(ctx.lib.uz_f32_bilinear2x(x.ptr, x.ld, x.N, x.H, x.W, x.C, out.ptr, out.ld, ctx.stream), "uz_f32_bilinear2x")
    ctx.launches += 1
    return out


def attention_gate(ctx, g, x, att, out):
    """AttentionBlock.forward (attention_unet.py:34-40): out = x * sigmoid(BN(w_psi . ReLU(BN(W_g g) + BN(W_x x))))."""
    conv_g, bn_g = att.w_g[0], att.w_g[1]
    conv_x, bn_x = att.w_x[0], att.w_x[1]
    conv_p, bn_p = att.psi[0], att.psi[1]
    F = conv_g.out_channels
    zg, zx = ctx.new_act(g.N, g.H, g.W, F), ctx.new_act(x.N, x.H, x.W, F)
    conv(ctx, L.CONV1X1, [g], pack(ctx, conv_g.weight, "conv"), F, conv_g.bias, zg.ptr, zg.ld)
    conv(ctx, L.CONV1X1, [x], pack(ctx, conv_x.weight, "conv"), F, conv_x.bias, zx.ptr, zx.ld)
    if ctx.training:
        bg, bx = batch_stats(ctx, zg, bn_g), batch_stats(ctx, zx, bn_x)
    else:
        bg, bx = E._running_affine(bn_g, ctx.device), E._running_affine(bn_x, ctx.device)
    g1 = ctx.new_act(g.N, g.H, g.W, F)
    affine_act(ctx, zg, bg[2].data_ptr(), bg[3].data_ptr(), False, g1)
    r = zx                                                        # r = ReLU(g1 + BN(zx)), in place over zx
    affine_act(ctx, zx, bx[2].data_ptr(), bx[3].data_ptr(), True, r, residual=g1)
    psi = ctx.new_act(x.N, x.H, x.W, 1)
    conv(ctx, L.CONV1X1, [r], pack(ctx, conv_p.weight, "conv"), 1, conv_p.bias, psi.ptr, psi.ld)
    bp = batch_stats(ctx, psi, bn_p) if ctx.training else E._running_affine(bn_p, ctx.device)
    L.check(ctx.lib.uz_f32_gate_apply(x.ptr, x.ld, psi.ptr, bp.data_ptr(), x.npix, x.C, out.ptr, out.ld, ctx.stream),
            "uz_f32_gate_apply")
    ctx.launches += 1
    return out
