"""`unet_transformer`: U-Net with a self-attention bottleneck and cross-attention decoders.

Parameters, buffers and wiring follow the reference (unet_zoo/models/unet_transformer.py:10-33,68-113,115-281; blocks
common_layers.py:20-37,130-159): ``inc`` (DoubleConv), three ``Down`` (MaxPool2d -> DoubleConvo), ``MHSA`` on the 1/8-scale
map, three ``TransformerUp`` (``MHCA`` + two conv3x3-BN-ReLU), ``OutConv``; 209 state_dict entries, bit-identical default
initialisation under the same seed.

Everything runs on the kernel library: the convolutions / BatchNorm / pooling / bilinear x2 on the fused operators the
four hot-path models use, the positional-encoding add (+ MaxPool2d of the sum), adaptive average pooling and the
bilinear resize from the attention resolution on the bandwidth kernels of ``csrc/transformer_ops.cu``, and the attention
products on ``uz_bmm_tn`` / ``uz_softmax_rows`` (``unet_zoo_b200/attention.py``) inside ``engine.token_island``.
The reference's softmax runs over dim=1 of the ``[b, queries, keys]`` score matrix - the QUERY axis (:130,216):
``uz_softmax_cols`` normalises every column over its rows, so the 4096 x 4096 matrix is never transposed in the forward.

bf16 mode only (there is no fp32 parity path for the attention operators); H and W must be multiples of 8.
"""
import math

import torch
import torch.nn as nn

from .. import _lib as L
from .. import engine as E
from ._runtime import HotPathModel
from .blocks import DoubleConv, OutConv, _Container
from .transatt_unet import Down


class MultiHeadDense(_Container):                    # unet_transformer.py:10-33: y = x @ weight, no bias
    def __init__(self, d):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(d, d))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))


class PositionalEncoding2D(_Container):              # unet_transformer.py:68-102
    def __init__(self, channels):
        super().__init__()
        channels = int(math.ceil(channels / 2))
        self.channels = channels
        self.register_buffer("inv_freq", 1.0 / (10000 ** (torch.arange(0, channels, 2).float() / channels)))


class PositionalEncodingPermute2D(_Container):       # unet_transformer.py:104-115
    def __init__(self, channels):
        super().__init__()
        self.penc = PositionalEncoding2D(channels)

    def table(self, c: int, h: int, w: int) -> torch.Tensor:
        """The encoding as an fp32 [h, w, c] table (channels last = the activation layout): the first `channels`
        features carry the row index (sin | cos halves), the next `channels` the column index (:83-101)."""
        f = self.penc.inv_freq.float()
        ch = self.penc.channels
        ang_h = torch.arange(h, device=f.device, dtype=f.dtype)[:, None] * f[None]
        ang_w = torch.arange(w, device=f.device, dtype=f.dtype)[:, None] * f[None]
        emb_h = torch.cat([ang_h.sin(), ang_h.cos()], dim=-1)            # [h, 2 * len(f)]
        emb_w = torch.cat([ang_w.sin(), ang_w.cos()], dim=-1)
        emb = torch.zeros(h, w, 2 * ch, device=f.device, dtype=torch.float32)
        emb[:, :, :emb_h.shape[1]] = emb_h[:, None, :]
        emb[:, :, ch:ch + emb_w.shape[1]] = emb_w[None, :, :]
        return emb[:, :, :c].contiguous()


class MultiHeadSelfAttention(_Container):            # unet_transformer.py:117-136
    def __init__(self, channel):
        super().__init__()
        self.query, self.key, self.value = MultiHeadDense(channel), MultiHeadDense(channel), MultiHeadDense(channel)
        self.softmax = nn.Softmax(dim=1)
        self.pe = PositionalEncodingPermute2D(channel)


class MultiHeadCrossAttention(_Container):           # unet_transformer.py:138-229
    def __init__(self, channelY, channelS, common_attn_res_for_QK_V=(64, 64)):
        super().__init__()
        cs = channelS
        self.common_attn_channels = cs
        self.common_attn_res_for_QK_V = tuple(common_attn_res_for_QK_V)
        self.Sconv_process = nn.Sequential(nn.MaxPool2d(2), nn.Conv2d(cs, cs, kernel_size=1), nn.BatchNorm2d(cs),
                                           nn.ReLU(inplace=True))
        self.Yconv_process = nn.Sequential(nn.Conv2d(channelY, cs, kernel_size=1), nn.BatchNorm2d(cs), nn.ReLU(inplace=True))
        self.query, self.key, self.value = MultiHeadDense(cs), MultiHeadDense(cs), MultiHeadDense(cs)
        self.conv_after_attention = nn.Sequential(nn.Conv2d(cs, cs, kernel_size=1), nn.BatchNorm2d(cs), nn.ReLU(inplace=True))
        self.Yconv2_process = nn.Sequential(nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True),
                                            nn.Conv2d(channelY, channelY, kernel_size=3, padding=1),
                                            nn.Conv2d(channelY, cs, kernel_size=1), nn.BatchNorm2d(cs), nn.ReLU(inplace=True))
        self.softmax = nn.Softmax(dim=1)
        self.Spe = PositionalEncodingPermute2D(channelS)
        self.Ype = PositionalEncodingPermute2D(channelY)


class TransformerUp(_Container):                     # unet_transformer.py:231-254
    def __init__(self, Ychannels, Schannels, common_attn_res_for_QK_V=(64, 64)):
        super().__init__()
        self.MHCA = MultiHeadCrossAttention(Ychannels, Schannels, common_attn_res_for_QK_V)
        s = Schannels
        self.conv = nn.Sequential(nn.Conv2d(s * 2, s, kernel_size=3, stride=1, padding=1, bias=True), nn.BatchNorm2d(s),
                                  nn.ReLU(inplace=True),
                                  nn.Conv2d(s, s, kernel_size=3, stride=1, padding=1, bias=True), nn.BatchNorm2d(s),
                                  nn.ReLU(inplace=True))


def _attention(q_src, v_src, att, channels):
    """softmax(Q K^T / sqrt(c), dim=1) V on bf16 tokens (unet_transformer.py:129-135, :209-217): Q, K from `q_src`, V from
    `v_src`, MultiHeadDense = x @ weight."""
    from .. import attention as A
    q = A.linear_tokens(q_src, att.query.weight.t())
    k = A.linear_tokens(q_src, att.key.weight.t())
    v = A.linear_tokens(v_src, att.value.weight.t())
    a = A.softmax_cols(A.bmm_tn(q, k), 1.0 / math.sqrt(channels))        # [N, queries, keys], normalised over the QUERIES
    return A.bmm_tn(a, A.transpose(v))                                   # out[i, c] = sum_j A[i, j] V[j, c]


class U_Transformer(HotPathModel):
    _spatial_multiple = 8

    def __init__(self, in_channels=3, num_classes=1, bilinear=True, common_attn_res_for_QK_V=(64, 64), **kwargs):
        super().__init__()
        self._in_channels = in_channels
        self.in_channels, self.classes, self.bilinear = in_channels, num_classes, bilinear
        self.inc = DoubleConv(in_channels, 64)
        self.down1, self.down2, self.down3 = Down(64, 128), Down(128, 256), Down(256, 512)
        self.MHSA = MultiHeadSelfAttention(512)
        self.up1 = TransformerUp(512, 256, common_attn_res_for_QK_V)
        self.up2 = TransformerUp(256, 128, common_attn_res_for_QK_V)
        self.up3 = TransformerUp(128, 64, common_attn_res_for_QK_V)
        self.outc = OutConv(64, num_classes)
        self._pe_tables = {}

    def set_precision(self, precision: str):
        if precision != "bf16":
            raise ValueError("U_Transformer: only the bf16 mode exists (no fp32 parity path for the attention operators)")
        return super().set_precision(precision)

    def _pe(self, mod: PositionalEncodingPermute2D, c: int, h: int, w: int) -> torch.Tensor:
        key = (id(mod), c, h, w, str(mod.penc.inv_freq.device))
        t = self._pe_tables.get(key)
        if t is None:
            with torch.no_grad():
                t = self._pe_tables[key] = mod.table(c, h, w)
        return t

    def _up(self, ctx: E.Ctx, up: TransformerUp, y: E.Act, s: E.Act, head=None):
        """TransformerUp.forward (unet_transformer.py:251-254) = MHCA.forward (:186-229) + two conv3x3-BN-ReLU."""
        m = up.MHCA
        cs, cy = m.common_attn_channels, y.C
        ah, aw = m.common_attn_res_for_QK_V
        N = y.N
        _, s_pool = E.add_posenc(ctx, s, self._pe(m.Spe, s.C, s.H, s.W), pooled=ctx.new_act(N, s.H // 2, s.W // 2, s.C))
        s_p = E.conv_bn_act(ctx, [s_pool], m.Sconv_process[1], m.Sconv_process[2], mode=L.CONV1X1)
        y_pe, _ = E.add_posenc(ctx, y, self._pe(m.Ype, cy, y.H, y.W))
        y_p = E.conv_bn_act(ctx, [y_pe], m.Yconv_process[0], m.Yconv_process[1], mode=L.CONV1X1)
        qk = E.adaptive_avgpool(ctx, y_p, ah, aw)
        vs = E.adaptive_avgpool(ctx, s_p, ah, aw)
        z = E.token_island(ctx, [qk, vs], lambda a, b: _attention(a, b, m, cs), [m.query.weight, m.key.weight, m.value.weight])
        cat = ctx.new_buf(N, 2 * y.H, 2 * y.W, 2 * cs)                       # torch.cat([Z_attn, Y2_processed], 1), :228
        zr = E.bilinear_resize(ctx, z, 2 * y.H, 2 * y.W)
        E.conv_bn_act(ctx, [zr], m.conv_after_attention[0], m.conv_after_attention[1], out=E.Act(cat, 0, cs), mode=L.CONV1X1)
        y_up = E.bilinear_up2x(ctx, y_pe, ctx.new_act(N, 2 * y.H, 2 * y.W, cy))
        y2 = E.conv_plain(ctx, [y_up], m.Yconv2_process[1])
        E.conv_bn_act(ctx, [y2], m.Yconv2_process[2], m.Yconv2_process[3], out=E.Act(cat, cs, cs), mode=L.CONV1X1)
        h = E.conv_bn_act(ctx, [E.Act(cat)], up.conv[0], up.conv[1])
        if head is not None:
            return E.conv_bn_act_head(ctx, [h], up.conv[3], up.conv[4], head)
        return E.conv_bn_act(ctx, [h], up.conv[3], up.conv[4])

    def _graph(self, ctx: E.Ctx, x):
        if ctx.fp32:
            raise RuntimeError("U_Transformer: only the bf16 mode exists")
        N, _, H, W = x.shape
        cur, stem = E.stem_input(ctx, x)
        skips = []
        blocks = [self.inc, self.down1.maxpool_conv[1], self.down2.maxpool_conv[1], self.down3.maxpool_conv[1]]
        for i, (blk, w) in enumerate(zip(blocks, (64, 128, 256, 512))):
            (c1, b1), (c2, b2) = blk.units()
            h = E.conv_bn_act(ctx, [cur], c1, b1, stem=stem if i == 0 else None, need_dx=i > 0)
            pooled = ctx.new_act(N, H >> (i + 1), W >> (i + 1), w) if i < 3 else None     # the next Down's MaxPool2d
            skips.append(E.conv_bn_act(ctx, [h], c2, b2, pooled=pooled))
            cur = pooled
        x1, x2, x3, x4 = skips
        t, _ = E.add_posenc(ctx, x4, self._pe(self.MHSA.pe, 512, x4.H, x4.W))               # MHSA.forward, :126-136
        y = E.token_island(ctx, t, lambda a: _attention(a, a, self.MHSA, 512),
                           [self.MHSA.query.weight, self.MHSA.key.weight, self.MHSA.value.weight])
        y = self._up(ctx, self.up1, y, x3)
        y = self._up(ctx, self.up2, y, x2)
        self._up(ctx, self.up3, y, x1, head=self.outc.conv)
        return ctx.outputs
