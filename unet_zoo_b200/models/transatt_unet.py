"""`transatt_unet`: a UNet whose encoder / decoder are the hot-path blocks and whose 1/16-scale
bottleneck carries two attention branches.

Parameters and wiring follow the reference (unet_zoo/models/transatt_unet.py:109-164; blocks
common_layers.py:130-180): ``inc`` + four ``Down`` (MaxPool2d -> DoubleConvo), four bilinear ``Up`` blocks whose
concat order is **[skip, upsampled]** (common_layers.py:179) and whose DoubleConvo narrows through
``in_channels // 2``, ``OutConv``.  All of that - 99.3 % of the FLOPs at 512^2 - runs on the fused sm_100a operators:
the skip half of each decoder buffer is written by the encoder's BatchNorm pass (together with the pooled tensor), the
upsampled half by ``uz_bilinear2x_fwd``; no cat / pool / pad kernels run.

The bottleneck (learned position embedding + PAM spatial attention + channel self-attention over a 512 x (H/16 * W/16)
map, transatt_unet.py:143-154) is an island of stock torch ops (``engine.torch_island``): batched matmuls and
softmaxes on a tensor 1/256 of the input resolution, outside the path BASELINE.json names.  Its three modules below
hold the reference's parameter names and restate its math.

``bilinear=False`` is rejected: in the reference that configuration cannot run either (``Up`` builds
``ConvTranspose2d(1536, 768)`` for a 1024-channel input and the 512-channel position embedding is added to a
1024-channel map, transatt_unet.py:121-135).
"""
import os

import torch
import torch.nn as nn

from .. import engine as E
from ._runtime import HotPathModel
from .blocks import OutConv, _Container, _cbr


class DoubleConvo(_Container):         # common_layers.py:130-147
    def __init__(self, in_channels, out_channels, mid_channels=None):
        super().__init__()
        mid = mid_channels or out_channels
        first, second = _cbr(in_channels, mid, n=1), _cbr(mid, out_channels, n=1)
        self.double_conv = nn.Sequential(*first, *second)

    def units(self):
        s = self.double_conv
        return (s[0], s[1]), (s[3], s[4])


class Down(_Container):                # common_layers.py:149-159
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.maxpool_conv = nn.Sequential(nn.MaxPool2d(2), DoubleConvo(in_channels, out_channels))


class Up(_Container):                  # common_layers.py:161-180 (bilinear branch)
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.up = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True)
        self.conv = DoubleConvo(in_channels, out_channels, in_channels // 2)


class PositionEmbeddingLearned(nn.Module):
    """Row / column embeddings concatenated along channels (transatt_unet.py:57-82): channel c < F of pixel (y, x) is
    col_embed[x, c], channel F + c is row_embed[y, c]."""

    def __init__(self, num_pos_feats=256, len_embedding=32):
        super().__init__()
        self.row_embed = nn.Embedding(len_embedding, num_pos_feats)
        self.col_embed = nn.Embedding(len_embedding, num_pos_feats)
        nn.init.uniform_(self.row_embed.weight)
        nn.init.uniform_(self.col_embed.weight)

    def forward(self, x):
        n, _, h, w = x.shape
        if h > self.row_embed.num_embeddings or w > self.col_embed.num_embeddings:
            raise RuntimeError(f"TransAttUNet: the bottleneck map {h}x{w} exceeds the learned position table "
                               f"({self.row_embed.num_embeddings}); inputs up to 512x512 are supported, as in the reference")
        cols = self.col_embed.weight[:w].t()[:, None, :].expand(-1, h, -1)      # [F, h, w], constant over rows
        rows = self.row_embed.weight[:h].t()[:, :, None].expand(-1, -1, w)      # [F, h, w], constant over columns
        return torch.cat([cols, rows], 0)[None].expand(n, -1, -1, -1)


class PAM_Module(nn.Module):
    """Spatial (position) attention (transatt_unet.py:26-55): softmax(Q^T K) over pixels applied to V, scaled by a
    learned gamma (zero at init) and added to the input."""

    def __init__(self, in_dim):
        super().__init__()
        self.chanel_in = in_dim
        self.query_conv = nn.Conv2d(in_dim, in_dim // 8, kernel_size=1)
        self.key_conv = nn.Conv2d(in_dim, in_dim // 8, kernel_size=1)
        self.value_conv = nn.Conv2d(in_dim, in_dim, kernel_size=1)
        self.gamma = nn.Parameter(torch.zeros(1))
        self.softmax = nn.Softmax(dim=-1)

    def forward(self, x):
        n, c, h, w = x.shape
        q = self.query_conv(x).flatten(2)                       # [n, c/8, hw]
        k = self.key_conv(x).flatten(2)
        v = self.value_conv(x).flatten(2)                       # [n, c, hw]
        att = self.softmax(q.transpose(1, 2) @ k)               # [n, hw, hw], rows sum to one
        return self.gamma * (v @ att.transpose(1, 2)).view(n, c, h, w) + x


class ScaledDotProductAttention(nn.Module):
    """Channel self-attention with q = k = v = the flattened map (transatt_unet.py:84-107): softmax over channels of
    (x x^T / sqrt(temperature)), dropout, applied back to x.  No parameters."""

    def __init__(self, temperature, attn_dropout=0.1):
        super().__init__()
        self.temperature = temperature ** 0.5
        self.dropout = nn.Dropout(attn_dropout)

    def forward(self, x):
        n, d, h, w = x.shape
        f = x.flatten(2)                                        # [n, d, hw]
        att = self.dropout(torch.softmax((f / self.temperature) @ f.transpose(1, 2), dim=-1))
        return (att @ f).view(n, d, h, w)


class TransAttUNet(HotPathModel):
    _any_size = True      # UpSample_UNet / Up pad the upsampled tensor to odd skip sizes

    def __init__(self, in_channels=3, num_classes=1, bilinear=True, **kwargs):
        super().__init__()
        if not bilinear:
            raise NotImplementedError("TransAttUNet(bilinear=False): the reference's transposed-convolution variant "
                                      "cannot run (channel mismatch in Up / the position embedding); use bilinear=True")
        self._in_channels = in_channels
        self.n_channels, self.n_classes, self.bilinear = in_channels, num_classes, bilinear
        self.inc = DoubleConvo(in_channels, 64)
        self.down1, self.down2, self.down3, self.down4 = Down(64, 128), Down(128, 256), Down(256, 512), Down(512, 512)
        self.up1, self.up2, self.up3, self.up4 = Up(1024, 256), Up(512, 128), Up(256, 64), Up(128, 64)
        self.outc = OutConv(64, num_classes)
        self.pos = PositionEmbeddingLearned(256)
        self.pam = PAM_Module(512)
        # the reference swallows **kwargs; `attn_dropout` is honoured here so that parity tests can switch the
        # (RNG-dependent) attention dropout off on both sides
        self.sdpa = ScaledDotProductAttention(512, attn_dropout=kwargs.get("attn_dropout", 0.1))

    def _check_input(self, x):
        # the learned position table has 32 rows / columns for the 1/16-scale map (transatt_unet.py:59,132): say so
        # before any kernel runs instead of failing inside the bottleneck
        if isinstance(x, torch.Tensor) and x.dim() == 4:
            h, w = (x.shape[1], x.shape[2]) if x.dtype == torch.uint8 else (x.shape[2], x.shape[3])
            lim = 16 * self.pos.row_embed.num_embeddings
            if h > lim or w > lim:
                raise RuntimeError(f"TransAttUNet: inputs up to {lim}x{lim} are supported (got {h}x{w}): the reference's "
                                   f"learned position embedding has {self.pos.row_embed.num_embeddings} entries per axis")
        return super()._check_input(x)

    def _bottleneck(self, x5):
        """transatt_unet.py:143-154 on an NCHW fp32 tensor (stock torch): the fp32 parity mode's path."""
        x5 = x5 + self.pos(x5)
        return self.sdpa(x5) + self.pam(x5)

    def _bottleneck_tokens(self, t, h, w):
        """The same bottleneck on bf16 tokens t [N, h*w, 512] with the library's own kernels (attention.py): every
        matrix product is uz_bmm_tn, both softmaxes uz_softmax_rows.
        Position embedding (transatt_unet.py:69-82), channel self-attention with q = k = v (:91-107), PAM (:40-55)."""
        from .. import attention as A
        n, hw, d = t.shape
        if h > self.pos.row_embed.num_embeddings or w > self.pos.col_embed.num_embeddings:
            raise RuntimeError(f"TransAttUNet: the bottleneck map {h}x{w} exceeds the learned position table")
        cols = self.pos.col_embed.weight[:w][None, :, :].expand(h, w, -1)           # [h, w, F], constant over rows
        rows = self.pos.row_embed.weight[:h][:, None, :].expand(h, w, -1)           # constant over columns
        pos = torch.cat([cols, rows], dim=-1).reshape(1, hw, d)
        t = (t.float() + pos).to(torch.bfloat16)                                     # x5 + pos
        # channel attention: att[a, b] = softmax_b(sum_p f[a, p] f[b, p] / sqrt(T)), out[p, a] = sum_b att[a, b] f[b, p]
        f = A.transpose(t)                                                           # [N, d, hw]
        att = A.softmax_rows(A.bmm_tn(f, f), 1.0 / self.sdpa.temperature)
        att = self.sdpa.dropout(att)
        sdpa = A.bmm_tn(t, att)                                                      # [N, hw, d]
        # PAM: energy[i, j] = q_i . k_j, att = softmax_j, out[i, c] = sum_j att[i, j] v[j, c]
        q = A.linear_tokens(t, self.pam.query_conv.weight, self.pam.query_conv.bias)
        k = A.linear_tokens(t, self.pam.key_conv.weight, self.pam.key_conv.bias)
        v = A.linear_tokens(t, self.pam.value_conv.weight, self.pam.value_conv.bias)
        patt = A.softmax_rows(A.bmm_tn(q, k), 1.0)                                   # [N, hw, hw]
        pam = A.bmm_tn(patt, A.transpose(v))                                         # [N, hw, d]
        return (self.pam.gamma * pam.float() + t.float() + sdpa.float()).to(torch.bfloat16)

    def _graph(self, ctx: E.Ctx, x):
        N, _, H, W = x.shape
        widths = (64, 128, 256, 512)
        # decoder level i owns [skip x_(i+1) | bilinear-up of the level below]
        below = (64, 128, 256, 512)
        cats = [ctx.new_buf(N, H >> i, W >> i, widths[i] + below[i]) for i in range(4)]
        cur, stem = E.stem_input(ctx, x)
        blocks = [self.inc, self.down1.maxpool_conv[1], self.down2.maxpool_conv[1], self.down3.maxpool_conv[1]]
        for i, w in enumerate(widths):
            (c1, b1), (c2, b2) = blocks[i].units()
            h = E.conv_bn_act(ctx, [cur], c1, b1, stem=stem if i == 0 else None, need_dx=i > 0)
            pooled = ctx.new_act(N, H >> (i + 1), W >> (i + 1), w)           # the MaxPool2d of the next Down
            E.conv_bn_act(ctx, [h], c2, b2, out=E.Act(cats[i], 0, w), pooled=pooled)
            cur = pooled
        (c1, b1), (c2, b2) = self.down4.maxpool_conv[1].units()
        x5 = E.conv_bn_act(ctx, [E.conv_bn_act(ctx, [cur], c1, b1)], c2, b2)
        island = [p for m in (self.pos, self.pam) for p in m.parameters()]
        if ctx.fp32 or os.environ.get("UZ_TRANSATT_TORCH_ISLAND"):
            cur = E.torch_island(ctx, x5, self._bottleneck, island)      # fp32 parity mode: strict fp32 torch math
        else:
            cur = E.token_island(ctx, x5, lambda t: self._bottleneck_tokens(t, x5.H, x5.W), island)
        for j, up in enumerate((self.up1, self.up2, self.up3, self.up4)):
            lvl = 3 - j
            E.bilinear_up2x(ctx, cur, E.Act(cats[lvl], widths[lvl], below[lvl]))
            (c1, b1), (c2, b2) = up.conv.units()
            h = E.conv_bn_act(ctx, [E.Act(cats[lvl])], c1, b1)
            if lvl > 0:
                cur = E.conv_bn_act(ctx, [h], c2, b2)
        E.conv_bn_act_head(ctx, [h], c2, b2, self.outc.conv)                 # transatt_unet.py:161-163
        return ctx.outputs
