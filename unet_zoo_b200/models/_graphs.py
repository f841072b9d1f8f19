"""Forward graphs of `attention_unet`, `nested_unet` and `resunet` on the fused sm_100a operators.

Each function mirrors one reference ``forward`` (cited below) but never materialises ``torch.cat`` / ``nn.Upsample``
/ ``nn.MaxPool2d`` results as separate passes: producers write channel slices of pre-allocated wide NHWC buffers,
nearest-x2 upsampling is folded into the following convolution's gather, pooling is a second output of the
BatchNorm-apply pass.
"""
from .. import _lib as L
from .. import engine as E


def attention_unet_graph(m, ctx: E.Ctx, x):
    """AttentionUNet.forward (attention_unet.py:73-110).  Decoder level buffer = [gated skip | upsampled]
    (attention_unet.py:90,95,100,105)."""
    N, _, H, W = x.shape
    widths = (64, 128, 256, 512, 1024)
    cur, stem = E.stem_input(ctx, x)
    feats = []
    for i, w in enumerate(widths):
        (c1, b1), (c2, b2) = getattr(m, f"conv{i + 1}").units()
        h = E.conv_bn_act(ctx, [cur], c1, b1, stem=stem if i == 0 else None, need_dx=i > 0)
        pooled = ctx.new_act(N, H >> (i + 1), W >> (i + 1), w) if i < 4 else None   # maxpool, attention_unet.py:77-87
        feats.append(E.conv_bn_act(ctx, [h], c2, b2, pooled=pooled))
        cur = pooled
    d = feats[4]
    for lvl in (5, 4, 3, 2):
        w = widths[lvl - 2]
        skip = feats[lvl - 2]
        cat = ctx.new_buf(N, skip.H, skip.W, 2 * w)
        up = getattr(m, f"up{lvl}").up                       # Sequential(Upsample, Conv2d, BatchNorm2d, ReLU)
        g = E.conv_bn_act(ctx, [d], up[1], up[2], out=E.Act(cat, w, w), mode=L.UP2X_CONV3X3)
        E.attention_gate(ctx, g, skip, getattr(m, f"att{lvl}"), E.Act(cat, 0, w))
        (c1, b1), (c2, b2) = getattr(m, f"upconv{lvl}").units()
        h = E.conv_bn_act(ctx, [E.Act(cat)], c1, b1)
        if lvl > 2:
            d = E.conv_bn_act(ctx, [h], c2, b2)
    E.conv_bn_act_head(ctx, [h], c2, b2, m.conv_1x1)          # attention_unet.py:106-108, activation never stored
    return ctx.outputs


def nested_unet_graph(m, ctx: E.Ctx, x):
    """NestedUNet.forward (nested_unet.py:71-106).  Level i owns one buffer laid out
    [x_i0 | x_i1 | ... | x_i(4-i) | up_1 | ... | up_(4-i)]: node x_ij reads the contiguous run x_i0..x_i(j-1) plus the
    slot holding bilinear-up(x_(i+1)(j-1)) - the reference's cat order (nested_unet.py:82-93) without any copy."""
    N, _, H, W = x.shape
    nb = (32, 64, 128, 256, 512)
    bufs = [ctx.new_buf(N, H >> i, W >> i, nb[i] * (5 - i) + (nb[i + 1] * (4 - i) if i < 4 else 0)) for i in range(5)]

    def node(i, j):
        return E.Act(bufs[i], nb[i] * j, nb[i])

    def upslot(i, j):
        return E.Act(bufs[i], nb[i] * (5 - i) + nb[i + 1] * (j - 1), nb[i + 1])

    def vgg(blk, xs, out, pooled=None, stem=None, need_dx=True, head=None):   # VGGBlock.forward, nested_unet.py:13-22
        (c1, b1), (c2, b2) = blk.units()
        h = E.conv_bn_act(ctx, xs, c1, b1, stem=stem, need_dx=need_dx)
        if head is not None:                                              # x0_4 feeds `final` only (nested_unet.py:104)
            return E.conv_bn_act_head(ctx, [h], c2, b2, head)
        return E.conv_bn_act(ctx, [h], c2, b2, out=out, pooled=pooled)

    cur, stem = E.stem_input(ctx, x)
    for i in range(5):                                                    # backbone column, nested_unet.py:73-79
        pooled = ctx.new_act(N, H >> (i + 1), W >> (i + 1), nb[i]) if i < 4 else None
        vgg(getattr(m, f"conv{i}_0"), [cur], node(i, 0), pooled=pooled, stem=stem if i == 0 else None, need_dx=i > 0)
        cur = pooled
    for j in range(1, 5):                                                 # dense columns, nested_unet.py:81-93
        for i in range(0, 5 - j):
            up = E.bilinear_up2x(ctx, node(i + 1, j - 1), upslot(i, j))
            vgg(getattr(m, f"conv{i}_{j}"), [E.Act(bufs[i], 0, nb[i] * j), up], node(i, j),
                head=m.final if (j == 4 and not m.deep_supervision) else None)
    if m.deep_supervision:                                                # nested_unet.py:95-102
        for k in range(1, 5):
            E.head1x1(ctx, node(0, k), getattr(m, f"final{k}"))
    return ctx.outputs


def _residual_conv(ctx: E.Ctx, x: E.Act, blk, out=None) -> E.Act:
    """ResidualConv.forward (common_layers.py:182-199): conv_block(x) + conv_skip(x).
    The skip branch (conv1x1(s) -> BN) is produced first; the block's last convolution accumulates onto it."""
    s2 = blk.stride == 2
    bn1, _, conv1, bn2, _, conv2 = blk.conv_block
    conv_s, bn_s = blk.conv_skip
    if not ctx.training and not ctx.fp32:
        # inference: the skip branch rides along in the last convolution's GEMM (engine.residual_conv_tail_fused)
        hh = E.conv_bn_act(ctx, [E.bn_act(ctx, x, bn1)], conv1, bn2, mode=L.CONV3X3_S2 if s2 else L.CONV3X3)
        fused = E.residual_conv_tail_fused(ctx, hh, x, conv2, conv_s, bn_s, out, s2)
        if fused is not None:
            return fused
        out = E.conv_bn_act(ctx, [x], conv_s, bn_s, out=out, relu=False, mode=L.CONV1X1_S2 if s2 else L.CONV1X1)
        return E.conv_plain(ctx, [hh], conv2, out=out, accumulate=True)
    out = E.conv_bn_act(ctx, [x], conv_s, bn_s, out=out, relu=False, mode=L.CONV1X1_S2 if s2 else L.CONV1X1)
    h = E.bn_act(ctx, x, bn1)
    h = E.conv_bn_act(ctx, [h], conv1, bn2, mode=L.CONV3X3_S2 if s2 else L.CONV3X3)
    return E.conv_plain(ctx, [h], conv2, out=out, accumulate=True)


def resunet_graph(m, ctx: E.Ctx, x):
    """ResUnet.forward (resunet.py:54-79).  Decoder level buffers are [upsampled | encoder skip] (resunet.py:63,68,73)."""
    N, _, H, W = x.shape
    f = [m.input_layer[0].out_channels, m.residual_conv_1.conv_block[2].out_channels,
         m.residual_conv_2.conv_block[2].out_channels, m.bridge.conv_block[2].out_channels]
    cats = [ctx.new_buf(N, H >> i, W >> i, 2 * f[i]) for i in range(3)]
    cur, stem = E.stem_input(ctx, x)
    h = E.conv_bn_act(ctx, [cur], m.input_layer[0], m.input_layer[1], stem=stem, need_dx=False)
    x1 = E.Act(cats[0], f[0], f[0])
    fused = None
    if not ctx.training and not ctx.fp32 and stem == "im2col":      # inference: both stem convolutions in one GEMM
        fused = E.stem_pair_fused(ctx, h, m.input_layer[3], cur, m.input_skip[0], x1)
    if fused is None:
        E.conv_plain(ctx, [h], m.input_layer[3], out=x1)                                    # resunet.py:56
        E.conv_plain(ctx, [cur], m.input_skip[0], out=x1, accumulate=True, stem=stem, need_dx=False)
    x2 = _residual_conv(ctx, x1, m.residual_conv_1, E.Act(cats[1], f[1], f[1]))
    x3 = _residual_conv(ctx, x2, m.residual_conv_2, E.Act(cats[2], f[2], f[2]))
    x4 = _residual_conv(ctx, x3, m.bridge)
    E.conv_transpose2x2(ctx, x4, m.upsample_1.upsample, E.Act(cats[2], 0, f[2]))
    x6 = _residual_conv(ctx, E.Act(cats[2]), m.up_residual_conv1)
    E.conv_transpose2x2(ctx, x6, m.upsample_2.upsample, E.Act(cats[1], 0, f[1]))
    x8 = _residual_conv(ctx, E.Act(cats[1]), m.up_residual_conv2)
    E.conv_transpose2x2(ctx, x8, m.upsample_3.upsample, E.Act(cats[0], 0, f[0]))
    x10 = _residual_conv(ctx, E.Act(cats[0]), m.up_residual_conv3)
    E.head1x1(ctx, x10, m.output_layer[0])
    return ctx.outputs
