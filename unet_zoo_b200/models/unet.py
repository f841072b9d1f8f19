"""`unet`: 4x(DoubleConv+MaxPool) -> DoubleConv(512,1024) -> 4x(ConvTranspose2d, cat, DoubleConv) -> 1x1 head.

Same parameters and wiring as the reference (unet_zoo/models/unet.py:12-44), executed by the fused sm_100a operators:
each level owns one NHWC buffer of 2C channels laid out [upsampled | skip] (common_layers.py:115 concatenates the
upsampled tensor first); the encoder's second conv unit writes its activation into the skip half (and the pooled tensor
for the next level), the decoder's transposed conv scatters into the other half - no cat / pad / pool kernels run.
"""
from .. import engine as E
from ._runtime import HotPathModel
from .blocks import DoubleConv, DownSample, OutConv, UpSample_UNet


class UNet(HotPathModel):
    _any_size = True      # UpSample_UNet / Up pad the upsampled tensor to odd skip sizes

    def __init__(self, in_channels: int, num_classes: int):
        super().__init__()
        self._in_channels = in_channels
        widths = (64, 128, 256, 512)
        cin = in_channels
        for i, w in enumerate(widths, start=1):
            setattr(self, f"down_convolution_{i}", DownSample(cin, w))
            cin = w
        self.bottle_neck = DoubleConv(512, 1024)
        for i, w in enumerate(reversed(widths), start=1):
            setattr(self, f"up_convolution_{i}", UpSample_UNet(2 * w, w))
        self.out = OutConv(in_channels=64, out_channels=num_classes)

    def _graph(self, ctx: E.Ctx, x):
        N, _, H, W = x.shape
        widths = (64, 128, 256, 512)
        cats = [ctx.new_buf(N, H >> i, W >> i, 2 * w) for i, w in enumerate(widths)]
        cur, stem = E.stem_input(ctx, x)
        for i, w in enumerate(widths):
            (c1, b1), (c2, b2) = getattr(self, f"down_convolution_{i + 1}").conv.units()
            h = E.conv_bn_act(ctx, [cur], c1, b1, stem=stem if i == 0 else None, need_dx=i > 0)
            pooled = ctx.new_act(N, H >> (i + 1), W >> (i + 1), w)
            E.conv_bn_act(ctx, [h], c2, b2, out=E.Act(cats[i], w, w), pooled=pooled)
            cur = pooled
        (c1, b1), (c2, b2) = self.bottle_neck.units()
        cur = E.conv_bn_act(ctx, [E.conv_bn_act(ctx, [cur], c1, b1)], c2, b2)
        for j in range(4):
            lvl = 3 - j
            up = getattr(self, f"up_convolution_{j + 1}")
            E.conv_transpose2x2(ctx, cur, up.up, E.Act(cats[lvl], 0, widths[lvl]))
            (c1, b1), (c2, b2) = up.conv.units()
            h = E.conv_bn_act(ctx, [E.Act(cats[lvl])], c1, b1)
            if lvl > 0:
                cur = E.conv_bn_act(ctx, [h], c2, b2)
        E.conv_bn_act_head(ctx, [h], c2, b2, self.out.conv)     # last half + OutConv: the activation is never stored
        return ctx.outputs
