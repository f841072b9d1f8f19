"""`attention_unet`: ConvBlock encoder, nearest-upsample UpConvBlock decoder with additive attention gates.

Parameters and wiring follow the reference (unet_zoo/models/attention_unet.py:42-110); like the reference class the
``depth`` argument is accepted and ignored (always five levels).  Concat order is [gated skip | upsampled]
(attention_unet.py:90).
"""
import torch.nn as nn

from .. import engine as E
from ._runtime import HotPathModel
from .blocks import AttentionBlock, ConvBlock, UpConvBlock


class AttentionUNet(HotPathModel):
    def __init__(self, in_channels=3, num_classes=1, depth=5):
        super().__init__()
        self._in_channels = in_channels
        self.maxpool = nn.MaxPool2d(kernel_size=2, stride=2)
        widths = (64, 128, 256, 512, 1024)
        cin = in_channels
        for i, w in enumerate(widths, start=1):
            setattr(self, f"conv{i}", ConvBlock(ch_in=cin, ch_out=w))
            cin = w
        for lvl in (5, 4, 3, 2):
            w = widths[lvl - 2]
            setattr(self, f"up{lvl}", UpConvBlock(ch_in=2 * w, ch_out=w))
            setattr(self, f"att{lvl}", AttentionBlock(f_g=w, f_l=w, f_int=w // 2))
            setattr(self, f"upconv{lvl}", ConvBlock(ch_in=2 * w, ch_out=w))
        self.conv_1x1 = nn.Conv2d(64, num_classes, kernel_size=1, stride=1, padding=0)

    def _graph(self, ctx: E.Ctx, x):
        from ._graphs import attention_unet_graph
        return attention_unet_graph(self, ctx, x)
