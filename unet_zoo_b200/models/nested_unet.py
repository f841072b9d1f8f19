"""`nested_unet` (UNet++): dense grid of VGG blocks with bilinear (align_corners=True) upsampling.

Parameters and wiring follow the reference (unet_zoo/models/nested_unet.py:24-106), including the reference's
argument order (``num_classes`` first), the swallowed ``**kwargs`` (``depth``) and the list of four logits returned
under ``deep_supervision=True``.
"""
import torch.nn as nn

from .. import engine as E
from ._runtime import HotPathModel
from .blocks import VGGBlock


class NestedUNet(HotPathModel):
    def __init__(self, num_classes, in_channels=3, deep_supervision=False, **kwargs):
        super().__init__()
        self._in_channels = in_channels
        nb = [32, 64, 128, 256, 512]
        self.deep_supervision = deep_supervision
        self.pool = nn.MaxPool2d(2, 2)
        self.up = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True)
        for i in range(5):                                   # backbone column (nested_unet.py:35-39)
            setattr(self, f"conv{i}_0", VGGBlock(in_channels if i == 0 else nb[i - 1], nb[i], nb[i]))
        # the reference registers the dense nodes in this (row-interleaved) order: nested_unet.py:41-59
        for (i, j) in ((0, 1), (1, 1), (2, 1), (3, 1), (0, 2), (1, 2), (2, 2), (0, 3), (1, 3), (0, 4)):
            setattr(self, f"conv{i}_{j}", VGGBlock(nb[i] * j + nb[i + 1], nb[i], nb[i]))
        if deep_supervision:
            for k in range(1, 5):
                setattr(self, f"final{k}", nn.Conv2d(nb[0], num_classes, kernel_size=1))
        else:
            self.final = nn.Conv2d(nb[0], num_classes, kernel_size=1)

    def _graph(self, ctx: E.Ctx, x):
        from ._graphs import nested_unet_graph
        return nested_unet_graph(self, ctx, x)

    def _pack_outputs(self, outs):
        return outs if self.deep_supervision else outs[0]
