"""`resunet`: pre-activation residual blocks, stride-2 conv downsampling, transposed-conv decoder.

Parameters and wiring follow the reference (unet_zoo/models/resunet.py:10-79): [up | skip] concat order, 1x1 head
without activation, and the (printed) warning for num_classes > 1 while still honouring num_classes.
"""
import torch.nn as nn

from .. import engine as E
from ._runtime import HotPathModel
from .blocks import ResidualConv, UpsampleResUnet


class ResUnet(HotPathModel):
    _spatial_multiple = 8

    def __init__(self, in_channels: int = 3, num_classes: int = 1, filters: list = None):
        super().__init__()
        self._in_channels = in_channels
        f = [64, 128, 256, 512] if filters is None else list(filters)
        if num_classes > 1:
            print(f"Warning: ResUnet output layer is set for 1 class by default. For {num_classes} classes, "
                  "consider changing the final Conv2d output channel.")
            self.final_conv_out_channels = num_classes
        else:
            self.final_conv_out_channels = 1
        self.input_layer = nn.Sequential(nn.Conv2d(in_channels, f[0], kernel_size=3, padding=1),
                                         nn.BatchNorm2d(f[0]), nn.ReLU(),
                                         nn.Conv2d(f[0], f[0], kernel_size=3, padding=1))
        self.input_skip = nn.Sequential(nn.Conv2d(in_channels, f[0], kernel_size=3, padding=1))
        self.residual_conv_1 = ResidualConv(f[0], f[1], 2, 1)
        self.residual_conv_2 = ResidualConv(f[1], f[2], 2, 1)
        self.bridge = ResidualConv(f[2], f[3], 2, 1)
        self.upsample_1 = UpsampleResUnet(f[3], f[2], 2, 2)
        self.up_residual_conv1 = ResidualConv(f[2] + f[2], f[2], 1, 1)
        self.upsample_2 = UpsampleResUnet(f[2], f[1], 2, 2)
        self.up_residual_conv2 = ResidualConv(f[1] + f[1], f[1], 1, 1)
        self.upsample_3 = UpsampleResUnet(f[1], f[0], 2, 2)
        self.up_residual_conv3 = ResidualConv(f[0] + f[0], f[0], 1, 1)
        self.output_layer = nn.Sequential(nn.Conv2d(f[0], self.final_conv_out_channels, 1, 1))

    def _graph(self, ctx: E.Ctx, x):
        from ._graphs import resunet_graph
        return resunet_graph(self, ctx, x)
