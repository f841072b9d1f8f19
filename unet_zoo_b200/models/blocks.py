"""Parameter containers of the hot-path blocks.

These modules hold the fp32 parameters / BatchNorm buffers under exactly the attribute paths the reference uses, so
``state_dict()`` keys, shapes and default initialisation order are identical and reference ``.pth`` files load
unchanged (unet_zoo/models/common_layers.py:20-128,182-207; attention_unet.py:6-40; nested_unet.py:4-22).
They never run ``torch.nn`` arithmetic: a model's ``forward`` hands them to the fused operators of
:mod:`unet_zoo_b200.engine`.  Calling a block on its own is therefore an error (there is no eager fallback).
"""
import torch.nn as nn


class _Container(nn.Module):
    def forward(self, *args, **kwargs):
        raise RuntimeError(
            f"unet_zoo_b200: {type(self).__name__} is a parameter container of the fused sm_100a executor; call the "
            "whole model (create_model(...)(x)) - blocks have no stand-alone eager path")


def _cbr(cin, cout, n=2, bias=True):
    """conv3x3-BN-ReLU repeated n times as an nn.Sequential with the reference's index layout (0,1,2,3,4,5)."""
    layers = []
    for i in range(n):
        layers += [nn.Conv2d(cin if i == 0 else cout, cout, kernel_size=3, stride=1, padding=1, bias=bias),
                   nn.BatchNorm2d(cout), nn.ReLU(inplace=True)]
    return nn.Sequential(*layers)


class DoubleConv(_Container):          # common_layers.py:20-37
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.conv_op = _cbr(in_channels, out_channels)

    def units(self):
        s = self.conv_op
        return (s[0], s[1]), (s[3], s[4])


class ConvBlock(_Container):           # common_layers.py:39-61
    def __init__(self, ch_in, ch_out):
        super().__init__()
        self.conv = _cbr(ch_in, ch_out)

    def units(self):
        s = self.conv
        return (s[0], s[1]), (s[3], s[4])


class UpConvBlock(_Container):         # common_layers.py:63-80 (nearest x2 -> conv3x3 -> BN -> ReLU)
    def __init__(self, ch_in, ch_out):
        super().__init__()
        self.up = nn.Sequential(nn.Upsample(scale_factor=2),
                                nn.Conv2d(ch_in, ch_out, kernel_size=3, stride=1, padding=1, bias=True),
                                nn.BatchNorm2d(ch_out), nn.ReLU(inplace=True))


class DownSample(_Container):          # common_layers.py:82-95
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.conv = DoubleConv(in_channels, out_channels)
        self.pool = nn.MaxPool2d(kernel_size=2, stride=2)


class UpSample_UNet(_Container):       # common_layers.py:97-116
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.up = nn.ConvTranspose2d(in_channels, in_channels // 2, kernel_size=2, stride=2)
        self.conv = DoubleConv(in_channels, out_channels)


class OutConv(_Container):             # common_layers.py:118-128
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size=1)


class ResidualConv(_Container):        # common_layers.py:182-199
    def __init__(self, in_channels, out_channels, stride, padding):
        super().__init__()
        self.stride = stride
        self.conv_block = nn.Sequential(
            nn.BatchNorm2d(in_channels), nn.ReLU(),
            nn.Conv2d(in_channels, out_channels, kernel_size=3, stride=stride, padding=padding, bias=False),
            nn.BatchNorm2d(out_channels), nn.ReLU(),
            nn.Conv2d(out_channels, out_channels, kernel_size=3, padding=1, bias=False))
        self.conv_skip = nn.Sequential(
            nn.Conv2d(in_channels, out_channels, kernel_size=1, stride=stride, bias=False),
            nn.BatchNorm2d(out_channels))


class UpsampleResUnet(_Container):     # common_layers.py:201-207
    def __init__(self, in_channels, out_channels, kernel_size, stride):
        super().__init__()
        self.upsample = nn.ConvTranspose2d(in_channels, out_channels, kernel_size=kernel_size, stride=stride)


class VGGBlock(_Container):            # nested_unet.py:4-22
    def __init__(self, in_channels, middle_channels, out_channels):
        super().__init__()
        self.relu = nn.ReLU(inplace=True)
        self.conv1 = nn.Conv2d(in_channels, middle_channels, 3, padding=1)
        self.bn1 = nn.BatchNorm2d(middle_channels)
        self.conv2 = nn.Conv2d(middle_channels, out_channels, 3, padding=1)
        self.bn2 = nn.BatchNorm2d(out_channels)

    def units(self):
        return (self.conv1, self.bn1), (self.conv2, self.bn2)


class AttentionBlock(_Container):      # attention_unet.py:6-40
    def __init__(self, f_g, f_l, f_int):
        super().__init__()
        self.w_g = nn.Sequential(nn.Conv2d(f_g, f_int, kernel_size=1, stride=1, padding=0, bias=True),
                                 nn.BatchNorm2d(f_int))
        self.w_x = nn.Sequential(nn.Conv2d(f_l, f_int, kernel_size=1, stride=1, padding=0, bias=True),
                                 nn.BatchNorm2d(f_int))
        self.psi = nn.Sequential(nn.Conv2d(f_int, 1, kernel_size=1, stride=1, padding=0, bias=True),
                                 nn.BatchNorm2d(1), nn.Sigmoid())
        self.relu = nn.ReLU(inplace=True)
