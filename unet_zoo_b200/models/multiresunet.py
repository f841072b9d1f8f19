"""`multiresunet`: MultiResUnet on the fused sm_100a operators - SURVEY.md 8(f) rank 4.

Parameters, attribute paths (``multiresblock1.conv2d_bn_3x3.conv1.weight`` ...), construction order (default
initialisation consumes the RNG exactly like the reference) and the returned tensor follow
unet_zoo/models/multiresunet.py:7-240.  What is particular about this model:

* channel counts come from ``int(int(f * 1.67) * 0.167 / 0.333 / 0.5)`` (:47-52): 8 + 17 + 26 = 51, 17 + 35 + 53 = 105 ...
  Internally every tensor is zero-padded to a multiple of 8 channels per concatenated source (51 -> [8 | 24 | 32]);
  parameters, buffers and gradients keep the reference's shapes and are scattered / gathered once per step
  (`models/_padded.py`);
* every BatchNorm2d is ``affine=False`` and ``batch_norm1`` of a Multiresblock is applied twice per forward (:79,84):
  in train mode each application uses its own batch statistics and updates the same running buffers, in that order;
* the model's output IS a BatchNorm: ``conv_final`` = Conv2d(1x1) -> BatchNorm2d(affine=False) without activation (:200,238).

A Multiresblock runs (training) as: three conv-BN-ReLU written side by side into one buffer (the concatenation), the
1x1 shortcut convolution with its batch statistics only, and ONE tail operator (`engine.dual_bn_add_relu_bn`):
relu(BN(cat) + BN(shortcut)) from the two raw tensors in one pass that also leaves the statistics of the second
`batch_norm1`, whose apply pass emits the stage's MaxPool2d.  A Respath step: two convolutions with statistics, the same
tail operator (with the inner ReLU of the 3x3 branch).  Eval mode (and UZ_MULTIRES_UNFUSED=1): conv-BN units with the
BatchNorm folded into the weights, `engine.residual_add(relu=True)`, `engine.bn_act`.  Inputs: multiples of 16 (the
reference's decoder concatenates a 2 * floor(s / 2) tensor with an s tensor, so it fails on other sizes too); bf16 mode only.
"""
import os

import torch
import torch.nn as nn

from .. import _lib as L
from .. import engine as E
from ._padded import Lay, Shadow, ceil8
from ._runtime import HotPathModel
from .blocks import _Container


# UZ_MULTIRES_UNFUSED=1: the training graph with every normalised branch materialised (conv-BN[-ReLU], add+ReLU as
# separate passes) - the A/B arm of the measurements in DESIGN.md; eval mode always runs that way (BatchNorm folded).
_UNFUSED = os.environ.get("UZ_MULTIRES_UNFUSED") == "1"


class Conv2d_batchnorm(_Container):                       # multiresunet.py:7-30
    def __init__(self, input_features, num_of_filters, kernel_size=(2, 2), stride=(1, 1), activation="relu", padding=0):
        super().__init__()
        self.activation = activation
        self.conv1 = nn.Conv2d(in_channels=input_features, out_channels=num_of_filters, kernel_size=kernel_size,
                               stride=stride, padding=padding)
        self.batchnorm = nn.BatchNorm2d(num_of_filters, affine=False)


def _widths(f: int, alpha: float = 1.67):
    """multiresunet.py:45-52."""
    w = int(f * alpha)
    return int(w * 0.167), int(w * 0.333), int(w * 0.5)


class Multiresblock(_Container):                          # multiresunet.py:32-85
    def __init__(self, input_features, corresponding_unet_filters, alpha=1.67):
        super().__init__()
        self.corresponding_unet_filters = corresponding_unet_filters
        self.alpha = alpha
        self.W = int(corresponding_unet_filters * alpha)
        w3, w5, w7 = _widths(corresponding_unet_filters, alpha)
        total = w3 + w5 + w7
        self.conv2d_bn_1x1 = Conv2d_batchnorm(input_features, total, kernel_size=(1, 1), activation="None", padding=0)
        self.conv2d_bn_3x3 = Conv2d_batchnorm(input_features, w3, kernel_size=(3, 3), activation="relu", padding=1)
        self.conv2d_bn_5x5 = Conv2d_batchnorm(w3, w5, kernel_size=(3, 3), activation="relu", padding=1)
        self.conv2d_bn_7x7 = Conv2d_batchnorm(w5, w7, kernel_size=(3, 3), activation="relu", padding=1)
        self.batch_norm1 = nn.BatchNorm2d(total, affine=False)


class Respath(_Container):                                # multiresunet.py:87-135
    def __init__(self, input_features, filters, respath_length):
        super().__init__()
        self.filters = filters
        self.respath_length = respath_length
        self.conv2d_bn_1x1_initial = Conv2d_batchnorm(input_features, filters, kernel_size=(1, 1), activation="None", padding=0)
        self.conv2d_bn_3x3_initial = Conv2d_batchnorm(input_features, filters, kernel_size=(3, 3), activation="relu", padding=1)
        self.batch_norm_initial = nn.BatchNorm2d(filters, affine=False)
        self.blocks = nn.ModuleList()
        for _ in range(respath_length - 1):
            self.blocks.append(nn.Sequential(
                Conv2d_batchnorm(filters, filters, kernel_size=(1, 1), activation="None", padding=0),
                Conv2d_batchnorm(filters, filters, kernel_size=(3, 3), activation="relu", padding=1),
                nn.BatchNorm2d(filters, affine=False)))

    def steps(self):
        out = [(self.conv2d_bn_1x1_initial, self.conv2d_bn_3x3_initial, self.batch_norm_initial)]
        return out + [(b[0], b[1], b[2]) for b in self.blocks]


class MultiResUnet(HotPathModel):                         # multiresunet.py:137-240
    _spatial_multiple = 16

    def __init__(self, in_channels, filters=32, num_classes=1, **kwargs):
        super().__init__()
        self.alpha = 1.67
        self.filters = filters
        self.nclasses = num_classes
        self.in_channels = in_channels
        self._in_channels = in_channels
        if num_classes > 8:
            raise NotImplementedError("MultiResUnet: more than 8 classes (the normalised 1x1 output runs in one 8-channel group)")
        f = filters
        out_of = lambda u: sum(_widths(u, self.alpha))            # noqa: E731  (get_mrb_output_channels, :154-159)
        self.multiresblock1 = Multiresblock(in_channels, f)
        self.in_filters1 = out_of(f)
        self.pool1 = nn.MaxPool2d(2, stride=2)
        self.respath1 = Respath(self.in_filters1, f, respath_length=4)
        self.multiresblock2 = Multiresblock(self.in_filters1, f * 2)
        self.in_filters2 = out_of(f * 2)
        self.pool2 = nn.MaxPool2d(2, 2)
        self.respath2 = Respath(self.in_filters2, f * 2, respath_length=3)
        self.multiresblock3 = Multiresblock(self.in_filters2, f * 4)
        self.in_filters3 = out_of(f * 4)
        self.pool3 = nn.MaxPool2d(2, 2)
        self.respath3 = Respath(self.in_filters3, f * 4, respath_length=2)
        self.multiresblock4 = Multiresblock(self.in_filters3, f * 8)
        self.in_filters4 = out_of(f * 8)
        self.pool4 = nn.MaxPool2d(2, 2)
        self.respath4 = Respath(self.in_filters4, f * 8, respath_length=1)
        self.multiresblock5 = Multiresblock(self.in_filters4, f * 16)
        self.in_filters5 = out_of(f * 16)
        prev = self.in_filters5
        for i, mult in zip(range(6, 10), (8, 4, 2, 1)):
            setattr(self, f"upsample{i}", nn.ConvTranspose2d(prev, f * mult, kernel_size=(2, 2), stride=(2, 2), padding=0))
            setattr(self, f"concat_filters{i - 5}", f * mult * 2)
            setattr(self, f"multiresblock{i}", Multiresblock(f * mult * 2, f * mult))
            prev = out_of(f * mult)
            setattr(self, f"in_filters{i}", prev)
        self.conv_final = Conv2d_batchnorm(self.in_filters9, num_classes, kernel_size=(1, 1), activation="None")

    def set_precision(self, precision: str):
        if precision != "bf16":
            raise ValueError("MultiResUnet: only the bf16 mode exists (its channel counts run zero-padded on the "
                             "tensor-core kernels; there is no fp32 parity path)")
        return super().set_precision(precision)

    # ------------------------------------------------------------------ padded stand-ins of parameters and buffers
    def _unit_specs(self, m: Conv2d_batchnorm, in_lay: Lay, out_lay: Lay):
        return [("conv", m.conv1, in_lay, out_lay), ("bn", m.batchnorm, out_lay)]

    def _block_specs(self, blk: Multiresblock, in_lay: Lay):
        w3, w5, w7 = _widths(blk.corresponding_unet_filters, blk.alpha)
        out = Lay(w3, w5, w7)
        specs = self._unit_specs(blk.conv2d_bn_1x1, in_lay, out)
        specs += self._unit_specs(blk.conv2d_bn_3x3, in_lay, Lay(w3))
        specs += self._unit_specs(blk.conv2d_bn_5x5, Lay(w3), Lay(w5))
        specs += self._unit_specs(blk.conv2d_bn_7x7, Lay(w5), Lay(w7))
        specs.append(("bn", blk.batch_norm1, out))
        return specs, out

    def _respath_specs(self, rp: Respath, in_lay: Lay):
        specs, lay = [], in_lay
        for (c1, c3, bn) in rp.steps():
            specs += self._unit_specs(c1, lay, Lay(rp.filters)) + self._unit_specs(c3, lay, Lay(rp.filters))
            specs.append(("bn", bn, Lay(rp.filters)))
            lay = Lay(rp.filters)
        return specs

    def _specs(self):
        f = self.filters
        specs, lay = [], Lay(self.in_channels)
        for i in range(1, 6):
            s, lay = self._block_specs(getattr(self, f"multiresblock{i}"), lay)
            specs += s
            if i < 5:
                specs += self._respath_specs(getattr(self, f"respath{i}"), lay)
        for i, mult in zip(range(6, 10), (8, 4, 2, 1)):
            specs.append(("convt", getattr(self, f"upsample{i}"), lay, Lay(f * mult)))
            s, lay = self._block_specs(getattr(self, f"multiresblock{i}"), Lay(f * mult, f * mult))
            specs += s
        specs += self._unit_specs(self.conv_final, lay, Lay(self.nclasses))
        return specs

    def _before_forward(self, device):
        sh = self._uz_shadow
        if sh is None or sh.device != device or [id(p) for p in sh.params] != [id(p) for p in self.parameters()]:
            sh = self._uz_shadow = Shadow(self, self._specs(), device)
            self._pack_plan = None
            self._eval_cache = None
        sh.refresh(self.training)

    # ------------------------------------------------------------------ graph
    def _unit(self, ctx, xs, m: Conv2d_batchnorm, out=None, need_dx=True):
        sh = self._uz_shadow.of
        mode = L.CONV3X3 if m.conv1.kernel_size[0] == 3 else L.CONV1X1
        return E.conv_bn_act(ctx, xs, sh[id(m.conv1)], sh[id(m.batchnorm)], out=out, relu=m.activation == "relu",
                             mode=mode, need_dx=need_dx)

    def _branch(self, ctx, xs, m: Conv2d_batchnorm, need_dx=True):
        """Convolution + batch statistics of a unit whose normalised output only feeds an add (engine.dual_bn_add_relu)."""
        sh = self._uz_shadow.of
        mode = L.CONV3X3 if m.conv1.kernel_size[0] == 3 else L.CONV1X1
        return E.bn_branch_conv(ctx, xs, sh[id(m.conv1)], sh[id(m.batchnorm)], relu=m.activation == "relu", mode=mode,
                                need_dx=need_dx)

    def _block(self, ctx, blk: Multiresblock, xs, out=None, pooled=None, need_dx=True):
        """Multiresblock.forward (multiresunet.py:70-85)."""
        a0 = xs[0]
        N, H, W = a0.N, a0.H, a0.W
        p3, p5, p7 = (ceil8(w) for w in _widths(blk.corresponding_unet_filters, blk.alpha))
        bn1 = self._uz_shadow.of[id(blk.batch_norm1)]
        cat = ctx.new_buf(N, H, W, p3 + p5 + p7)                   # torch.cat([a, b, c]) with one padded segment each
        fused = ctx.training and not _UNFUSED
        if fused:     # the shortcut's normalised tensor is never stored: relu(BN(cat) + BN(conv1x1(x))) in one pass
            temp = self._branch(ctx, xs, blk.conv2d_bn_1x1, need_dx=need_dx)
        else:
            temp = self._unit(ctx, xs, blk.conv2d_bn_1x1, need_dx=need_dx)
        a = self._unit(ctx, xs, blk.conv2d_bn_3x3, out=E.Act(cat, 0, p3), need_dx=need_dx)
        b = self._unit(ctx, [a], blk.conv2d_bn_5x5, out=E.Act(cat, p3, p5))
        self._unit(ctx, [b], blk.conv2d_bn_7x7, out=E.Act(cat, p3 + p5, p7))
        if fused:     # the whole tail: BN(cat), + shortcut, ReLU, BN again (+ the stage's pooling)
            return E.dual_bn_add_relu_bn(ctx, E.bn_branch_act(ctx, E.Act(cat), bn1), temp, bn1, out=out, pooled=pooled)
        if not ctx.training and os.environ.get("UZ_MULTIRES_EVAL_UNFUSED") != "1":
            s = E.bn_add_relu_eval(ctx, E.Act(cat), bn1, temp)       # running statistics: one pass
        else:
            x1 = E.bn_act(ctx, E.Act(cat), bn1, relu=False)
            s = E.residual_add(ctx, x1, temp, relu=True)
        return E.bn_act(ctx, s, bn1, relu=False, out=out, pooled=pooled)

    def _respath(self, ctx, rp: Respath, x, out):
        """Respath.forward (multiresunet.py:121-135); the last BatchNorm writes into the decoder's concat buffer."""
        steps = rp.steps()
        for i, (c1, c3, bn) in enumerate(steps):
            dst = out if i == len(steps) - 1 else None
            if ctx.training and not _UNFUSED:
                x = E.dual_bn_add_relu_bn(ctx, self._branch(ctx, [x], c3), self._branch(ctx, [x], c1),
                                          self._uz_shadow.of[id(bn)], out=dst)
            else:
                shortcut = self._unit(ctx, [x], c1)
                y = self._unit(ctx, [x], c3)
                s = E.residual_add(ctx, y, shortcut, relu=True)
                x = E.bn_act(ctx, s, self._uz_shadow.of[id(bn)], relu=False, out=dst)
        return x

    def _graph(self, ctx: E.Ctx, x):
        if ctx.fp32:
            raise RuntimeError("MultiResUnet: only the bf16 mode exists")
        if x.dtype == torch.uint8:
            raise RuntimeError("MultiResUnet: uint8 input is served by the im2col stem of the other models only")
        N, cin, H, W = x.shape
        f = self.filters
        cur = ctx.new_act(N, H, W, ceil8(cin))
        L.check(ctx.lib.uz_nchw_to_nhwc(x.data_ptr(), N, cin, H, W, cur.ptr, cur.ld, ctx.stream), "uz_nchw_to_nhwc")
        ctx.launches += 1
        # decoder level i (1 .. 4) at H >> (i - 1): [ConvTranspose2d output | Respath output] (multiresunet.py:222-232)
        cats = [ctx.new_buf(N, H >> i, W >> i, 2 * ceil8(f << i)) for i in range(4)]
        for i in range(1, 5):
            blk = getattr(self, f"multiresblock{i}")
            pooled = ctx.new_act(N, H >> i, W >> i, sum(ceil8(w) for w in _widths(blk.corresponding_unet_filters, blk.alpha)))
            m = self._block(ctx, blk, [cur], pooled=pooled, need_dx=i > 1)
            half = ceil8(f << (i - 1))
            self._respath(ctx, getattr(self, f"respath{i}"), m, E.Act(cats[i - 1], half, half))
            cur = pooled
        cur = self._block(ctx, self.multiresblock5, [cur])
        for i in range(6, 10):
            lvl = 9 - i                                            # 3, 2, 1, 0
            half = ceil8(f << lvl)
            up = self._uz_shadow.of[id(getattr(self, f"upsample{i}"))]
            E.conv_transpose2x2(ctx, cur, up, E.Act(cats[lvl], 0, half))
            cur = self._block(ctx, getattr(self, f"multiresblock{i}"), [E.Act(cats[lvl])])
        y = self._unit(ctx, [cur], self.conv_final)
        E.nhwc_logits(ctx, y, self.nclasses)
        if ctx.training:
            self._uz_shadow.write_back_stats()
        return ctx.outputs
