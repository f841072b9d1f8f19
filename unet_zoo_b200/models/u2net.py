"""`u2net` / `u2netp`: U^2-Net (nested RSU blocks) on the fused sm_100a operators - SURVEY.md 8(f) rank 4.

Parameters, attribute paths (``stage1.rebnconvin.conv_s1.weight`` ...), construction order (default initialisation
consumes the RNG exactly like the reference) and the returned dict follow unet_zoo/models/u2net.py:6-381.  Every
REBNCONV (Conv2d 3x3 with padding = dilation = d -> BatchNorm2d -> ReLU, :6-17) is one `engine.conv_bn_act`; dilated
ones run on the per-tap kernel (``UZ_DILATION``).  Nothing is concatenated or pooled in a pass of its own:

* inside an RSU block level i owns one NHWC buffer ``[decoder input | encoder skip]`` (the reference concatenates the
  upsampled / deeper tensor FIRST, e.g. :66 ``torch.cat((hx7, hx6), 1)``); encoder convs write the skip half (and the
  pooled tensor of the next level from the same BatchNorm pass), the bilinear upsample (align_corners=False, :19-23)
  writes the other half;
* the block's residual ``hx1d + hxin`` (:78) is one pass that also emits the stage's MaxPool2d (:222) and lands in
  the skip half of the OUTER decoder's buffer;
* the six side convolutions (Conv2d(C, out_ch, 3), :238-243) run with their output channels zero-padded to 8, are
  upsampled into 8-channel slots of ONE 48-channel full-resolution buffer, and a single pass over it (``uz_side_head_fwd``)
  produces all seven fp32 logit maps: ``outconv`` (:244) over the six side maps, and the side maps themselves.

Input sizes: multiples of 32 (five 2x poolings at every nesting level; the reference's ceil-mode pooling and
size-matching upsample also accept other sizes - not implemented, raises).  bf16 mode only.
"""
import torch.nn as nn

from .. import engine as E
from ._runtime import HotPathModel
from .blocks import _Container


class REBNCONV(_Container):                      # u2net.py:6-17
    def __init__(self, in_ch=3, out_ch=3, dirate=1):
        super().__init__()
        self.dirate = dirate
        self.conv_s1 = nn.Conv2d(in_ch, out_ch, 3, padding=1 * dirate, dilation=1 * dirate)
        self.bn_s1 = nn.BatchNorm2d(out_ch)
        self.relu_s1 = nn.ReLU(inplace=True)


class RSU(_Container):
    """RSU-L for L = 7, 6, 5, 4 (u2net.py:26-78, :81-127, :130-171, :174-210): the four classes differ only in depth."""

    def __init__(self, L_, in_ch=3, mid_ch=12, out_ch=3):
        super().__init__()
        self.depth = L_
        self.rebnconvin = REBNCONV(in_ch, out_ch, dirate=1)
        self.rebnconv1 = REBNCONV(out_ch, mid_ch, dirate=1)
        for i in range(2, L_):
            setattr(self, f"pool{i - 1}", nn.MaxPool2d(2, stride=2, ceil_mode=True))
            setattr(self, f"rebnconv{i}", REBNCONV(mid_ch, mid_ch, dirate=1))
        setattr(self, f"rebnconv{L_}", REBNCONV(mid_ch, mid_ch, dirate=2))
        for i in range(L_ - 1, 1, -1):
            setattr(self, f"rebnconv{i}d", REBNCONV(mid_ch * 2, mid_ch, dirate=1))
        self.rebnconv1d = REBNCONV(mid_ch * 2, out_ch, dirate=1)


class RSU4F(_Container):                         # u2net.py:213-237
    def __init__(self, in_ch=3, mid_ch=12, out_ch=3):
        super().__init__()
        self.rebnconvin = REBNCONV(in_ch, out_ch, dirate=1)
        self.rebnconv1 = REBNCONV(out_ch, mid_ch, dirate=1)
        self.rebnconv2 = REBNCONV(mid_ch, mid_ch, dirate=2)
        self.rebnconv3 = REBNCONV(mid_ch, mid_ch, dirate=4)
        self.rebnconv4 = REBNCONV(mid_ch, mid_ch, dirate=8)
        self.rebnconv3d = REBNCONV(mid_ch * 2, mid_ch, dirate=4)
        self.rebnconv2d = REBNCONV(mid_ch * 2, mid_ch, dirate=2)
        self.rebnconv1d = REBNCONV(mid_ch * 2, out_ch, dirate=1)


def _rebn(ctx, xs, m: REBNCONV, out=None, pooled=None, stem=None, need_dx=True):
    return E.conv_bn_act(ctx, xs, m.conv_s1, m.bn_s1, out=out, pooled=pooled, stem=stem, need_dx=need_dx,
                         dilation=m.dirate)


def _rsu(ctx: E.Ctx, blk: RSU, xs, out=None, pooled=None, stem=None, need_dx=True) -> E.Act:
    """RSU-L forward (u2net.py:46-78 for L = 7)."""
    depth = blk.depth
    a0 = xs[0]
    N, H, W = a0.N, a0.H, a0.W
    mid = blk.rebnconv1.conv_s1.out_channels
    hxin = _rebn(ctx, xs, blk.rebnconvin, stem=stem, need_dx=need_dx)
    # level i (1 .. L-1) at H >> (i - 1): [decoder input (mid) | encoder output hx_i (mid)]
    cats = [ctx.new_buf(N, H >> i, W >> i, 2 * mid) for i in range(depth - 1)]
    cur = hxin
    for i in range(1, depth):
        skip = E.Act(cats[i - 1], mid, mid)
        nxt = ctx.new_act(N, H >> i, W >> i, mid) if i < depth - 1 else None      # pool_i, ceil_mode moot: even sizes
        _rebn(ctx, [cur], getattr(blk, f"rebnconv{i}"), out=skip, pooled=nxt)
        cur = nxt if nxt is not None else skip
    _rebn(ctx, [cur], getattr(blk, f"rebnconv{depth}"), out=E.Act(cats[depth - 2], 0, mid))          # dilation 2
    for i in range(depth - 1, 0, -1):
        d = _rebn(ctx, [E.Act(cats[i - 1])], getattr(blk, f"rebnconv{i}d"))
        if i > 1:                                                                 # _upsample_like(hx_id, hx_(i-1))
            E.bilinear_resize(ctx, d, 2 * d.H, 2 * d.W, out=E.Act(cats[i - 2], 0, mid), align_corners=False)
    return E.residual_add(ctx, d, hxin, out=out, pooled=pooled)


def _rsu4f(ctx: E.Ctx, blk: RSU4F, xs, out=None, pooled=None, stem=None, need_dx=True) -> E.Act:
    """RSU-4F forward (u2net.py:226-237)."""
    a0 = xs[0]
    N, H, W = a0.N, a0.H, a0.W
    mid = blk.rebnconv1.conv_s1.out_channels
    hxin = _rebn(ctx, xs, blk.rebnconvin, stem=stem, need_dx=need_dx)
    c1, c2, c3 = (ctx.new_buf(N, H, W, 2 * mid) for _ in range(3))
    hx1 = _rebn(ctx, [hxin], blk.rebnconv1, out=E.Act(c1, mid, mid))
    hx2 = _rebn(ctx, [hx1], blk.rebnconv2, out=E.Act(c2, mid, mid))
    hx3 = _rebn(ctx, [hx2], blk.rebnconv3, out=E.Act(c3, mid, mid))
    _rebn(ctx, [hx3], blk.rebnconv4, out=E.Act(c3, 0, mid))
    _rebn(ctx, [E.Act(c3)], blk.rebnconv3d, out=E.Act(c2, 0, mid))
    _rebn(ctx, [E.Act(c2)], blk.rebnconv2d, out=E.Act(c1, 0, mid))
    hx1d = _rebn(ctx, [E.Act(c1)], blk.rebnconv1d)
    return E.residual_add(ctx, hx1d, hxin, out=out, pooled=pooled)


class _U2NetBase(HotPathModel):
    _spatial_multiple = 32
    _spec = None      # name -> (kind, in, mid, out), in the reference's construction order

    def __init__(self, in_ch=3, out_ch=1):
        super().__init__()
        self._in_channels = in_ch
        self.out_ch = out_ch
        if out_ch > 4:
            raise NotImplementedError("U2NET: more than 4 classes (uz_side_head_fwd handles K <= 4 per side map)")
        for name, (kind, cin, mid, cout) in self._spec:
            cin = in_ch if name == "stage1" else cin
            setattr(self, name, RSU4F(cin, mid, cout) if kind == "4F" else RSU(kind, cin, mid, cout))
            if name.startswith("stage") and not name.endswith("d") and name != "stage6":
                i = int(name[5:])
                setattr(self, f"pool{i}{i + 1}", nn.MaxPool2d(2, stride=2, ceil_mode=True))
        for i, c in enumerate(self._side_channels, start=1):
            setattr(self, f"side{i}", nn.Conv2d(c, out_ch, 3, padding=1))
        self.outconv = nn.Conv2d(6 * out_ch, out_ch, 1)

    def set_precision(self, precision: str):
        if precision != "bf16":
            raise ValueError(f"{type(self).__name__}: only the bf16 mode exists (no fp32 parity path for the dilated "
                             "convolutions)")
        return super().set_precision(precision)

    def _stage(self, ctx, name, xs, **kw):
        blk = getattr(self, name)
        return _rsu4f(ctx, blk, xs, **kw) if isinstance(blk, RSU4F) else _rsu(ctx, blk, xs, **kw)

    def _graph(self, ctx: E.Ctx, x):
        if ctx.fp32:
            raise RuntimeError(f"{type(self).__name__}: only the bf16 mode exists")
        N, _, H, W = x.shape
        K = self.out_ch
        width = [getattr(self, f"stage{i}").rebnconvin.conv_s1.out_channels for i in range(1, 7)]
        # outer decoder level i (1 .. 5): [upsampled deeper output | encoder output hx_i], u2net.py:263-271
        cats = [ctx.new_buf(N, H >> i, W >> i, 2 * width[i]) for i in range(5)]
        cur, stem = E.stem_input(ctx, x)
        for i in range(6):
            out = E.Act(cats[i], width[i], width[i]) if i < 5 else None
            pooled = ctx.new_act(N, H >> (i + 1), W >> (i + 1), width[i]) if i < 5 else None
            out = self._stage(ctx, f"stage{i + 1}", [cur], out=out, pooled=pooled, stem=stem if i == 0 else None,
                              need_dx=i > 0)
            cur = pooled
        dec = {6: out}
        prev = out
        for i in range(5, 0, -1):
            E.bilinear_resize(ctx, prev, 2 * prev.H, 2 * prev.W, out=E.Act(cats[i - 1], 0, width[i - 1]),
                              align_corners=False)
            prev = dec[i] = self._stage(ctx, f"stage{i}d", [E.Act(cats[i - 1])])
        # side outputs (u2net.py:273-284) into 8-channel slots of one full-resolution buffer, then one head pass
        fuse = ctx.new_buf(N, H, W, 48)
        for i in range(1, 7):
            slot = E.Act(fuse, 8 * (i - 1), 8)
            conv = getattr(self, f"side{i}")
            if i == 1:
                E.conv_padded_out(ctx, [dec[1]], conv, slot)
            else:
                s = E.conv_padded_out(ctx, [dec[i]], conv, ctx.new_act(N, dec[i].H, dec[i].W, 8))
                E.bilinear_resize(ctx, s, H, W, out=slot, align_corners=False)
        E.fused_side_head(ctx, E.Act(fuse), self.outconv, K, 6)
        return ctx.outputs

    def _pack_outputs(self, outs):
        o, K = outs[0], self.out_ch
        keys = ("main", "side1", "side2", "side3", "side4", "side5", "side6")       # u2net.py:287-295
        return {k: o[:, i * K:(i + 1) * K] for i, k in enumerate(keys)}


class U2NET(_U2NetBase):                         # u2net.py:214-295
    _spec = (("stage1", (7, 3, 32, 64)), ("stage2", (6, 64, 32, 128)), ("stage3", (5, 128, 64, 256)),
             ("stage4", (4, 256, 128, 512)), ("stage5", ("4F", 512, 256, 512)), ("stage6", ("4F", 512, 256, 512)),
             ("stage5d", ("4F", 1024, 256, 512)), ("stage4d", (4, 1024, 128, 256)), ("stage3d", (5, 512, 64, 128)),
             ("stage2d", (6, 256, 32, 64)), ("stage1d", (7, 128, 16, 64)))
    _side_channels = (64, 64, 128, 256, 512, 512)


class U2NETP(_U2NetBase):                        # u2net.py:298-379
    _spec = (("stage1", (7, 3, 16, 64)), ("stage2", (6, 64, 16, 64)), ("stage3", (5, 64, 16, 64)),
             ("stage4", (4, 64, 16, 64)), ("stage5", ("4F", 64, 16, 64)), ("stage6", ("4F", 64, 16, 64)),
             ("stage5d", ("4F", 128, 16, 64)), ("stage4d", (4, 128, 16, 64)), ("stage3d", (5, 128, 16, 64)),
             ("stage2d", (6, 128, 16, 64)), ("stage1d", (7, 128, 16, 64)))
    _side_channels = (64, 64, 64, 64, 64, 64)
