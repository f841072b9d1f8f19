"""Zero-padded stand-ins ("shadows") for the parameters and BatchNorm buffers of models whose channel counts are not
multiples of 8 (MultiResUnet: 17, 26, 35, 51, 53, 105 ... channels, multiresunet.py:47-52).

The kernels work on NHWC bf16 tensors whose channel slices are multiples of 8 (16-byte vectors, TMA boxes).  Such a model
keeps its parameters and buffers in the reference's shapes (state_dict, optimizer and checkpoints see nothing else),
while the executor is handed objects that quack like ``nn.Conv2d`` / ``nn.ConvTranspose2d`` / ``nn.BatchNorm2d`` and
hold the same numbers scattered into padded tensors: every real tensor of ``C`` channels lives in ``ceil8(C)`` channels
whose tail is zero, concatenations keep one padded segment per source.  A zero-padded convolution writes exact zeros
into its padding channels (zero weight rows, zero bias), BatchNorm2d(affine=False) maps an all-zero channel to zeros,
ReLU / add / max-pool keep them - so the padding never mixes into real channels, forward or backward.

Per step this costs two gathers/scatters that are pure data movement on index tensors built once:
``refresh`` (parameters -> padded fp32 operands, running statistics -> padded buffers; before the forward),
``write_back_stats`` (padded running statistics -> the module buffers; after a training forward) and
``end_backward`` (padded gradients -> the flat gradient buffer the parameters' ``.grad`` are views of).
"""
from typing import Dict, List, Sequence, Tuple

import torch


def ceil8(c: int) -> int:
    return (c + 7) // 8 * 8


class Lay:
    """Channel layout of a (virtual) concatenation: one padded segment per real tensor."""

    def __init__(self, *widths: int):
        self.widths = tuple(int(w) for w in widths)
        self.real = sum(self.widths)
        self.padded = sum(ceil8(w) for w in self.widths)

    def index(self) -> torch.Tensor:
        """Padded position of every real channel."""
        pos, off = [], 0
        for w in self.widths:
            pos.append(torch.arange(w, dtype=torch.int64) + off)
            off += ceil8(w)
        return torch.cat(pos) if pos else torch.zeros(0, dtype=torch.int64)

    def offsets(self) -> List[Tuple[int, int]]:
        """(padded offset, padded width) of each segment."""
        out, off = [], 0
        for w in self.widths:
            out.append((off, ceil8(w)))
            off += ceil8(w)
        return out


class ShadowConv:
    """What engine.conv_bn_act / conv_transpose2x2 read from a convolution module, on padded tensors."""
    __slots__ = ("weight", "bias", "in_channels", "out_channels", "kernel_size")


class ShadowBN:
    """What engine.bn_finalize / _fold_bn / _running_affine read from an nn.BatchNorm2d(affine=False)."""
    __slots__ = ("weight", "bias", "eps", "momentum", "track_running_stats", "running_mean", "running_var",
                 "num_batches_tracked")


def _align4(n: int) -> int:
    return (n + 3) // 4 * 4


class Shadow:
    """All shadows of one model on one device.

    specs: ("conv", module, in_lay, out_lay) | ("convt", module, in_lay, out_lay) | ("bn", module, lay)
    """

    def __init__(self, model: torch.nn.Module, specs: Sequence[tuple], device):
        self.device = device
        self.params = list(model.parameters())
        self.of: Dict[int, object] = {}
        where: Dict[int, torch.Tensor] = {}      # id(real parameter) -> positions in wflat (cpu int64)
        views = []                               # (shadow object, attribute, offset, shape)
        total = 0
        bn_idx, self.bn_real, bn_total = [], [], 0
        bn_views = []
        self._bn_mods = []
        for spec in specs:
            kind, mod = spec[0], spec[1]
            if kind in ("conv", "convt"):
                in_lay, out_lay = spec[2], spec[3]
                w = mod.weight
                T = w.shape[2] * w.shape[3]
                im, om = in_lay.index(), out_lay.index()
                t = torch.arange(T, dtype=torch.int64)
                if kind == "conv":      # [Co][Ci][kh][kw]
                    if (w.shape[0], w.shape[1]) != (out_lay.real, in_lay.real):
                        raise ValueError(f"shadow: conv weight {tuple(w.shape)} vs layouts {out_lay.real} <- {in_lay.real}")
                    shape = (out_lay.padded, in_lay.padded, w.shape[2], w.shape[3])
                    idx = (om[:, None, None] * in_lay.padded + im[None, :, None]) * T + t[None, None, :]
                else:                   # ConvTranspose2d: [Ci][Co][kh][kw]
                    if (w.shape[0], w.shape[1]) != (in_lay.real, out_lay.real):
                        raise ValueError(f"shadow: convT weight {tuple(w.shape)} vs layouts {in_lay.real} -> {out_lay.real}")
                    shape = (in_lay.padded, out_lay.padded, w.shape[2], w.shape[3])
                    idx = (im[:, None, None] * out_lay.padded + om[None, :, None]) * T + t[None, None, :]
                s = ShadowConv()
                s.in_channels, s.out_channels, s.kernel_size = in_lay.padded, out_lay.padded, tuple(w.shape[2:])
                s.weight = s.bias = None
                where[id(w)] = idx.reshape(-1) + total
                views.append((s, "weight", total, shape))
                total += _align4(shape[0] * shape[1] * T)
                if mod.bias is not None:
                    where[id(mod.bias)] = om + total
                    views.append((s, "bias", total, (out_lay.padded,)))
                    total += _align4(out_lay.padded)
                self.of[id(mod)] = s
            elif kind == "bn":
                lay = spec[2]
                if mod.affine or not mod.track_running_stats:
                    raise NotImplementedError("shadow: only BatchNorm2d(affine=False, track_running_stats=True)")
                s = ShadowBN()
                s.weight = s.bias = None
                s.eps, s.momentum, s.track_running_stats = mod.eps, mod.momentum, True
                pos = lay.index()
                for name in ("running_mean", "running_var"):
                    bn_idx.append(pos + bn_total)
                    self.bn_real.append((mod, name))
                    bn_views.append((s, name, bn_total, lay.padded))
                    bn_total += _align4(lay.padded)
                s.num_batches_tracked = None
                self.of[id(mod)] = s
                self._bn_mods.append((mod, s))
            else:
                raise ValueError(kind)
        missing = [i for i, p in enumerate(self.params) if id(p) not in where]
        if missing:
            raise RuntimeError(f"shadow: {len(missing)} parameters have no padded stand-in")
        self.wflat = torch.zeros(max(total, 1), dtype=torch.float32, device=device)
        self.gflat = None
        self.scatter_idx = torch.cat([where[id(p)] for p in self.params]).to(device)
        self.tensors = []                        # (shadow tensor, offset): gradient destinations
        with torch.no_grad():
            for (s, attr, off, shape) in views:
                n = 1
                for d in shape:
                    n *= d
                v = self.wflat[off:off + n].view(shape)
                v.requires_grad_(True)           # the executor computes gradients for operands that ask for one
                setattr(s, attr, v)
                self.tensors.append((v, off))
        # gradient gather: position in the executor's flat gradient buffer (engine.ParamGrads: registration order,
        # every tensor aligned to 4 elements) -> position in the padded gradient buffer; alignment gaps read element 0
        g, off = [], 0
        for p in self.params:
            n = p.numel()
            g.append(where[id(p)])
            if _align4(n) != n:
                g.append(torch.zeros(_align4(n) - n, dtype=torch.int64))
            off += _align4(n)
        self.gather_idx = torch.cat(g).to(device)
        # running statistics: padded mean 0 / variance 1 (an all-zero channel stays zero in eval mode, too)
        self.bflat = torch.zeros(max(bn_total, 1), dtype=torch.float32, device=device)
        for (s, name, off, n) in bn_views:
            v = self.bflat[off:off + n]
            if name == "running_var":
                v.fill_(1.0)
            setattr(s, name, v)
        self.bn_idx = torch.cat(bn_idx).to(device) if bn_idx else None
        self.bn_sizes = [getattr(m, n).numel() for (m, n) in self.bn_real]
        self._sig = None

    # ------------------------------------------------------------------ forward side
    def _signature(self):
        sig = [(p.data_ptr(), p._version) for p in self.params]
        sig += [(getattr(m, n).data_ptr(), getattr(m, n)._version) for (m, n) in self.bn_real]
        return sig

    def refresh(self, training: bool):
        """Parameters and running statistics -> padded operands.  In eval mode skipped while nothing changed (the
        executor's inference memo keys on the version counters of the padded tensors)."""
        sig = None
        if not training:
            sig = self._signature()
            if sig == self._sig:
                return
        with torch.no_grad():
            self.wflat.index_copy_(0, self.scatter_idx, torch.cat([p.detach().reshape(-1) for p in self.params]))
            if self.bn_idx is not None:
                bufs = [getattr(m, n).detach().reshape(-1).float() for (m, n) in self.bn_real]
                self.bflat.index_copy_(0, self.bn_idx, torch.cat(bufs))
            for (mod, s) in self._bn_mods:
                s.num_batches_tracked = mod.num_batches_tracked
        self._sig = sig

    def write_back_stats(self):
        """Padded running statistics (updated by uz_bn_finalize during a training forward) -> the module buffers."""
        if self.bn_idx is None:
            return
        with torch.no_grad():
            vals = self.bflat.index_select(0, self.bn_idx)
            torch._foreach_copy_([getattr(m, n) for (m, n) in self.bn_real], list(vals.split(self.bn_sizes)))

    def pack_sources(self) -> List[torch.Tensor]:
        return [t for (t, _off) in self.tensors if t.dim() == 4]

    # ------------------------------------------------------------------ backward side
    def begin_backward(self, pgrads):
        self.gflat = torch.zeros_like(self.wflat)
        for (t, off) in self.tensors:
            pgrads.add_extern(t, off, self.gflat)

    def end_backward(self, pgrads, params):
        if [id(p) for p in params] != [id(p) for p in self.params]:
            raise RuntimeError("unet_zoo_b200: the model's parameter list changed since its padded operands were built")
        torch.index_select(self.gflat, 0, self.gather_idx, out=pgrads.flat)
        self.gflat = None
        for p in self.params:
            if p.requires_grad:
                pgrads.ready(p)
