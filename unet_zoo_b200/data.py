"""Input pipeline in front of the stem: ``transforms.Resize((512, 512))`` of the reference's
dataset (unet_zoo/data/datasets.py:40-49) on the DEVICE, bit-exact with the PIL resize it replaces.

The reference decodes every image with PIL, resizes it on a CPU worker (``transforms.Resize`` on a PIL image is
``Image.resize(size, BILINEAR)``: Pillow's two-pass fixed-point resampler with support scaling, i.e. anti-aliased when
shrinking), converts to float, normalises and ships 12 bytes per pixel.  Here the decoded bytes (any size) go to the
GPU as they are; :func:`resize_u8` runs Pillow's algorithm there (``uz_resize_bilinear_u8``: integer arithmetic, so the
result equals PIL's byte for byte), and the model's stem kernel does ToTensor + Normalize
(``HotPathModel.set_input_normalization``), the mask path ``(m / 255 > 0.5)`` (``training.to_device_batch``).
Decoding (PNG / JPEG) stays on the host.

The window / coefficient tables are built on the host exactly as Pillow builds them (``precompute_coeffs`` and
``normalize_coeffs_8bpc`` in libImaging/Resample.c: double precision, bilinear filter with support 1.0 scaled by the
shrink factor, 22 fractional bits) and cached per (input size, output size).
"""
import math
from functools import lru_cache
from typing import Sequence, Tuple

import torch

from . import _lib as L

PRECISION_BITS = 32 - 8 - 2


@lru_cache(maxsize=256)
def resample_coeffs(in_size: int, out_size: int) -> Tuple[Tuple[Tuple[int, int], ...], Tuple[Tuple[int, ...], ...], int]:
    """Pillow's bilinear resampling windows for one axis: (bounds[out] = (first, count), kk[out][ksize] int, ksize)."""
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale                  # bilinear filter support = 1.0
    ksize = int(math.ceil(support)) * 2 + 1
    ss = 1.0 / filterscale
    bounds, kk = [], []
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = []
        for x in range(xmax):
            v = (x + xmin - center + 0.5) * ss
            v = -v if v < 0.0 else v
            w.append(1.0 - v if v < 1.0 else 0.0)
        ww = 0.0
        for v in w:                               # same left-to-right accumulation as the C loop
            ww += v
        if ww != 0.0:
            w = [v / ww for v in w]
        row = [int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS)) for v in w]
        row += [0] * (ksize - xmax)
        bounds.append((xmin, xmax))
        kk.append(tuple(row))
    return tuple(bounds), tuple(kk), ksize


_TABLES = {}


def _device_tables(in_size: int, out_size: int, device):
    key = (in_size, out_size, str(device))
    t = _TABLES.get(key)
    if t is None:
        bounds, kk, ksize = resample_coeffs(in_size, out_size)
        t = _TABLES[key] = (torch.tensor(bounds, dtype=torch.int32, device=device).contiguous(),
                            torch.tensor(kk, dtype=torch.int32, device=device).contiguous(), ksize)
    return t


def resize_u8(img: torch.Tensor, size: Sequence[int]) -> torch.Tensor:
    """``transforms.Resize(size)`` for decoded bytes on the GPU: uint8 ``[H, W, C]`` or ``[N, H, W, C]`` (C <= 4, as PIL
    lays an "RGB" / "L" image out) -> uint8 of the same rank with spatial size ``size = (H_out, W_out)``.  Bit-exact
    with ``PIL.Image.resize((W_out, H_out), PIL.Image.BILINEAR)``.  CUDA tensors only (there is no CPU path)."""
    if img.dtype != torch.uint8 or img.dim() not in (3, 4):
        raise RuntimeError("unet_zoo_b200.data.resize_u8: expected a uint8 tensor [H, W, C] or [N, H, W, C]")
    if not img.is_cuda:
        raise RuntimeError("unet_zoo_b200.data.resize_u8: the image must be on the GPU (there is no CPU path; the "
                           "reference's CPU path is PIL itself)")
    batched = img.dim() == 4
    x = (img if batched else img[None]).contiguous()
    N, Hin, Win, C = x.shape
    Hout, Wout = int(size[0]), int(size[1])
    if C > 4 or Hout < 1 or Wout < 1:
        raise RuntimeError("unet_zoo_b200.data.resize_u8: at most 4 channels, positive output size")
    lib = L.load()
    dev = x.device
    out = torch.empty((N, Hout, Wout, C), dtype=torch.uint8, device=dev)
    tmp = torch.empty((N, Hin, Wout, C), dtype=torch.uint8, device=dev) if (Hout != Hin and Wout != Win) else None
    bh, kh, ksh = _device_tables(Win, Wout, dev) if Wout != Win else (None, None, 0)
    bv, kv, ksv = _device_tables(Hin, Hout, dev) if Hout != Hin else (None, None, 0)
    with torch.cuda.device(dev):
        rc = lib.uz_resize_bilinear_u8(x.data_ptr(), N, Hin, Win, C, tmp.data_ptr() if tmp is not None else None,
                                       out.data_ptr(), Hout, Wout, bh.data_ptr() if bh is not None else None,
                                       kh.data_ptr() if kh is not None else None, ksh,
                                       bv.data_ptr() if bv is not None else None,
                                       kv.data_ptr() if kv is not None else None, ksv,
                                       torch.cuda.current_stream(dev).cuda_stream)
    L.check(rc, "uz_resize_bilinear_u8")
    return out if batched else out[0]
