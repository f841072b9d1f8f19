"""Host logic of the device-side Resize (unet_zoo_b200/data.py) against Pillow itself - the third-party library whose
`Image.resize(size, BILINEAR)` the reference's dataset calls through `transforms.Resize((512, 512))`
(unet_zoo/data/datasets.py:41,47; the reference pins pillow==11.0.0, this image has a newer one with the same
resampler).  The window / coefficient tables of `resample_coeffs`, run through a plain numpy restatement of Pillow's two
fixed-point passes (libImaging/Resample.c: ImagingResampleHorizontal_8bpc / Vertical_8bpc), must reproduce PIL's bytes
exactly; tests/test_resize_gpu.py then holds the CUDA kernels to the same bytes."""
import numpy as np
import pytest

from unet_zoo_b200.data import PRECISION_BITS, resample_coeffs

Image = pytest.importorskip("PIL.Image")


def resize_numpy(a, h_out, w_out):
    """The two passes of Pillow's resampler in numpy int64 (horizontal first, uint8 intermediate)."""
    x = a.astype(np.int64)
    half = 1 << (PRECISION_BITS - 1)
    if w_out != x.shape[1]:
        bounds, kk, _ = resample_coeffs(x.shape[1], w_out)
        o = np.empty((x.shape[0], w_out, x.shape[2]), np.int64)
        for i, (x0, n) in enumerate(bounds):
            o[:, i] = half + (x[:, x0:x0 + n] * np.array(kk[i][:n], np.int64)[None, :, None]).sum(1)
        x = np.clip(o >> PRECISION_BITS, 0, 255)
    if h_out != x.shape[0]:
        bounds, kk, _ = resample_coeffs(x.shape[0], h_out)
        o = np.empty((h_out, x.shape[1], x.shape[2]), np.int64)
        for i, (y0, n) in enumerate(bounds):
            o[i] = half + (x[y0:y0 + n] * np.array(kk[i][:n], np.int64)[:, None, None]).sum(0)
        x = np.clip(o >> PRECISION_BITS, 0, 255)
    return x.astype(np.uint8)


def pil_resize(a, h_out, w_out):
    im = Image.fromarray(a if a.shape[2] == 3 else a[..., 0], "RGB" if a.shape[2] == 3 else "L")
    return np.asarray(im.resize((w_out, h_out), Image.BILINEAR)).reshape(h_out, w_out, a.shape[2])


CASES = [(300, 400, 3, 512, 512), (1000, 777, 3, 512, 512), (512, 512, 1, 512, 512), (600, 512, 1, 512, 512),
         (97, 1301, 3, 64, 48), (2048, 1536, 3, 512, 512), (5, 7, 3, 16, 16), (1, 1, 1, 4, 3)]


@pytest.mark.parametrize("h,w,c,ho,wo", CASES)
def test_coefficient_tables_reproduce_pil_bytes(h, w, c, ho, wo):
    rng = np.random.default_rng(h * 7 + w)
    a = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
    assert np.array_equal(resize_numpy(a, ho, wo), pil_resize(a, ho, wo))


def test_tables_are_normalised_windows():
    bounds, kk, ksize = resample_coeffs(1000, 512)
    assert len(bounds) == len(kk) == 512 and all(len(r) == ksize for r in kk)
    for (x0, n), row in zip(bounds, kk):
        assert 0 <= x0 and x0 + n <= 1000 and 0 < n <= ksize
        assert abs(sum(row) - (1 << PRECISION_BITS)) <= n       # coefficients sum to one (fixed point, rounding)
        assert all(v == 0 for v in row[n:])


def test_cpu_tensor_is_rejected():
    import torch

    from unet_zoo_b200.data import resize_u8
    with pytest.raises(RuntimeError, match="GPU"):
        resize_u8(torch.zeros(8, 8, 3, dtype=torch.uint8), (4, 4))
