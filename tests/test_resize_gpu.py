"""uz_resize_bilinear_u8 against Pillow: byte-exact on images and masks of assorted sizes, batched and single, and the
whole input path  decoded bytes -> device Resize -> uint8 stem (ToTensor + Normalize inside the kernel) -> logits
against the reference's CPU pipeline (PIL resize -> ToTensor -> Normalize -> fp32 tensor -> same model)."""
import numpy as np
import pytest
import torch

from conftest import seeded_model
from test_resize_cpu import CASES, pil_resize
from unet_zoo_b200.data import resize_u8
from unet_zoo_b200.training import IMAGENET_MEAN, IMAGENET_STD, to_device_batch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("h,w,c,ho,wo", CASES)
def test_device_resize_is_byte_exact_with_pil(h, w, c, ho, wo):
    rng = np.random.default_rng(h * 7 + w)
    a = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
    got = resize_u8(torch.from_numpy(a).cuda(), (ho, wo)).cpu().numpy()
    assert got.shape == (ho, wo, c) and np.array_equal(got, pil_resize(a, ho, wo))


def test_batched_resize_and_smooth_images():
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:700, 0:900]
    imgs = []
    for k in range(3):                              # smooth content + noise: exercises the rounding of mid-range sums
        base = 127.5 + 120 * np.sin(xx / (17.0 + k)) * np.cos(yy / (23.0 - k))
        imgs.append(np.clip(base[..., None] + rng.normal(0, 6, (700, 900, 3)), 0, 255).astype(np.uint8))
    batch = torch.from_numpy(np.stack(imgs)).cuda()
    got = resize_u8(batch, (512, 512)).cpu().numpy()
    for k in range(3):
        assert np.array_equal(got[k], pil_resize(imgs[k], 512, 512))


def test_input_path_matches_the_reference_cpu_pipeline():
    """datasets.py:40-61 on the CPU (PIL resize -> ToTensor -> Normalize; mask: resize -> ToTensor -> > 0.5) against the
    device path on the decoded bytes: identical mask, logits bit-identical to feeding the CPU-normalised tensor."""
    rng = np.random.default_rng(9)
    raw = rng.integers(0, 256, (2, 300, 260, 3), dtype=np.uint8)
    raw_mask = rng.integers(0, 256, (2, 300, 260, 1), dtype=np.uint8)
    S = 128
    # reference pipeline (CPU)
    ref_img = np.stack([pil_resize(r, S, S) for r in raw]).astype(np.float32) / 255.0
    mean, std = np.array(IMAGENET_MEAN, np.float32), np.array(IMAGENET_STD, np.float32)
    ref_x = torch.from_numpy(((ref_img - mean) / std).transpose(0, 3, 1, 2).copy())
    ref_m = torch.from_numpy((np.stack([pil_resize(m, S, S) for m in raw_mask]).astype(np.float32) / 255.0 > 0.5)
                             .astype(np.float32).transpose(0, 3, 1, 2).copy())
    # device pipeline
    img8 = resize_u8(torch.from_numpy(raw).cuda(), (S, S))
    m8 = resize_u8(torch.from_numpy(raw_mask).cuda(), (S, S))
    x_dev, m_dev = to_device_batch(img8, m8.permute(0, 3, 1, 2).contiguous(), torch.device("cuda", 0))
    assert torch.equal(m_dev.cpu(), ref_m)
    model = seeded_model("unet").cuda().eval()
    model.set_input_normalization(IMAGENET_MEAN, IMAGENET_STD)
    with torch.no_grad():
        a = model(x_dev)
        model.set_input_normalization(None, None)
        b = model(ref_x.cuda())
    assert torch.equal(a, b)
