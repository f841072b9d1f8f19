"""The drop-in epoch loops on the GPU with the hot-path model, the fused criterion and the flat optimizer, against the
reference's loop body run with torch's own loss / clip / AdamW on a second copy of the same model."""
import copy

import pytest
import torch

from conftest import seeded_model
from oracle import reference_port as O
from unet_zoo_b200 import training as T
from unet_zoo_b200.loss import BCEDiceWithLogits
from unet_zoo_b200.optim import FlatAdamW

pytestmark = pytest.mark.gpu


def _batches(n):
    out = []
    for i in range(n):
        g = torch.Generator().manual_seed(100 + i)
        out.append((torch.randn(2, 3, 64, 64, generator=g), (torch.rand(2, 1, 64, 64, generator=g) > 0.5).float(), f"p{i}"))
    return out


def test_train_and_validate_epoch_with_fused_tail_match_the_torch_tail():
    dev = torch.device("cuda", 0)
    data = _batches(3)
    a = seeded_model("unet").to(dev)
    b = seeded_model("unet").to(dev)
    opt_a = FlatAdamW(a.parameters(), lr=1e-4, weight_decay=1e-5, max_norm=1.0)
    opt_b = torch.optim.AdamW(b.parameters(), lr=1e-4, weight_decay=1e-5)
    got = T.train_one_epoch(a, data, opt_a, BCEDiceWithLogits(), dev, None, 0, "unet", None)
    crit = torch.nn.BCEWithLogitsLoss()
    b.train()
    rl = rd = 0.0
    for img, mask, _ in data:                                   # training_loop.py:108-124
        img, mask = img.to(dev), mask.to(dev)
        opt_b.zero_grad()
        out = b(img)
        loss = crit(out, mask)
        dc = O.dice_coefficient(out.detach().cpu(), mask.cpu())
        loss.backward()
        torch.nn.utils.clip_grad_norm_(b.parameters(), 1.0)
        opt_b.step()
        rl += loss.item()
        rd += float(dc)
    # two optimizers with 1e-6-level arithmetic differences: the losses agree to ~1e-5; the Dice of random-init logits
    # (|z| ~ 0.1, straddling the threshold) moves by a few flipped pixels
    assert abs(got[0] - rl / 3) < 5e-5 and abs(got[1] - rd / 3) < 5e-3, (got, rl / 3, rd / 3)
    # Adam's first updates are ~lr*sign(g): a near-zero gradient whose sign differs between the two runs (bf16 weight
    # rounding after the first, 1e-9-different, update) moves a parameter by up to 2*lr per step - bound, not equality
    # (the optimizers themselves are compared on IDENTICAL gradients in test_tail_gpu.py)
    for (k, p), q in zip(a.named_parameters(), b.parameters()):
        d = (p.detach() - q.detach()).abs()
        assert float(d.max()) <= 3 * 2 * 1e-4 and float(d.mean()) <= 2e-5, k
    va = T.validate_one_epoch(a, data, BCEDiceWithLogits(), dev, "unet", None)
    vb = T.validate_one_epoch(b, data, crit, dev, "unet", None)           # torch criterion + torch-op Dice path
    assert abs(va[0] - vb[0]) < 1e-4 and abs(va[1] - vb[1]) < 2e-3, (va, vb)
    # the eval numbers are those of the fp32 oracle on the same weights within the bf16 bar
    sd = {k: v.detach().cpu() for k, v in a.state_dict().items()}
    ref = sum(float(O.bce_with_logits(O.FORWARDS["unet"](copy.deepcopy(sd), i, False), m)) for i, m, _ in data) / 3
    assert abs(va[0] - ref) < 2e-2 * max(1.0, abs(ref))


def test_manager_wraps_nothing_on_one_gpu_and_checkpoints_roundtrip(tmp_path):
    mgr = T.MultiGPUManager(device="cuda:0")
    m = mgr.setup_model_for_gpu(seeded_model("resunet"))
    assert next(m.parameters()).is_cuda and not hasattr(m, "module")
    path = str(tmp_path / "resunet_last.pth")
    mgr.save_model_state(m, path)
    sd = torch.load(path)
    assert len(sd) == 127 and all(v.device.type == "cpu" for v in sd.values())
    fresh = mgr.setup_model_for_gpu(seeded_model("resunet"))
    with torch.no_grad():
        for p in fresh.parameters():
            p.add_(1.0)
    mgr.load_model_state(fresh, path, torch.device("cuda", 0))
    for k, v in fresh.state_dict().items():
        assert torch.equal(v.cpu(), sd[k]), k


def test_decoded_uint8_batches_match_the_reference_dataset_transform():
    """datasets.py:40-44,59: ToTensor + Normalize(ImageNet) on the CPU, mask > 0.5 - against the same bytes sent as
    uint8 and normalised inside the stem kernel / thresholded on the device."""
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(9)
    img8 = torch.randint(0, 256, (2, 64, 48, 3), generator=g, dtype=torch.uint8)          # NHWC, as decoded
    mask8 = torch.randint(0, 256, (2, 1, 64, 48), generator=g, dtype=torch.uint8)
    mean, std = torch.tensor(T.IMAGENET_MEAN), torch.tensor(T.IMAGENET_STD)
    ref_img = (img8.permute(0, 3, 1, 2).float().div(255) - mean[None, :, None, None]) / std[None, :, None, None]
    ref_mask = (mask8.float().div(255) > 0.5).float()
    model = seeded_model("unet").to(dev).eval()
    with pytest.raises(RuntimeError, match="set_input_normalization"):
        model(img8.to(dev))
    model.set_input_normalization(T.IMAGENET_MEAN, T.IMAGENET_STD)
    x, m = T.to_device_batch(img8, mask8, dev)
    assert x.dtype == torch.uint8 and torch.equal(m.cpu(), ref_mask)
    with torch.no_grad():
        a = model(x)
        b = model(ref_img.to(dev))
    assert torch.equal(a, b)                                    # same bf16 stem operand -> bit-identical logits
    # and through a training step: gradients identical as well
    model.train()
    outs = []
    for inp in (x, ref_img.to(dev)):
        model.zero_grad(set_to_none=True)
        loss = torch.nn.functional.binary_cross_entropy_with_logits(model(inp), m)
        loss.backward()
        outs.append((float(loss.detach()), model.out.conv.weight.grad.clone()))
    assert outs[0][0] == outs[1][0] and torch.equal(outs[0][1], outs[1][1])
    with pytest.raises(RuntimeError):
        seeded_model("unet", precision="fp32").to(dev).set_input_normalization(T.IMAGENET_MEAN, T.IMAGENET_STD)(x)


@pytest.mark.parametrize("N,cin,H,W", [(2, 3, 32, 48), (1, 1, 16, 16), (2, 7, 24, 72), (1, 3, 9, 130)])
def test_uint8_stem_operand_is_bit_identical_to_the_normalised_fp32_operand(N, cin, H, W, monkeypatch):
    """uz_stem_im2col_u8 (tiled kernel and the first-generation per-pixel one) produces exactly the im2col operand that
    uz_stem_im2col builds from the tensor the reference's dataset transform would have produced on the CPU."""
    from unet_zoo_b200 import engine as E
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(13)
    img8 = torch.randint(0, 256, (N, H, W, cin), generator=g, dtype=torch.uint8)
    mean = [0.1 + 0.1 * c for c in range(cin)]
    std = [0.2 + 0.05 * c for c in range(cin)]
    ref = (img8.permute(0, 3, 1, 2).float().div(255) - torch.tensor(mean)[None, :, None, None]) \
        / torch.tensor(std)[None, :, None, None]
    ctx = E.Ctx(dev, training=False, record=False)
    want, _ = E.stem_input(ctx, ref.to(dev).contiguous())
    got = []
    for gen1 in (False, True):
        if gen1:
            monkeypatch.setenv("UZ_STEM_GEN1", "1")
        ctx = E.Ctx(dev, training=False, record=False)
        ctx.input_norm = (mean, std)
        a, kind = E.stem_input(ctx, img8.to(dev).contiguous().permute(0, 3, 1, 2))
        torch.cuda.synchronize()
        assert kind == "im2col" and a.C == want.C
        got.append(a.buf.data.clone())
    assert torch.equal(got[0], got[1])
    assert torch.equal(got[0], want.buf.data)
