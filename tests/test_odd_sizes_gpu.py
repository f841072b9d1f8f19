"""Input sizes that are NOT multiples of 16 (`unet`, `transatt_unet`): the reference accepts them - MaxPool2d(2,2) floors,
and UpSample_UNet / Up zero-pad the upsampled tensor on the bottom / right to the skip tensor's size
(common_layers.py:110-113, :173-178) - so the drop-in must too.  The CUDA path against the fp32 oracle, same bars as
everywhere else (bf16: 2e-2 on logits; fp32 mode: 1e-4), on the conditioned state (see conftest.conditioned_model)."""
import pytest
import torch

from conftest import conditioned_model, grad_errors, parity_report, rel_err, seeded_model
from oracle import reference_port as O

pytestmark = pytest.mark.gpu


def _batch(shape, k=1, seed=5):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(*shape, generator=g)
    m = (torch.rand(shape[0], k, *shape[2:], generator=g) > 0.5).float()
    return x, m


@pytest.mark.parametrize("name,kw,shape", [
    ("unet", {}, (2, 3, 50, 70)),            # even sizes whose halvings turn odd: 50, 25, 12, 6, 3
    ("unet", {}, (1, 3, 33, 47)),            # odd at every level
    ("unet", {}, (3, 3, 17, 16)),            # smallest: 17 -> 8 -> 4 -> 2 -> 1
    ("unet", {"in_channels": 9, "num_classes": 2}, (1, 9, 45, 38)),   # NHWC stem path, K = 2 head
    ("transatt_unet", {"attn_dropout": 0.0}, (2, 3, 52, 44)),
    ("transatt_unet", {"attn_dropout": 0.0}, (1, 3, 35, 61)),
])
def test_train_step_and_eval_forward_on_odd_sizes(name, kw, shape):
    model, sd = conditioned_model(name, **kw)
    x, mask = _batch(shape, kw.get("num_classes", 1))
    model.cuda().train()
    out = model(x.cuda())
    O.gaussian_projection_loss(out, mask.cuda()).backward()
    torch.cuda.synchronize()
    fkw = {k: v for k, v in kw.items() if k in ("attn_dropout", "deep_supervision")}     # forward-time keywords only
    ref_out, _, ref_grads = O.train_step(name, sd, x, mask, loss_fn=O.gaussian_projection_loss, **fkw)
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters()}
    e_all, per = grad_errors(grads, ref_grads)
    e_train = rel_err(out.detach().cpu(), ref_out.detach())
    sd_eval = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        ev = model(x.cuda()).cpu()
        ref_ev = O.FORWARDS[name](sd_eval, x, False, **fkw)
    e_eval = rel_err(ev, ref_ev)
    parity_report(f"odd/{name}{shape}", train=e_train, eval=e_eval, grad_all=e_all, grad_worst=per[:2])
    assert out.shape == ref_out.shape
    assert e_train < 2e-2, e_train
    assert e_eval < 2e-2, e_eval
    for k, v in sd.items():                                   # running statistics after the step (oracle updated sd)
        if "running_" in k:
            assert rel_err(sd_eval[k], v) < 2e-2, k
    assert all(bool(torch.isfinite(g).all()) for g in grads.values())
    # fixed bars; the worst-tensor bar only where every BatchNorm sees at least 16 samples (17 x 16 images reach a
    # 1 x 1 bottleneck: statistics over the 3 images of the batch, gradients through them are noise-dominated in bf16)
    assert e_all < 0.3, (e_all, per[:3])
    if shape[0] * (shape[2] // 16) * (shape[3] // 16) >= 16:
        assert per[0][0] < 0.3, per[:3]


def test_fp32_mode_on_odd_sizes():
    model = seeded_model("unet", precision="fp32").cuda()
    x, _ = _batch((2, 3, 37, 52))
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    for training in (True, False):
        model.train(training)
        with torch.no_grad():
            out = model(x.cuda()).cpu()
            ref = O.unet_forward({k: v.clone() for k, v in sd.items()}, x, training)
        assert rel_err(out, ref) < 1e-4, (training, rel_err(out, ref))
        model.load_state_dict(sd)                              # undo the train-mode running-stat update


def test_models_whose_reference_rejects_odd_sizes_say_so():
    for name in ("attention_unet", "nested_unet", "resunet"):
        model = seeded_model(name).cuda()
        with pytest.raises(RuntimeError, match="multiples of"):
            model(torch.randn(1, 3, 50, 64, device="cuda"))
