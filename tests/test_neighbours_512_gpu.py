"""The neighbouring models AT THE BASELINE RESOLUTION, 2 x 3 x 512 x 512: one training step and the
eval forward on the statistics it produced, against the fp32 CPU oracle on the same (conditioned) weights and inputs -
the same protocol as tests/test_fullres_parity_gpu.py uses for the four BASELINE models, with each model's documented
bars (tests/test_models_gpu.py, tests/test_u2net_gpu.py).  `multiresunet` has its own 512 x 512 test
in tests/test_multiresunet_gpu.py (it needs the rounding-matched oracle).

Different code runs here than at the 64 / 128-pixel fixtures: multi-wave persistent scheduling, split-K plans, the
4096-token attention of unet_transformer (at 64 x 64 input it sees 64 tokens), 32x bilinear upsampling of U^2-Net's
deepest side map."""
import copy
import json
import os

import pytest
import torch

from conftest import grad_errors, rel_err, seeded_model
from oracle import reference_port as O

pytestmark = pytest.mark.gpu
N, S = 2, 512
# (train logits, train masks, eval logits, eval masks, all-gradient rel. L2) - each model's documented bars
BARS = {"transatt_unet": (2e-2, 0.999, 2e-2, 0.999, 0.3),
        "unet_transformer": (1e-1, 0.98, 2e-2, 0.999, 0.7),      # train: see tests/test_models_gpu.py (documented exception)
        # u2netp at THIS shape: torch's own bf16 autocast of the reference deviates from fp32 by 3.5 / 3.9 / 2.6 / 2.5 /
        # 4.4 / 3.7 / 5.6 e-2 on the seven train-mode maps (masks 99.06 - 100 %) and up to 2.5e-2 in eval mode
        # (tools/parity_floor.py u2netp 2 512 gauss 0.8,1.2,1.0,2.0) - the max-norm criterion grows with the number of
        # pixels on a 112-layer model; this path measures 4.0 / 4.1 / 2.9 / 2.6 / 4.4 / 3.6 / 6.7 e-2 with the same masks
        # (profiles/r02_cw_neighbours_512.json): the same profile map by map, i.e. the function, not the kernels
        "u2netp": (8e-2, 0.985, 3e-2, 0.998, 0.6)}
KW = {"transatt_unet": {"attn_dropout": 0.0}}
_REPORT = {}


def _report(name, **kv):
    _REPORT.setdefault(name, {}).update(kv)
    path = os.environ.get("UZ_PARITY_REPORT_N512")
    if path:
        with open(path, "w") as f:
            json.dump(_REPORT, f, indent=1, sort_keys=True)


def _as_list(o):
    return list(o.values()) if isinstance(o, dict) else (list(o) if isinstance(o, (list, tuple)) else [o])


def _agree(a, b):
    return float((O.threshold_mask(a) == O.threshold_mask(b)).float().mean())


@pytest.mark.parametrize("name", list(BARS))
def test_train_step_and_eval_forward_at_512(name):
    torch.set_num_threads(16)
    kw = KW.get(name, {})
    t_tol, t_agree, e_tol, e_agree, g_all = BARS[name]
    model = seeded_model(name, **kw)             # transatt_unet: attention dropout off (comparable across RNG engines)
    sd = O.randomize_bn_affine(copy.deepcopy(model.state_dict()))
    model.load_state_dict(sd)
    sd = copy.deepcopy(sd)
    x, mask = O.synthetic_batch(N, 3, S, S)

    model.cuda().train()
    outs = _as_list(model(x.cuda()))
    sum(O.gaussian_projection_loss(o, mask.cuda()) for o in outs).backward()
    torch.cuda.synchronize()
    outs = [o.detach().cpu() for o in outs]
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters()}

    refs, _, ref_grads = O.train_step(name, sd, x, mask, loss_fn=O.gaussian_projection_loss, **kw)
    refs = [r.detach() for r in _as_list(refs)]
    errs = [rel_err(o, r) for o, r in zip(outs, refs)]
    agree = [_agree(o, r) for o, r in zip(outs, refs)]
    e_all, per = grad_errors(grads, ref_grads)
    worst_stat = max(rel_err(v.detach().cpu(), sd[k]) for k, v in model.state_dict().items() if "running_" in k)
    _report(name, train_logits=errs, train_agree=agree, grad_all=e_all, grad_worst=[[round(e, 4), k] for e, k in per[:3]],
            running_stats=worst_stat)
    assert len(outs) == len(refs) and all(o.shape == r.shape for o, r in zip(outs, refs))
    assert max(errs) < t_tol, errs
    assert min(agree) >= t_agree, agree
    assert e_all < g_all, (e_all, per[:3])
    assert worst_stat < 2e-2
    for k, g in grads.items():
        assert bool(torch.isfinite(g).all()), k

    state = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        oe = [o.cpu() for o in _as_list(model(x.cuda()))]
        re = _as_list(O.FORWARDS[name](state, x, False, **kw))
    ee = [rel_err(o, r) for o, r in zip(oe, re)]
    ea = [_agree(o, r) for o, r in zip(oe, re)]
    _report(name, eval_logits=ee, eval_agree=ea)
    assert max(ee) < e_tol, ee
    assert min(ea) >= e_agree, ea
