"""CPU experiment: how far does torch's own bf16 autocast of the oracle deviate from fp32 at a given shape / protocol?
Used to choose well-conditioned shapes for the full-resolution parity tests (the CUDA path's noise is comparable)."""
import copy
import sys
import time

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
sys.path.insert(0, __file__.rsplit("/", 2)[0] + "/tests")
from conftest import rel_err, seeded_model  # noqa: E402
from oracle import reference_port as O  # noqa: E402


def as_list(o):
    return list(o) if isinstance(o, (list, tuple)) else [o]


def condition(sd, glo, ghi, blo, bhi):
    return O.randomize_bn_affine(sd, (glo, ghi), (blo, bhi))


def main():
    names = sys.argv[1].split(",")
    n, s = int(sys.argv[2]), int(sys.argv[3])
    loss_name = sys.argv[4] if len(sys.argv) > 4 else "bce"
    loss_fn = {"proj": O.projection_loss, "gauss": O.gaussian_projection_loss}.get(loss_name, O.bce_with_logits)
    for name in names:
        kw = {"attn_dropout": 0.0} if name == "transatt_unet" else {}
        model = seeded_model(name, **kw)
        sd = copy.deepcopy(model.state_dict())
        if len(sys.argv) > 5:
            sd = condition(sd, *[float(v) for v in sys.argv[5].split(",")])
        x, mask = O.synthetic_batch(n, 3, s, s)
        t0 = time.time()
        sd32 = copy.deepcopy(sd)
        out32, loss32, g32 = O.train_step(name, sd32, x, mask, loss_fn=loss_fn, **kw)
        t1 = time.time()
        sd16 = copy.deepcopy(sd)
        with torch.autocast("cpu", dtype=torch.bfloat16):
            out16, loss16, g16 = O.train_step(name, sd16, x, mask, loss_fn=loss_fn, **kw)
        errs = [rel_err(a.detach().float(), b.detach()) for a, b in zip(as_list(out16), as_list(out32))]
        agree = [float((O.threshold_mask(a.detach().float()) == O.threshold_mask(b.detach())).float().mean())
                 for a, b in zip(as_list(out16), as_list(out32))]
        worst = []
        for k, ref in g32.items():
            if not k.endswith("weight") or ref.dim() < 2:
                continue
            den = ref.double().norm().clamp_min(1e-20)
            worst.append((float((g16[k].double() - ref.double()).norm() / den), k))
        worst.sort(reverse=True)
        # eval with the updated running stats
        with torch.no_grad():
            e32 = as_list(O.FORWARDS[name](copy.deepcopy(sd32), x, False, **kw))
            with torch.autocast("cpu", dtype=torch.bfloat16):
                e16 = as_list(O.FORWARDS[name](copy.deepcopy(sd32), x, False, **kw))
        eerr = [rel_err(a.float(), b) for a, b in zip(e16, e32)]
        eagree = [float((O.threshold_mask(a.float()) == O.threshold_mask(b)).float().mean()) for a, b in zip(e16, e32)]
        print(f"{name} n{n} {s}^2 {loss_name}: fp32 step {t1 - t0:.1f}s | train logits err {errs} agree {agree} | "
              f"eval err {eerr} agree {eagree} | loss {float(loss32):.5f}/{float(loss16):.5f} | "
              f"grad worst {[(round(e, 4), k) for e, k in worst[:4]]} median {worst[len(worst) // 2][0]:.4f}", flush=True)


if __name__ == "__main__":
    main()
