"""ORACLE tooling - generates tests/golden/*.pt by running the UNMODIFIED reference modules on CPU.

Runs only in the build container (needs /root/reference).  The reference package cannot be imported as a whole
(its registry eagerly imports models that need `timm`, models/__init__.py:5-25), so the five files of the hot path
are loaded by file path under a synthetic package (route B).  Nothing is copied.

    python oracle/make_golden.py            # writes tests/golden/<model>.pt

Each fixture holds: the seed recipe, a fingerprint of the default-initialised state_dict (so tests can check that
`torch.manual_seed(0); create_model(name)` reproduces the same weights), the train-mode logits and BCE loss of one
synthetic batch, per-parameter gradient norms (+ full gradients of the small tensors), BatchNorm running statistics
after that step, and the eval-mode logits computed afterwards.
"""
import importlib.util
import os
import sys
import types

import torch

REF = "/root/reference/unet_zoo/models"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.dont_write_bytecode = True


def load_reference():
    pkg = types.ModuleType("refmodels")
    pkg.__path__ = [REF]
    sys.modules["refmodels"] = pkg
    mods = {}
    for name in ("common_layers", "unet", "attention_unet", "nested_unet", "resunet", "transatt_unet", "unet_transformer", "u2net", "multiresunet"):
        spec = importlib.util.spec_from_file_location(f"refmodels.{name}", os.path.join(REF, name + ".py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules[f"refmodels.{name}"] = m
        spec.loader.exec_module(m)
        mods[name] = m
    return mods


def build(mods, name, **kw):
    if name == "unet":
        return mods["unet"].UNet(in_channels=3, num_classes=1)
    if name == "attention_unet":
        return mods["attention_unet"].AttentionUNet(in_channels=3, num_classes=1, depth=5)
    if name == "nested_unet":
        return mods["nested_unet"].NestedUNet(in_channels=3, num_classes=1, depth=5,
                                              deep_supervision=kw.get("deep_supervision", False))
    if name == "resunet":
        return mods["resunet"].ResUnet(in_channels=3, num_classes=1, filters=[64, 128, 256, 512])
    if name == "transatt_unet":
        # ScaledDotProductAttention's nn.Dropout(0.1) draws from the global RNG in train mode; the main fixture is
        # taken with the knob of that nn.Dropout INSTANCE set to kw["attn_dropout"] (0.0: comparable across RNG
        # engines); the as-is (p = 0.1) train logits are stored next to it under a fixed seed
        model = mods["transatt_unet"].TransAttUNet(in_channels=3, num_classes=1, depth=5)
        model.sdpa.dropout.p = kw.get("attn_dropout", 0.1)
        return model
    if name == "unet_transformer":   # oracle groundwork for the next model (no CUDA path yet)
        return mods["unet_transformer"].U_Transformer(in_channels=3, num_classes=1)
    if name == "u2net":
        return mods["u2net"].U2NET(in_ch=3, out_ch=1)        # the factory's arguments (models/__init__.py:156-158)
    if name == "u2netp":
        return mods["u2net"].U2NETP(in_ch=3, out_ch=1)
    if name == "multiresunet":   # the factory's arguments (models/__init__.py:134-137; `depth` lands in **kwargs)
        return mods["multiresunet"].MultiResUnet(in_channels=3, num_classes=1)
    raise ValueError(name)


def as_list(out):
    """Model output -> list of logit maps: U2NET returns a dict (u2net.py:287-295), kept in its own key order."""
    if isinstance(out, dict):
        return list(out.values())
    return out if isinstance(out, list) else [out]


def fingerprint(sd):
    return {k: (float(v.double().sum()), float(v.double().abs().sum())) for k, v in sd.items()}


def conditioned_case(mods, name, kw, x, mask):
    """Second fixture per model on a WELL-CONDITIONED state: default init + randomised BatchNorm affine parameters
    (oracle.reference_port.randomize_bn_affine), one train step of the unmodified reference under the
    Gaussian projection loss, then its eval-mode forward.  The bf16 GPU tests hold the CUDA path to the north_star's
    plain bars against THESE outputs (train mode at gamma = 1 / beta = 0 amplifies rounding 1.2x per layer)."""
    from oracle.reference_port import bce_with_logits, gaussian_projection_loss, randomize_bn_affine
    torch.manual_seed(0)
    model = build(mods, name, **kw)
    model.load_state_dict(randomize_bn_affine({k: v.clone() for k, v in model.state_dict().items()}))
    model.train()
    outs = as_list(model(x))
    sum(gaussian_projection_loss(o, mask) for o in outs).backward()
    grads = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
    cond = {
        "logits_train": [o.detach().clone() for o in outs],
        "bce": float(sum(bce_with_logits(o.detach(), mask) for o in outs)),
        "grad_norm": {k: float(g.double().norm()) for k, g in grads.items()},
        "grad_small": {k: g for k, g in grads.items() if g.numel() <= 2048},
        "bn_after": {k: v.detach().clone() for k, v in model.state_dict().items()
                     if "running_" in k or "num_batches" in k},
    }
    model.eval()
    with torch.no_grad():
        oe = model(x)
    cond["logits_eval"] = [o.clone() for o in as_list(oe)]
    return cond


def main():
    from oracle.reference_port import synthetic_batch
    torch.set_num_threads(8)
    mods = load_reference()
    os.makedirs(OUT, exist_ok=True)
    cases = [("unet", {}), ("attention_unet", {}), ("nested_unet", {}), ("resunet", {}),
             ("nested_unet_ds", {"deep_supervision": True}), ("transatt_unet", {"attn_dropout": 0.0}), ("unet_transformer", {}),
             ("u2net", {}), ("u2netp", {}), ("multiresunet", {})]
    only = sys.argv[1:]
    if only:
        cases = [c for c in cases if c[0] in only]
    for tag, kw in cases:
        name = tag.replace("_ds", "")
        torch.manual_seed(0)
        model = build(mods, name, **kw)
        fp = fingerprint(model.state_dict())
        # u2net: 128 x 128 so that the innermost maps (six poolings) are 4 x 4, not 2 x 2
        size = 128 if name in ("u2net", "u2netp") else 64
        x, mask = synthetic_batch(2, 3, size, size)
        model.train()
        extra = {}
        if name == "transatt_unet":            # the reference as is: attention dropout p = 0.1 under a fixed seed
            sd0 = {k: v.clone() for k, v in model.state_dict().items()}
            model.sdpa.dropout.p = 0.1
            torch.manual_seed(77)
            with torch.no_grad():
                extra["logits_train_dropout"] = model(x).clone()
            extra["dropout_seed"] = 77
            model.sdpa.dropout.p = kw["attn_dropout"]
            model.load_state_dict(sd0)         # undo that forward's running-statistics update
        outs = as_list(model(x))
        crit = torch.nn.BCEWithLogitsLoss()
        loss = sum(crit(o, mask) for o in outs)
        loss.backward()
        grads = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
        gold = {
            "recipe": {"seed": 0, "batch": [2, 3, size, size], "seed_x": 1234, "seed_m": 4321, "kwargs": kw,
                       "torch": str(torch.__version__)},
            "fingerprint": fp,
            "logits_train": [o.detach().clone() for o in outs],
            "loss": float(loss),
            "grad_norm": {k: float(g.double().norm()) for k, g in grads.items()},
            "grad_small": {k: g for k, g in grads.items() if g.numel() <= 2048},
            "bn_after": {k: v.detach().clone() for k, v in model.state_dict().items()
                         if "running_" in k or "num_batches" in k},
        }
        if name in ("u2net", "u2netp"):     # fixture size: the default-init case keeps main + side1 (the conditioned case
            gold["logits_train"] = gold["logits_train"][:2]      # below keeps all seven maps); "loss" is over all seven
        gold.update(extra)
        model.eval()
        with torch.no_grad():
            oe = model(x)
        gold["logits_eval"] = [o.clone() for o in as_list(oe)][:2 if name in ("u2net", "u2netp") else None]
        gold["cond"] = conditioned_case(mods, name, kw, x, mask)
        # README example (README.md:74-90): eval forward on randn(1,3,256,256) -> shape only (fresh default weights)
        if tag == "unet":
            gold["readme_shape"] = list(model(torch.randn(1, 3, 256, 256)).shape)
        path = os.path.join(OUT, tag + ".pt")
        torch.save(gold, path)
        print(tag, "loss", gold["loss"], "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
