"""Debug helper (GPU box): per-model logits error of the CUDA path vs the fp32 oracle, next to the bf16-autocast floor."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import unet_zoo_b200
from oracle import reference_port as O

def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())

def aslist(o):
    return list(o) if isinstance(o, (list, tuple)) else [o]

for name, kw, shape in [("attention_unet", {}, (2, 3, 64, 64)), ("nested_unet", {}, (2, 3, 64, 64)), ("resunet", {}, (2, 3, 64, 64)),
                        ("unet", {}, (2, 3, 64, 64)), ("attention_unet", {}, (4, 3, 128, 128)), ("nested_unet", {}, (4, 3, 128, 128))]:
    torch.manual_seed(0)
    model = unet_zoo_b200.create_model(name, **kw)
    sd = copy.deepcopy(model.state_dict())
    x, mask = O.synthetic_batch(*shape)
    model.cuda().train()
    outs = [o.detach().cpu() for o in aslist(model(x.cuda()))]
    with torch.no_grad():
        ref = aslist(O.FORWARDS[name](copy.deepcopy(sd), x, True, **kw))
        with torch.autocast("cpu", dtype=torch.bfloat16):
            r16 = aslist(O.FORWARDS[name](copy.deepcopy(sd), x, True, **kw))
    print(name, shape, "train logits: ours", [round(rel(o, r), 4) for o, r in zip(outs, ref)],
          "autocast floor", [round(rel(a.float(), r), 4) for a, r in zip(r16, ref)], flush=True)
    # eval mode after the one training step above (running statistics updated by the CUDA path)
    sd2 = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        oe = [o.cpu() for o in aslist(model(x.cuda()))]
        re_ = aslist(O.FORWARDS[name](copy.deepcopy(sd2), x, False, **kw))
        with torch.autocast("cpu", dtype=torch.bfloat16):
            r16e = aslist(O.FORWARDS[name](copy.deepcopy(sd2), x, False, **kw))
    agree = lambda a, b: float((O.threshold_mask(a) == O.threshold_mask(b)).float().mean())
    print("    eval logits: ours", [round(rel(o, r), 4) for o, r in zip(oe, re_)], "floor", [round(rel(a.float(), r), 4) for a, r in zip(r16e, re_)],
          "mask agree ours", [round(agree(o, r), 5) for o, r in zip(oe, re_)], "floor", [round(agree(a.float(), r), 5) for a, r in zip(r16e, re_)], flush=True)
