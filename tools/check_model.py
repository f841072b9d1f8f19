"""Ad-hoc GPU diagnostic: CUDA path vs the CPU oracle for one model (prints errors; tests/ hold the real gates)."""
import copy
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import unet_zoo_b200  # noqa: E402
from oracle import reference_port as O  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "unet"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2
HW = int(sys.argv[3]) if len(sys.argv) > 3 else 64
kw = {}
torch.manual_seed(0)
model = unet_zoo_b200.create_model(name, **kw)
sd = copy.deepcopy(model.state_dict())
x, mask = O.synthetic_batch(N, 3, HW, HW)
LOSS = O.projection_loss if os.environ.get("UZ_PROJ") else O.bce_with_logits
out_o, loss_o, grads_o = O.train_step(name, sd, x, mask, loss_fn=LOSS)
model.cuda().train()
xg, mg = x.cuda(), mask.cuda()
out = model(xg)
loss = LOSS(out, mg)
loss.backward()
torch.cuda.synchronize()
rel = lambda a, b: float((a.double().cpu() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))
print("train logits rel err", rel(out.detach(), out_o.detach()), "loss", float(loss), float(loss_o))
agree = ((torch.sigmoid(out.detach().cpu()) > 0.5) == (torch.sigmoid(out_o.detach()) > 0.5)).float().mean()
print("mask agreement", float(agree))
worst = []
for k, p in model.named_parameters():
    g, go = p.grad.detach().cpu().double(), grads_o[k].double()
    e = float((g - go).norm() / go.norm().clamp_min(1e-30))
    worst.append((e, k, float(go.norm())))
if os.environ.get("UZ_ALL"):
    for e, k, n in worst:
        print(f"  grad rel-l2 err {e:.4f}  {k}  |g|={n:.3e}")
worst = [w for w in worst if w[2] > 1e-6]
worst.sort(reverse=True)
for e, k, n in worst[:12]:
    print(f"  grad rel-l2 err {e:.4f}  {k}  |g|={n:.3e}")
print("median grad err", sorted(w[0] for w in worst)[len(worst) // 2])
sdg = model.state_dict()
bn_err = max(rel(sdg[k], v) for k, v in sd.items() if "running_" in k)
print("running stats rel err", bn_err)
model.eval()
with torch.no_grad():
    oe = model(xg)
    oo = O.FORWARDS[name](sd, x, False)
print("eval logits rel err (oracle stats)", rel(oe, oo))
print("launches fwd/bwd", model._last_launches, getattr(model, "_last_launches_bwd", None))
