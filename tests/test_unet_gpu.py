"""Whole-model parity of `unet` on the GPU: CUDA path (through create_model / nn.Module) vs the CPU oracle and the
golden fixtures generated from the reference.  Tolerances are BASELINE.json's: bf16 compute within 2e-2 relative on
logits (max|a-b| / max|b|), thresholded-mask agreement, threshold indexing bit-exact on identical logits."""
import copy

import pytest
import torch

from conftest import conditioned_model, grad_errors, load_golden, parity_report, rel_err, seeded_model
from oracle import reference_port as O

pytestmark = pytest.mark.gpu
LOGIT_TOL_BF16 = 2e-2


def _cuda_train_step(model, x, mask, loss_fn=O.bce_with_logits):
    model.cuda().train()
    out = model(x.cuda())
    loss = loss_fn(out, mask.cuda())
    loss.backward()
    torch.cuda.synchronize()
    return out.detach().cpu(), float(loss)


def test_unet_train_step_against_golden_and_oracle():
    gold = load_golden("unet")
    model = seeded_model("unet")
    sd = copy.deepcopy(model.state_dict())
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    out, loss = _cuda_train_step(model, x, mask)
    assert out.shape == gold["logits_train"][0].shape and out.dtype == torch.float32
    assert rel_err(out, gold["logits_train"][0]) < LOGIT_TOL_BF16            # vs the reference's own output
    assert abs(loss - gold["loss"]) < 2e-3
    ref_out, ref_loss, ref_grads = O.train_step("unet", sd, x, mask)
    assert rel_err(out, ref_out.detach()) < LOGIT_TOL_BF16
    # BatchNorm running statistics after the step (momentum 0.1, unbiased running_var)
    got = model.state_dict()
    for k, v in gold["bn_after"].items():
        if "running_" in k:
            assert rel_err(got[k].cpu(), v) < 2e-2, k
        else:
            assert int(got[k]) == int(v), k
    # the well-conditioned gradients (head, last BatchNorm) match the fp32 reference tightly
    for k in ("out.conv.weight", "out.conv.bias", "up_convolution_4.conv.conv_op.4.weight"):
        g = dict(model.named_parameters())[k].grad.cpu()
        assert float((g - gold["grad_small"][k]).norm() / gold["grad_small"][k].norm()) < 2e-2, k
    # every parameter received a finite gradient of the right shape
    for k, p in model.named_parameters():
        assert p.grad is not None and p.grad.shape == p.shape and bool(torch.isfinite(p.grad).all()), k
    # the same on the conditioned fixture state (randomised BatchNorm affine parameters, Gaussian projection loss)
    cond = gold["cond"]
    model2, _ = conditioned_model("unet")
    out2, _ = _cuda_train_step(model2, x, mask, O.gaussian_projection_loss)
    assert rel_err(out2, cond["logits_train"][0]) < LOGIT_TOL_BF16
    assert abs(float(O.bce_with_logits(out2, mask)) - cond["bce"]) < 2e-3
    got2 = model2.state_dict()
    for k, v in cond["bn_after"].items():
        if "running_" in k:
            assert rel_err(got2[k].cpu(), v) < 2e-2, k
    g2 = dict(model2.named_parameters())
    for k in ("out.conv.weight", "out.conv.bias"):
        assert float((g2[k].grad.cpu() - cond["grad_small"][k]).norm() / cond["grad_small"][k].norm()) < 2e-2, k


def test_unet_gradients_against_fp32_reference_fixed_bars():
    """Every weight gradient of one train step (4 x 3 x 128 x 128, conditioned state, Gaussian projection loss) against
    the fp32 oracle, FIXED bars: all conv / convT weights concatenated < 0.12 rel. L2, worst single tensor < 0.3, the
    head's gradient < 2e-2 and the last block's (little accumulated bf16 noise; measured 2.6 % / 4.7 %) < 6e-2."""
    model, sd = conditioned_model("unet")
    x, mask = O.synthetic_batch(4, 3, 128, 128)
    out, _ = _cuda_train_step(model, x, mask, O.gaussian_projection_loss)
    ref_out, _, g32 = O.train_step("unet", sd, x, mask, loss_fn=O.gaussian_projection_loss)
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters()}
    e_all, per = grad_errors(grads, g32)
    near = {k: float((grads[k].double() - g32[k].double()).norm() / g32[k].double().norm())
            for k in ("out.conv.weight", "up_convolution_4.conv.conv_op.3.weight", "up_convolution_4.conv.conv_op.0.weight")}
    parity_report("unet128_grads", logits=rel_err(out, ref_out.detach()), grad_all=e_all, grad_worst=per[:3], near=near)
    assert rel_err(out, ref_out.detach()) < LOGIT_TOL_BF16
    assert e_all < 0.12 and per[0][0] < 0.3, (e_all, per[:3])
    assert near["out.conv.weight"] < 2e-2 and max(near.values()) < 6e-2, near


def test_unet_eval_forward_and_mask_agreement():
    gold = load_golden("unet")
    model = seeded_model("unet")
    x, mask = O.synthetic_batch(*gold["recipe"]["batch"])
    _cuda_train_step(model, x, mask)                    # one step of running-stat updates, like the fixture
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    model.eval()
    with torch.no_grad():
        out = model(x.cuda()).cpu()
        ref = O.unet_forward(sd, x, False)              # oracle on the SAME (CUDA-updated) statistics
    assert rel_err(out, ref) < LOGIT_TOL_BF16
    assert rel_err(out, gold["logits_eval"][0]) < 3e-2  # fixture used the reference's own fp32 statistics
    agree = float((O.threshold_mask(out) == O.threshold_mask(ref)).float().mean())
    assert agree >= 0.999, agree
    # bit-exact indexing: identical logits -> identical mask through the device-side formulation
    dev_mask = (torch.sigmoid(ref.cuda()) > 0.5).cpu()
    assert torch.equal(dev_mask, O.threshold_mask(ref))


def test_unet_mask_agreement_train_mode_larger_batch():
    """Train-mode (batch statistics) logits and masks at 4 x 3 x 128 x 128 on the conditioned state: plain bars."""
    model, sd = conditioned_model("unet")
    x, mask = O.synthetic_batch(4, 3, 128, 128)
    out, _ = _cuda_train_step(model, x, mask)
    with torch.no_grad():
        ref = O.unet_forward(sd, x, True)
    agree = float((O.threshold_mask(out) == O.threshold_mask(ref)).float().mean())
    parity_report("unet128_train", logits=rel_err(out, ref), agree=agree, positive=float(O.threshold_mask(ref).float().mean()))
    assert rel_err(out, ref) < LOGIT_TOL_BF16
    assert agree >= 0.999, agree


def test_unet_module_contract():
    model = seeded_model("unet").cuda()
    assert type(model).__name__ == "UNet"
    with pytest.raises(RuntimeError, match="at least 16"):
        model(torch.randn(1, 3, 12, 64, device="cuda"))
    with pytest.raises(RuntimeError, match="input channels"):
        model(torch.randn(1, 1, 64, 64, device="cuda"))
    model.eval()
    with torch.no_grad():
        y = model(torch.randn(1, 3, 256, 256, device="cuda"))      # README.md:74-90
    assert tuple(y.shape) == (1, 1, 256, 256)
    # optimizer / clipping work on the parameters exactly as in training_loop.py:119-121
    model.train()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-5)
    x, m = O.synthetic_batch(2, 3, 64, 64)
    losses = []
    for _ in range(3):
        opt.zero_grad()
        loss = O.bce_with_logits(model(x.cuda()), m.cuda())
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0]
    with pytest.raises(RuntimeError, match="backward called twice|retain_graph"):
        out = model(x.cuda())
        l2 = out.sum()
        l2.backward(retain_graph=True)
        l2.backward()


def test_graphed_train_step_matches_eager_steps():
    """GraphedTrainStep replays exactly the kernels of the eager step: same losses, same parameters after 3 steps."""
    from unet_zoo_b200.train_step import GraphedTrainStep
    x, m = O.synthetic_batch(2, 3, 64, 64)
    x, m = x.cuda(), m.cuda()
    crit = torch.nn.BCEWithLogitsLoss()
    results = []
    for graphed in (False, True):
        model = seeded_model("unet").cuda().train()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=1e-5, fused=True, capturable=True)
        losses = []
        if graphed:
            sd0 = copy.deepcopy(model.state_dict())
            step = GraphedTrainStep(model, crit, opt, x, m, max_norm=1.0, warmup=1)
            model.load_state_dict(sd0)                       # undo the warm-up / capture updates
            opt.state.clear() if False else None
            for st in opt.state.values():                    # reset AdamW moments and step counters in place
                for v in st.values():
                    if torch.is_tensor(v):
                        v.zero_()
            for _ in range(3):
                losses.append(float(step(x, m)))
        else:
            for _ in range(3):
                opt.zero_grad(set_to_none=True)
                loss = crit(model(x), m)
                loss.backward()
                torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
                opt.step()
                losses.append(float(loss))
        results.append((losses, {k: v.detach().clone() for k, v in model.state_dict().items()}))
    (l0, s0), (l1, s1) = results
    assert max(abs(a - b) for a, b in zip(l0, l1)) < 1e-6, (l0, l1)
    for k in s0:
        if s0[k].dtype.is_floating_point:
            assert torch.allclose(s0[k], s1[k], rtol=0, atol=1e-6), k


def test_data_parallel_and_side_stream_gradients_match_the_plain_step(monkeypatch):
    """Three ways to run the same step must give the same gradients bit for bit: everything on one stream, weight
    gradients on the side stream (engine.Ctx.on_side), and the bucketed NCCL wrapper (world size 1: the all-reduce is
    an identity, but buckets are launched on the communication stream while backward is still enqueueing)."""
    import torch.distributed as dist
    from unet_zoo_b200 import parallel as P
    x, m = O.synthetic_batch(2, 3, 64, 64)
    x, m = x.cuda(), m.cuda()
    crit = torch.nn.BCEWithLogitsLoss()

    def grads(wrap, overlap):
        monkeypatch.setenv("UZ_OVERLAP_WGRAD", "1" if overlap else "0")
        model = seeded_model("unet").cuda().train()
        net = P.DataParallel(model, bucket_mb=4.0) if wrap else model
        crit(net(x), m).backward()
        torch.cuda.synchronize()
        return {k: p.grad.clone() for k, p in model.named_parameters()}

    plain = grads(False, False)
    side = grads(False, True)
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29533", rank=0, world_size=1)
    try:
        wrapped = grads(True, True)
    finally:
        if created:
            dist.destroy_process_group()
    for k in plain:
        assert torch.equal(plain[k], side[k]), k
        assert torch.equal(plain[k], wrapped[k]), k


def test_split_graph_data_parallel_step_matches_eager_steps():
    """N > 1 launch mode: graph A (forward + backward into the flat gradient buffer), ONE eager NCCL all-reduce of that
    buffer (DataParallel.defer_sync / sync_gradients), graph B (clip + FlatAdamW).  With a world of one the all-reduce
    is an identity, so three replayed steps must reproduce three eager steps of the un-wrapped model."""
    import torch.distributed as dist
    from unet_zoo_b200 import parallel as P
    from unet_zoo_b200.loss import BCEDiceWithLogits
    from unet_zoo_b200.optim import FlatAdamW
    from unet_zoo_b200.train_step import GraphedTrainStep
    x, m = O.synthetic_batch(2, 3, 64, 64)
    x, m = x.cuda(), m.cuda()
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29534", rank=0, world_size=1)
    try:
        results = []
        for graphed in (False, True):
            model = seeded_model("unet").cuda().train()
            net = model
            if graphed:
                net = P.DataParallel(model)
                net.defer_sync = True
            params = list(model.parameters())
            opt = FlatAdamW(params, lr=1e-3, weight_decay=1e-5, max_norm=1.0)
            crit = BCEDiceWithLogits()
            losses = []
            if graphed:
                sd0 = copy.deepcopy(model.state_dict())
                step = GraphedTrainStep(net, crit, opt, x, m, max_norm=None, params=params, warmup=1,
                                        grad_sync=net.sync_gradients)
                assert step.graph_tail is not None and type(model._last_param_grads) is P.ParamGrads
                model.load_state_dict(sd0)                   # undo the warm-up updates (copies into the flat buffer)
                opt.exp_avg.zero_()
                opt.exp_avg_sq.zero_()
                opt.state.zero_()
                for _ in range(3):
                    losses.append(float(step(x, m)))
            else:
                for _ in range(3):
                    opt.zero_grad(set_to_none=True)
                    loss = crit(net(x), m)
                    loss.backward()
                    opt.step()
                    losses.append(float(loss))
            torch.cuda.synchronize()
            results.append((losses, {k: v.detach().clone() for k, v in model.state_dict().items()}))
    finally:
        if created:
            dist.destroy_process_group()
    (l0, s0), (l1, s1) = results
    assert max(abs(a - b) for a, b in zip(l0, l1)) < 1e-6, (l0, l1)
    for k in s0:
        if s0[k].dtype.is_floating_point:
            assert torch.allclose(s0[k], s1[k], rtol=0, atol=1e-6), k


def test_overlapped_graph_data_parallel_step_matches_eager_steps():
    """bench.py's default N > 1 launch mode: ONE CUDA graph per step with the bucketed NCCL all-reduces captured on the
    communication stream (GraphedTrainStep(capture_nccl=True)).  World of one (the all-reduce is an identity), small
    buckets so that several are launched mid-backward: three replayed steps reproduce three eager steps of the
    un-wrapped model; the warm-up's side effects are undone by the constructor (restore_after_warmup)."""
    import os

    import torch.distributed as dist
    from unet_zoo_b200 import parallel as P
    from unet_zoo_b200.loss import BCEDiceWithLogits
    from unet_zoo_b200.optim import FlatAdamW
    from unet_zoo_b200.train_step import GraphedTrainStep
    x, m = O.synthetic_batch(2, 3, 64, 64)
    x, m = x.cuda(), m.cuda()
    os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29535", rank=0, world_size=1)
    try:
        results = []
        for graphed in (False, True):
            model = seeded_model("unet").cuda().train()
            net = P.DataParallel(model, bucket_mb=8.0) if graphed else model
            params = list(model.parameters())
            opt = FlatAdamW(params, lr=1e-3, weight_decay=1e-5, max_norm=1.0)
            crit = BCEDiceWithLogits()
            losses = []
            if graphed:
                step = GraphedTrainStep(net, crit, opt, x, m, max_norm=None, params=params, warmup=2, capture_nccl=True)
                assert step.graph_tail is None and type(model._last_param_grads) is P.BucketedGrads
                assert len(model._last_param_grads.buckets) > 3
                assert float(opt.step_count) == 0.0          # warm-up undone: the optimizer has not stepped yet
                for _ in range(3):
                    losses.append(float(step(x, m)))
            else:
                for _ in range(3):
                    opt.zero_grad(set_to_none=True)
                    loss = crit(net(x), m)
                    loss.backward()
                    opt.step()
                    losses.append(float(loss))
            torch.cuda.synchronize()
            results.append((losses, {k: v.detach().clone() for k, v in model.state_dict().items()}))
    finally:
        if created:
            dist.destroy_process_group()
    (l0, s0), (l1, s1) = results
    assert max(abs(a - b) for a, b in zip(l0, l1)) < 1e-6, (l0, l1)
    for k in s0:
        if s0[k].dtype.is_floating_point:
            assert torch.allclose(s0[k], s1[k], rtol=0, atol=1e-6), k
        else:
            assert torch.equal(s0[k], s1[k]), k


def test_batched_weight_packs_equal_per_layer_packs(monkeypatch):
    """engine.PackPlan (all weight packs of a step in one launch, learned in step 1 and used from step 2 on) gives bit-identical
    logits and gradients to the per-layer packs, follows in-place weight updates (optimizer steps, load_state_dict) and
    is rebuilt when the parameters move."""
    x, mask = O.synthetic_batch(2, 3, 64, 64)
    x, mask = x.cuda(), mask.cuda()

    def run(plan: bool):
        monkeypatch.setenv("UZ_NO_PACK_PLAN", "0" if plan else "1")
        model = seeded_model("unet").cuda().train()
        opt = torch.optim.SGD(model.parameters(), lr=0.05)
        outs = []
        for step in range(3):
            opt.zero_grad(set_to_none=True)
            out = model(x)
            O.bce_with_logits(out, mask).backward()
            outs.append((out.detach().clone(), model.out.conv.weight.grad.detach().clone(),
                         model.down_convolution_2.conv.conv_op[0].weight.grad.detach().clone()))
            opt.step()
        assert (model._pack_plan is not None) == plan
        if plan:
            assert model._last_launches < 120          # one pack launch instead of one per convolution
            sd = {k: v.clone() for k, v in seeded_model("unet").state_dict().items()}
            model.load_state_dict(sd)                  # in place: same storage, the plan stays valid and sees the new values
            a = model(x).detach().clone()
            fresh = seeded_model("unet").cuda().train()
            monkeypatch.setenv("UZ_NO_PACK_PLAN", "1")
            b = fresh(x).detach()
            assert torch.equal(a, b)
        return outs

    for (o1, g1, h1), (o2, g2, h2) in zip(run(True), run(False)):
        assert torch.equal(o1, o2) and torch.equal(g1, g2) and torch.equal(h1, h2)
