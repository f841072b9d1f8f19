"""Gradient clipping + AdamW over flat buffers.

The reference's step ends with ``torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)`` and
``optim.AdamW(lr=1e-4, weight_decay=1e-5).step()`` (training_loop.py:120-121, scripts/train.py:156): multi-tensor
kernels over ~140 parameter tensors plus a handful of small reductions.  The hot-path models already deliver every
parameter gradient as a slice of ONE flat fp32 buffer (engine.ParamGrads), so :class:`FlatAdamW` re-seats the
parameters as slices of one flat buffer with the same layout and runs the whole tail as three launches of
``uz_adamw_flat``: partial sums of squares, one-thread prepare (norm, clip coefficient, bias corrections, step counter
in device memory), one fused update pass.  Nothing synchronises with the host; the call is CUDA-graph capturable.

Semantics are those of ``clip_grad_norm_`` + ``torch.optim.AdamW`` (decoupled weight decay, bias correction, no
amsgrad); ``param_groups[0]['lr']`` stays the knob the reference's ``DiceScheduler`` turns (lr_scheduler.py) - call
:meth:`sync_hyper` after changing it when the step is replayed from a CUDA graph.
"""
from typing import Iterable, Optional

import torch

from . import _lib as L


class FlatAdamW:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, max_norm: Optional[float] = None):
        self.params = list(params)
        if not self.params:
            raise ValueError("FlatAdamW: no parameters")
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FlatAdamW: parameters must be on a CUDA device (there is no CPU path) - move the model "
                               "to the GPU before creating the optimizer")
        for p in self.params:
            if p.device != dev or p.dtype != torch.float32:
                raise RuntimeError("FlatAdamW: all parameters must be fp32 tensors on one device")
        self.lib = L.load()
        self.device = dev
        self.offsets, total = [], 0
        for p in self.params:                       # same layout as engine.ParamGrads: 16-byte aligned slices
            self.offsets.append(total)
            total += (p.numel() + 3) // 4 * 4
        self.total = total
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, off in zip(self.params, self.offsets):
                view = self.flat[off:off + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view                        # the module's parameters now live inside the flat buffer
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.state = torch.zeros(8, dtype=torch.float32, device=dev)   # step, total_norm, clip_coef, bc1, sqrt(bc2)
        self.hyper = torch.zeros(8, dtype=torch.float32, device=dev)
        self.ws = torch.empty(self.lib.uz_tail_workspace_bytes(), dtype=torch.uint8, device=dev)
        self.param_groups = [dict(params=self.params, lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay,
                                  max_norm=max_norm)]
        self._hyper_host = None
        self.sync_hyper()

    # ------------------------------------------------------------------ hyper-parameters
    def sync_hyper(self):
        """Copy lr / betas / eps / weight_decay / max_norm of ``param_groups[0]`` to the device (no-op if unchanged)."""
        g = self.param_groups[0]
        mn = g.get("max_norm")
        host = (float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]),
                float(mn) if mn is not None else 0.0)
        if host != self._hyper_host:
            self.hyper[:6].copy_(torch.tensor(host, dtype=torch.float32), non_blocking=False)
            self._hyper_host = host

    # ------------------------------------------------------------------ torch.optim-like surface
    def zero_grad(self, set_to_none: bool = True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    def _flat_grad_ptr(self) -> int:
        g0 = self.params[0].grad
        if g0 is None:
            raise RuntimeError("FlatAdamW.step: parameters have no gradient (call backward first)")
        base = g0.data_ptr() - 4 * self.offsets[0]
        for p, off in zip(self.params, self.offsets):
            g = p.grad
            if g is None or g.dtype != torch.float32 or not g.is_contiguous() or g.data_ptr() != base + 4 * off:
                raise RuntimeError(
                    "FlatAdamW.step: gradients are not slices of one flat buffer in registration order - this "
                    "optimizer serves the unet_zoo_b200 hot-path models (engine.ParamGrads layout) and has no "
                    "per-tensor fallback; pass exactly model.parameters()")
            if p.data_ptr() != self.flat.data_ptr() + 4 * off:
                raise RuntimeError("FlatAdamW.step: a parameter was moved after the optimizer was created "
                                   "(model.to(...) / load into new storage); re-create the optimizer")
        return base

    @torch.no_grad()
    def step(self):
        """clip_grad_norm_(max_norm) + AdamW update; returns nothing, `total_norm` / `step_count` are device views."""
        if not torch.cuda.is_current_stream_capturing():
            self.sync_hyper()
        gptr = self._flat_grad_ptr()
        with torch.cuda.device(self.device):
            rc = self.lib.uz_adamw_flat(self.flat.data_ptr(), gptr, self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                                        self.total, self.hyper.data_ptr(), self.state.data_ptr(), self.ws.data_ptr(),
                                        self.ws.numel(), torch.cuda.current_stream(self.device).cuda_stream)
        L.check(rc, "uz_adamw_flat")

    @property
    def total_norm(self) -> torch.Tensor:
        """Gradient norm before clipping of the last step (what clip_grad_norm_ returns), device scalar."""
        return self.state[1]

    @property
    def step_count(self) -> torch.Tensor:
        return self.state[0]

    def state_dict(self):
        g = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        return {"flat_adamw": 1, "hyper": g, "step": float(self.state[0]), "exp_avg": self.exp_avg.clone(),
                "exp_avg_sq": self.exp_avg_sq.clone()}

    def load_state_dict(self, sd):
        if "flat_adamw" not in sd or sd["exp_avg"].numel() != self.total:
            raise ValueError("FlatAdamW.load_state_dict: not a FlatAdamW state for this parameter list")
        self.param_groups[0].update(sd["hyper"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.state.zero_()
        self.state[0] = float(sd["step"])
        self.sync_hyper()
