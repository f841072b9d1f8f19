"""nn.Module <-> fused-executor boundary shared by the four hot-path models.

``forward(x)`` keeps the reference contract (NCHW float input -> raw NCHW logits, unet.py:29-44) while the whole
network runs as one ``torch.autograd.Function``: forward launches the fused kernels and records a tape, backward
replays the tape and hands one fp32 gradient per ``nn.Parameter`` back to autograd, so ``loss.backward()``,
``clip_grad_norm_`` and any ``torch.optim`` optimizer work unchanged (training_loop.py:112-121).
"""
import os
from typing import List, Optional

import torch
import torch.nn as nn

from .. import engine as E


class _HotPathFn(torch.autograd.Function):
    @staticmethod
    def forward(fctx, model, x, *params):
        ctx = E.Ctx(x.device, training=True, record=True)
        ctx.input_norm = model._input_norm
        plan = model._pack_plan
        cache = plan.run(ctx) if plan is not None and plan.device == x.device else None
        if cache is not None:
            ctx.pack_cache = cache
        elif os.environ.get("UZ_NO_PACK_PLAN") != "1":
            ctx.pack_log = []              # learn the plan from this step (finished at the end of backward)
        outs = model._graph(ctx, x)
        fctx.ectx = ctx
        fctx.model = model
        fctx.params = params
        model._last_launches = ctx.launches
        return tuple(outs)

    @staticmethod
    def backward(fctx, *douts):
        ctx: E.Ctx = fctx.ectx
        if ctx is None:
            raise RuntimeError("unet_zoo_b200: backward called twice on the same forward (retain_graph is not "
                               "supported: gradients overwrite the saved activations in place)")
        fctx.ectx = None
        model = fctx.model
        params = fctx.params
        ctx.stream = torch.cuda.current_stream(ctx.device).cuda_stream
        ctx.launches = 0
        make = getattr(model, "_uz_make_param_grads", None)
        ctx.pgrads = make(params, ctx.device) if make is not None else E.ParamGrads(params, ctx.device)
        model._last_param_grads = ctx.pgrads     # the flat gradient buffer of this step (parallel.sync_gradients)
        shadow = model._uz_shadow      # zero-padded operands standing in for the parameters (models/multiresunet.py)
        if shadow is not None:
            shadow.begin_backward(ctx.pgrads)
        ctx.out_grads = list(douts)
        for step in reversed(ctx.tape):
            step(ctx)
        ctx.join_side()                      # weight gradients ran on the side stream (engine.Ctx.on_side)
        ctx.tape.clear()
        if shadow is not None:
            shadow.end_backward(ctx.pgrads, params)
        if ctx.pack_log is not None and not torch.cuda.is_current_stream_capturing():
            sources = list(params) + (shadow.pack_sources() if shadow is not None else [])
            model._pack_plan = E.PackPlan(ctx.pack_log, sources, ctx.device)    # one launch packs them all from now on
            ctx.pack_log = None
        model._last_launches_bwd = ctx.launches
        finish = getattr(ctx.pgrads, "finish", None)
        if finish is not None:
            finish()
        grads = [ctx.pgrads.view(p) if p.requires_grad else None for p in params]
        return (None, None) + tuple(grads)


class HotPathModel(nn.Module):
    """Base class: input checks, train/eval dispatch, autograd wiring.  Subclasses implement ``_graph(ctx, x)``."""

    _spatial_multiple = 16
    # True: any H, W >= _spatial_multiple is accepted, as by the reference (whose Up / UpSample_UNet blocks pad the
    # upsampled tensor to an odd skip size, common_layers.py:110-113, :173-178).  False: H and W must be multiples of
    # _spatial_multiple - the reference's own forward raises on anything else for these models (their decoders
    # concatenate / add tensors of 2 * floor(s / 2) and s pixels without padding: attention_unet.py:90,
    # nested_unet.py:82, resunet.py:63).
    _any_size = False
    # "bf16": tensor-core kernels, forward + backward (default).  "fp32": forward-only parity mode on the FFMA kernels
    # of csrc/f32.cu (BASELINE.json: fp32 logits within 1e-4 of the reference) - see set_precision().
    precision = "bf16"

    def set_precision(self, precision: str):
        """Select the arithmetic of ``forward``: "bf16" (training and inference) or "fp32" (forward only)."""
        if precision not in ("bf16", "fp32"):
            raise ValueError(f"{type(self).__name__}: precision must be 'bf16' or 'fp32' (got {precision!r})")
        self.precision = precision
        return self

    _input_norm = None
    _uz_shadow = None           # models whose tensors need zero-padding to multiples of 8 channels keep their padded
                                # stand-ins here (refreshed from the parameters before every forward)
    _pack_plan = None           # engine.PackPlan: every weight pack of a training step in one launch (learned in step 1)
    _eval_cache = None          # engine.Ctx.cached: inference memo of packed / BatchNorm-folded weights

    def train(self, mode: bool = True):
        """nn.Module.train; entering train mode drops the inference memo (the optimizer is about to change the weights -
        FlatAdamW and CUDA-graph replays do so without touching torch's version counters)."""
        if mode:
            self._eval_cache = None
        return super().train(mode)

    def set_input_normalization(self, mean, std):
        """Let ``forward`` take the DECODED image bytes: a uint8 tensor [N, H, W, C] (as PIL / the dataset's Resize
        leaves it).  ToTensor (/255) and Normalize(mean, std) of the reference's dataset (datasets.py:40-44) then run
        inside the stem kernel (uz_stem_im2col_u8) and the host->device copy shrinks 4x.  None, None switches it off."""
        if mean is None or std is None:
            self._input_norm = None
            return self
        mean, std = [float(v) for v in mean], [float(v) for v in std]
        if len(mean) != self._in_channels or len(std) != self._in_channels or min(std) <= 0:
            raise ValueError(f"{type(self).__name__}: mean / std need {self._in_channels} entries, std > 0")
        self._input_norm = (mean, std)
        return self

    def _check_input(self, x: torch.Tensor) -> torch.Tensor:
        if not isinstance(x, torch.Tensor) or x.dim() != 4:
            raise RuntimeError(f"{type(self).__name__}: expected a 4-D NCHW tensor")
        if x.dtype == torch.uint8:
            if self._input_norm is None:
                raise RuntimeError(f"{type(self).__name__}: uint8 input needs set_input_normalization(mean, std)")
            if self.precision != "bf16" or 9 * self._in_channels > 64:
                raise RuntimeError(f"{type(self).__name__}: uint8 input is served by the bf16 im2col stem "
                                   "(in_channels <= 7) only")
            if x.shape[3] != self._in_channels:
                raise RuntimeError(f"{type(self).__name__}: uint8 input must be [N, H, W, {self._in_channels}] "
                                   f"(got {tuple(x.shape)})")
            x = x.contiguous().permute(0, 3, 1, 2)     # NCHW-shaped view of the NHWC bytes; checks below apply as is
        if not x.is_cuda:
            raise RuntimeError(f"{type(self).__name__}: input must be a CUDA tensor - this is the sm_100a build and "
                               "has no CPU path (move the model and the batch to the GPU)")
        p = next(self.parameters())
        if p.device != x.device:
            raise RuntimeError(f"{type(self).__name__}: parameters are on {p.device}, input on {x.device}")
        if p.dtype != torch.float32:
            raise RuntimeError(f"{type(self).__name__}: parameters must stay fp32 master weights (got {p.dtype}); "
                               "bf16 compute is internal")
        if x.shape[1] != self._in_channels:
            raise RuntimeError(f"{type(self).__name__}: expected {self._in_channels} input channels, got {x.shape[1]}")
        m = self._spatial_multiple
        if x.shape[2] < m or x.shape[3] < m:
            raise RuntimeError(f"{type(self).__name__}: H and W must be at least {m} (got {tuple(x.shape[2:])})")
        if not self._any_size and (x.shape[2] % m or x.shape[3] % m):
            raise RuntimeError(f"{type(self).__name__}: H and W must be multiples of {m} (got {tuple(x.shape[2:])}); "
                               "the reference's forward fails on such sizes too (its decoder joins tensors of "
                               "2 * floor(s / 2) and s pixels without padding)")
        if x.requires_grad:
            raise RuntimeError(f"{type(self).__name__}: gradients w.r.t. the input image are not implemented")
        return x if x.dtype == torch.uint8 else x.float().contiguous()

    def forward(self, x):
        x = self._check_input(x)
        with torch.cuda.device(x.device):
            self._before_forward(x.device)
            if self.precision == "fp32":
                if self.training and torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
                    raise RuntimeError(f"{type(self).__name__}: the fp32 mode is a forward-only parity mode - call it "
                                       "under torch.no_grad() (train-mode BatchNorm statistics are supported), or "
                                       "set_precision('bf16') to train")
                with torch.no_grad():
                    ctx = E.Ctx(x.device, training=self.training, record=False, precision="fp32")
                    outs = self._graph(ctx, x)
                    self._last_launches = ctx.launches
            elif self.training and torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
                outs = _HotPathFn.apply(self, x, *self.parameters())
            else:
                with torch.no_grad():
                    ctx = E.Ctx(x.device, training=self.training, record=False)
                    ctx.input_norm = self._input_norm
                    if not self.training:           # packed / folded weights are constants of an eval-mode forward
                        if self._eval_cache is None:
                            self._eval_cache = {}
                        ctx.eval_cache = self._eval_cache
                    outs = self._graph(ctx, x)
                    self._last_launches = ctx.launches
        return self._pack_outputs(list(outs))

    def _before_forward(self, device):
        """Hook that runs on the input's device before anything is launched (padded stand-ins of the parameters)."""

    def _pack_outputs(self, outs: List[torch.Tensor]):
        return outs[0]
