"""Loss + metric of the reference's training step in one device pass.

The reference computes, per batch, ``nn.BCEWithLogitsLoss()(logits, mask)`` (scripts/train.py:135,
training_loop.py:78), ``dice_coefficient(logits, mask)`` (metrics.py:7-24, including a host-synchronising
``if union == 0``) and later the loss gradient in autograd - about a dozen ATen kernels over the logits.
:class:`BCEDiceWithLogits` is a drop-in ``criterion``: calling it returns the same mean BCE loss (differentiable
w.r.t. the logits), and the Dice score of the same batch is left on the device in ``last_dice`` - both from ONE
kernel (``uz_bce_dice``) that also writes the loss gradient, so backward is a scalar multiply.
"""
import torch

from . import _lib as L


class _BCEDiceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, owner):
        if not logits.is_cuda:
            raise RuntimeError("unet_zoo_b200.loss: logits must be a CUDA tensor (there is no CPU path)")
        if logits.shape != target.shape:
            raise RuntimeError(f"unet_zoo_b200.loss: logits {tuple(logits.shape)} vs target {tuple(target.shape)}")
        lib = L.load()
        z = logits.detach().float().contiguous()
        y = target.detach().float().contiguous()
        n = z.numel()
        need_grad = logits.requires_grad
        dl = torch.empty_like(z) if need_grad else None
        result = torch.empty(8, dtype=torch.float32, device=z.device)
        ws = torch.empty(lib.uz_tail_workspace_bytes(), dtype=torch.uint8, device=z.device)
        with torch.cuda.device(z.device):
            rc = lib.uz_bce_dice(z.data_ptr(), y.data_ptr(), n, 1.0 / n, owner.epsilon,
                                 dl.data_ptr() if need_grad else None, result.data_ptr(), ws.data_ptr(), ws.numel(),
                                 torch.cuda.current_stream(z.device).cuda_stream)
        L.check(rc, "uz_bce_dice")
        owner.last_dice = result[1]
        owner.last_stats = result[2:5]          # intersection, prediction sum, target sum
        ctx.dl = dl
        ctx.shape = logits.shape
        ctx.dtype = logits.dtype
        return result[0].clone()

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.dl is None:
            return None, None, None
        g = (ctx.dl * grad_out).view(ctx.shape)
        return g.to(ctx.dtype), None, None


class BCEDiceWithLogits(torch.nn.Module):
    """``criterion(logits, mask) -> loss`` like ``nn.BCEWithLogitsLoss()``; ``last_dice`` holds
    ``dice_coefficient(logits, mask)`` of the same call as a device scalar (no host sync)."""

    def __init__(self, epsilon: float = 1e-7):
        super().__init__()
        self.epsilon = float(epsilon)     # metrics.py:7 default; the 0.5 threshold is fixed like the reference's default
        self.last_dice = None
        self.last_stats = None

    def forward(self, logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        return _BCEDiceFn.apply(logits, target, self)
