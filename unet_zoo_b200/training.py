"""The callers either side of the hot path: drop-ins for the reference's per-epoch loops and
its multi-GPU manager, re-shaped for one process per GPU.

* :func:`train_one_epoch` / :func:`validate_one_epoch` keep the reference signatures and return values
  (unet_zoo/utils/training_loop.py:86-145,147-180) but never synchronise with the host inside the batch loop: the
  reference reads ``loss.item()`` and ``dc.item()`` every batch (:123-124) and branches on ``union == 0`` on the host
  (metrics.py:18); here loss and Dice are accumulated in device scalars and read once per epoch (and at the
  reference's own 50- / 100-batch logging points).
* :class:`MultiGPUManager` keeps ``setup_model_for_gpu / save_model_state / load_model_state /
  get_model_for_inference`` (unet_zoo/utils/multi_gpu.py:8-97) with the reference's checkpoint key format, but wraps
  the model in :class:`unet_zoo_b200.parallel.DataParallel` (bucketed NCCL all-reduce overlapped with backward) when a
  ``torch.distributed`` process group is up, instead of the single-process ``nn.DataParallel``.

Host-side orchestration only - the arithmetic is the model's forward/backward (C-ABI kernels) and, when the criterion
is :class:`unet_zoo_b200.loss.BCEDiceWithLogits` and the optimizer :class:`unet_zoo_b200.optim.FlatAdamW`, the
device-side tail kernels.
"""
import os
from typing import Optional, Tuple

import torch
import torch.nn as nn

MAX_GRAD_NORM = 1.0       # training_loop.py:101


def dice_coefficient(prediction: torch.Tensor, target: torch.Tensor, epsilon: float = 1e-7,
                     threshold: float = 0.5) -> torch.Tensor:
    """metrics.py:7-24 without the host round trip of ``if union == 0``: returns a 0-d tensor on the input's device."""
    pred = (torch.sigmoid(prediction) > threshold).float().reshape(-1)
    tgt = target.reshape(-1).float()
    inter = (pred * tgt).sum()
    union = pred.sum() + tgt.sum()
    return torch.where(union == 0, torch.ones_like(union), (2.0 * inter + epsilon) / (union + epsilon))


IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)   # datasets.py:43


def to_device_batch(img: torch.Tensor, mask: torch.Tensor, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """training_loop.py:109-110 (`img.float().to(device)`, `mask.float().to(device)`).  uint8 tensors - the decoded
    bytes: image [N,H,W,C], mask any shape with values 0..255 - are copied as they are (4x fewer bytes); the image is
    normalised inside the model's stem kernel (HotPathModel.set_input_normalization) and the mask becomes
    `(m/255 > 0.5).float()` (datasets.py:59) in one device pass (uz_mask_u8_to_f32)."""
    if img.dtype == torch.uint8:
        img = img.to(device, non_blocking=True)
    else:
        img = img.float().to(device, non_blocking=True)
    if mask.dtype == torch.uint8:
        from . import _lib as L
        m8 = mask.to(device, non_blocking=True).contiguous()
        if not m8.is_cuda:
            raise RuntimeError("unet_zoo_b200.training: uint8 masks are converted on the GPU (there is no CPU path)")
        out = torch.empty(m8.shape, dtype=torch.float32, device=m8.device)
        with torch.cuda.device(m8.device):
            L.check(L.load().uz_mask_u8_to_f32(m8.data_ptr(), m8.numel(), out.data_ptr(),
                                               torch.cuda.current_stream(m8.device).cuda_stream), "uz_mask_u8_to_f32")
        mask = out
    else:
        mask = mask.float().to(device, non_blocking=True)
    return img, mask


def _unwrap(model: nn.Module) -> nn.Module:
    return model.module if hasattr(model, "module") and isinstance(model.module, nn.Module) else model


def _loss_and_dice(outputs, masks, criterion) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """training_loop.py:43-83 for the hot-path models (plain tensor; NestedUNet deep supervision returns a list whose
    per-head losses are summed and whose last head is the main prediction, nested_unet.py:95-102)."""
    if isinstance(outputs, dict):
        # U2NET / U2NETP (training_loop.py:23-31, :61-64): every map against the mask (resized to the map's size - the
        # identity for these models), weights U2NET_LOSS_WEIGHTS = 1.0 for all seven keys; 'main' is the prediction
        outs = list(outputs.values())
        main = outputs["main"]
        sized = {}
    else:
        outs = list(outputs) if isinstance(outputs, (list, tuple)) else [outputs]
        main = outs[-1]
        sized = None
    loss = None
    dice = None
    for o in outs:
        tgt = masks
        if sized is not None and o.shape[2:] != masks.shape[2:]:
            key = tuple(o.shape[2:])
            if key not in sized:
                sized[key] = torch.nn.functional.interpolate(masks, size=key, mode="bilinear", align_corners=False)
            tgt = sized[key]
        term = criterion(o, tgt)
        loss = term if loss is None else loss + term
        if o is main and getattr(criterion, "last_dice", None) is not None:
            dice = criterion.last_dice              # uz_bce_dice already produced it in the same pass
    if dice is None:
        dice = dice_coefficient(main.detach(), masks)
    return loss, main, dice


def train_one_epoch(model: nn.Module, dataloader, optimizer, criterion, device, writer=None, epoch: int = 0,
                    model_name: str = "", logger=None) -> Tuple[float, float]:
    """One training epoch: per batch ``zero_grad -> forward -> loss (+Dice) -> backward -> clip(1.0) -> step``
    (training_loop.py:108-121).  Returns (average loss, average Dice) like the reference."""
    model.train()
    run_loss = torch.zeros((), dtype=torch.float32, device=device)
    run_dc = torch.zeros((), dtype=torch.float32, device=device)
    clips_itself = getattr(optimizer, "param_groups", [{}])[0].get("max_norm") is not None
    n = 0
    for idx, batch in enumerate(dataloader):
        img, mask = to_device_batch(batch[0], batch[1], device)
        optimizer.zero_grad()
        outputs = model(img)
        loss, main_logits, dc = _loss_and_dice(outputs, mask, criterion)
        loss.backward()
        if not clips_itself:                          # FlatAdamW(max_norm=...) clips inside its fused update
            torch.nn.utils.clip_grad_norm_(model.parameters(), MAX_GRAD_NORM)
        optimizer.step()
        run_loss += loss.detach()
        run_dc += dc.detach()
        n += 1
        if idx % 50 == 0 and logger is not None:      # training_loop.py:126-136 (the only per-batch host reads left)
            with torch.no_grad():
                ps = torch.sigmoid(main_logits.detach())
                logger.log_file_only(f"{model_name} - Batch {idx}: Loss={float(loss):.4f}, Dice={float(dc):.4f}")
                logger.log_file_only(f"  Pred stats (sigmoid): mean={float(ps.mean()):.4f}, max={float(ps.max()):.4f}, "
                                     f"mask_mean={float(mask.mean()):.4f}")
        if idx % 100 == 0 and writer is not None:     # training_loop.py:138-141
            step = epoch * len(dataloader) + idx
            writer.add_scalar(f"Batch/{model_name}_Train_Loss", float(loss), step)
            writer.add_scalar(f"Batch/{model_name}_Train_Dice", float(dc), step)
    n = max(n, 1)
    return float(run_loss) / n, float(run_dc) / n


def validate_one_epoch(model: nn.Module, dataloader, criterion, device, model_name: str = "",
                       logger=None) -> Tuple[float, float]:
    """training_loop.py:147-180: eval-mode forward, loss and Dice averaged over the batches."""
    model.eval()
    run_loss = torch.zeros((), dtype=torch.float32, device=device)
    run_dc = torch.zeros((), dtype=torch.float32, device=device)
    n = 0
    with torch.no_grad():
        for batch in dataloader:
            img, mask = to_device_batch(batch[0], batch[1], device)
            loss, _, dc = _loss_and_dice(model(img), mask, criterion)
            run_loss += loss
            run_dc += dc
            n += 1
    n = max(n, 1)
    return float(run_loss) / n, float(run_dc) / n


def init_distributed(backend: Optional[str] = None) -> Tuple[int, int, torch.device]:
    """One process per GPU under ``torchrun`` (replaces the reference's single-process ``nn.DataParallel`` setup,
    multi_gpu.py:20-37 / scripts/train.py:60-75): reads RANK / WORLD_SIZE / LOCAL_RANK, binds the process to its GPU and
    brings up the process group.  Returns (rank, world size, device); a plain ``python`` launch gives (0, 1, cuda:0)."""
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
        device = torch.device("cuda", local)
    else:
        device = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        # the step graph captures the bucketed NCCL all-reduces (train_step.GraphedTrainStep(capture_nccl=True))
        os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")
        backend = backend or ("nccl" if device.type == "cuda" else "gloo")
        if backend == "nccl":
            dist.init_process_group(backend, device_id=device)
        else:
            dist.init_process_group(backend)
    return rank, world, device


def train_model(model: nn.Module, train_dataloader, val_dataloader, optimizer, criterion, config, model_name: str,
                best_checkpoint_path: str, last_checkpoint_path: str, logger=None, gpu_manager=None, writer=None,
                early_stopping=None, lr_scheduler=None):
    """The reference's epoch driver (training_loop.py:182-285) on the drop-in loops above: per epoch
    ``train_one_epoch -> validate_one_epoch -> scheduler / early stopping -> save last checkpoint`` and the same return
    value ``(train_losses, train_dcs, val_losses, val_dcs, early_stopped)``.

    ``config`` needs ``EPOCHS`` and ``DEVICE``.  The reference builds its ``EarlyStopping``, ``DiceScheduler`` and
    TensorBoard ``SummaryWriter`` inside this function (:204-224); they are not part of the hot path, so here they are
    optional injected objects with the reference's call shapes (``early_stopping(val_dc, model, epoch)`` /
    ``.early_stop`` / ``.get_best_score()``, ``lr_scheduler.step(val_dc, epoch)``, ``writer.add_scalar(tag, v, step)``)
    - pass the reference's own instances, or nothing.  One fix over the reference: the BEST checkpoint is written
    whenever the validation Dice improves (the reference computes ``best_checkpoint_path`` but never writes it,
    training_loop.py:252 vs scripts/train_all.py:125)."""
    gpu_manager = gpu_manager or MultiGPUManager(config)
    device = getattr(config, "DEVICE", None) or next(model.parameters()).device
    epochs = int(getattr(config, "EPOCHS"))
    log = (lambda m: logger.log_both(m)) if logger is not None and hasattr(logger, "log_both") else (lambda m: None)
    train_losses, train_dcs, val_losses, val_dcs = [], [], [], []
    early_stopped, best = False, float("-inf")
    log(f"\nStarting training for {model_name} - {epochs} epochs on {device}")
    for epoch in range(epochs):
        sampler = getattr(train_dataloader, "sampler", None)
        if hasattr(sampler, "set_epoch"):                       # DistributedSampler: reshuffle per epoch
            sampler.set_epoch(epoch)
        tl, td = train_one_epoch(model, train_dataloader, optimizer, criterion, device, writer, epoch, model_name, logger)
        vl, vd = validate_one_epoch(model, val_dataloader, criterion, device, model_name, logger)
        train_losses.append(tl)
        train_dcs.append(td)
        val_losses.append(vl)
        val_dcs.append(vd)
        if lr_scheduler is not None:
            lr_scheduler.step(vd, epoch + 1)
            if hasattr(optimizer, "sync_hyper"):               # FlatAdamW keeps its hyper-parameters on the device
                optimizer.sync_hyper()
        if early_stopping is not None:
            early_stopping(vd, model, epoch + 1)
        gpu_manager.save_model_state(model, last_checkpoint_path)
        if vd > best:
            best = vd
            gpu_manager.save_model_state(model, best_checkpoint_path)
        if writer is not None:
            for tag, v in (("Train_Loss", tl), ("Train_Dice", td), ("Val_Loss", vl), ("Val_Dice", vd)):
                writer.add_scalar(f"Epoch/{model_name}_{tag}", v, epoch + 1)
            writer.add_scalar(f"{model_name}_Learning_Rate", optimizer.param_groups[0]["lr"], epoch + 1)
        log(f"{'-' * 60}\n{model_name} - Epoch {epoch + 1}/{epochs}\n  Train Loss: {tl:.6f} | Train DICE: {td:.6f}"
            f"\n  Val Loss:   {vl:.6f} | Val DICE:   {vd:.6f}\n  Learning Rate: {optimizer.param_groups[0]['lr']:.8f}"
            f"\n  Best Val Dice: {best:.6f}\n{'-' * 60}")
        if early_stopping is not None and getattr(early_stopping, "early_stop", False):
            log(f"\nEarly stopping triggered for {model_name} at epoch {epoch + 1}")
            early_stopped = True
            break
    log(f"Training {'stopped early' if early_stopped else 'completed'} for {model_name} after {len(train_losses)} epochs")
    return train_losses, train_dcs, val_losses, val_dcs, early_stopped


class MultiGPUManager:
    """multi_gpu.py:8-97 for one process per GPU.  ``config`` needs ``DEVICE`` only; when ``torch.distributed`` is
    initialised with more than one rank the model is wrapped in the bucketed-all-reduce data-parallel wrapper."""

    def __init__(self, config=None, device: Optional[torch.device] = None, bucket_mb: float = 25.0):
        self.config = config
        self.device = torch.device(device if device is not None else getattr(config, "DEVICE", "cuda"))
        self.bucket_mb = bucket_mb

    def _get_unwrapped_model(self, model: nn.Module) -> nn.Module:
        return _unwrap(model)

    def setup_model_for_gpu(self, model: nn.Module) -> nn.Module:
        import torch.distributed as dist
        model = model.to(self.device)
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            from .parallel import DataParallel
            model = DataParallel(model, bucket_mb=self.bucket_mb)
            print(f"Model wrapped for data-parallel training: rank {dist.get_rank()} of {dist.get_world_size()} "
                  f"on {self.device}")
        else:
            print(f"Model running on single GPU: {self.device}")
        return model

    def save_model_state(self, model: nn.Module, path: str):
        """Rank 0 writes the un-prefixed state_dict the reference writes (multi_gpu.py:39-42)."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_rank() != 0:
            return
        sd = {k: v.detach().cpu().clone() for k, v in _unwrap(model).state_dict().items()}
        torch.save(sd, path)

    @staticmethod
    def _strip_module_prefix(state_dict: dict) -> dict:
        return {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state_dict.items()}

    def load_model_state(self, model: nn.Module, path: str, device=None):
        """multi_gpu.py:55-87: tolerant of a leading ``module.``; strict first, then ``strict=False``."""
        if not os.path.exists(path):
            print(f"Warning: Checkpoint file not found at {path}. Model weights not loaded.")
            return model
        sd = self._strip_module_prefix(torch.load(path, map_location=device or self.device))
        target = _unwrap(model)
        try:
            target.load_state_dict(sd)
            print(f"Model weights loaded successfully from {path} onto {device or self.device}.")
        except RuntimeError as e:
            print(f"Error loading state_dict: {e}\nAttempting to load with `strict=False` (might load partial weights).")
            try:
                target.load_state_dict(sd, strict=False)
            except Exception as e2:   # noqa: BLE001 - mirror the reference: report, keep the random weights
                print(f"Failed to load even with strict=False: {e2}")
        return model

    def get_model_for_inference(self, model: nn.Module, device=None) -> nn.Module:
        m = _unwrap(model).to(device or self.device)
        m.eval()
        return m
