"""Data-parallel training step: one process per GPU, bucketed gradient all-reduce overlapped with backward.

Replaces the reference's single-process ``nn.DataParallel`` wrap (unet_zoo/utils/multi_gpu.py:28-31): the batch is
sharded across ranks (each rank normalises BatchNorm over its local shard, exactly like DataParallel replicas do),
and gradients are averaged with NCCL over NVLink/NVSwitch.  The weight-gradient kernels write straight into one flat
fp32 buffer; it is cut into ~25 MB buckets in reverse registration order (the order backward completes them), and as
soon as the tape reports the last gradient of a bucket an all-reduce is launched on a side stream while the remaining
backward kernels keep running.  ``clip_grad_norm_`` / the optimizer then see averaged gradients
(training_loop.py:119-121 order).

Second mode (``DataParallel.defer_sync = True``): backward only fills the flat buffer and the caller averages it
afterwards with ONE all-reduce (``sync_gradients()``).  That is what lets the step run from CUDA graphs at N > 1
(``train_step.GraphedTrainStep(..., grad_sync=dp.sync_gradients)``): forward + backward replay as one graph, the
all-reduce is issued eagerly behind it, clip + AdamW replay as a second graph - no NCCL call is ever captured.  The
all-reduce is no longer hidden (124 MB over NVSwitch for unet), but the ~300 eager launches per step are.
"""
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn as nn

from .engine import ParamGrads, side_stream


def plan_buckets(numels: Sequence[int], bucket_bytes: int) -> List[List[int]]:
    """Parameter indices grouped into buckets, walking the registration order backwards."""
    buckets, cur, size = [], [], 0
    for i in reversed(range(len(numels))):
        cur.append(i)
        size += numels[i] * 4
        if size >= bucket_bytes:
            buckets.append(cur)
            cur, size = [], 0
    if cur:
        buckets.append(cur)
    return buckets


class BucketedGrads(ParamGrads):
    """ParamGrads whose contiguous bucket ranges are all-reduced (averaged) as soon as they are complete."""

    def __init__(self, params, device, group, bucket_bytes, comm_stream):
        super().__init__(params, device)
        self.group = group
        self.world = dist.get_world_size(group)
        self.comm_stream = comm_stream
        self.params = list(params)
        numels = [(p.numel() + 3) // 4 * 4 for p in self.params]
        self.buckets = plan_buckets(numels, bucket_bytes)
        self.bucket_of = {}
        self.pending = []
        self.ranges = []
        for b, idxs in enumerate(self.buckets):
            lo = min(self.offsets[id(self.params[i])][0] for i in idxs)
            hi = max(self.offsets[id(self.params[i])][0] + numels[i] for i in idxs)
            self.ranges.append((lo, hi))
            need = 0
            for i in idxs:
                self.bucket_of[id(self.params[i])] = b
                if self.params[i].requires_grad:
                    need += 1
            self.pending.append(need)
        self.launched = [False] * len(self.buckets)
        self.on_ready = self._param_ready
        if comm_stream is not None:
            self.flat.record_stream(comm_stream)

    def _param_ready(self, p):
        b = self.bucket_of[id(p)]
        self.pending[b] -= 1
        if self.pending[b] == 0:
            self._launch(b)

    def _launch(self, b):
        if self.launched[b]:
            return
        self.launched[b] = True
        lo, hi = self.ranges[b]
        chunk = self.flat[lo:hi]
        if self.comm_stream is not None:
            dev = self.flat.device
            cur = torch.cuda.current_stream(dev)
            side = side_stream(dev)                       # weight gradients are produced there
            self.comm_stream.wait_stream(cur)
            if side != cur:
                # under CUDA-graph capture the side stream is only part of the capture once a weight gradient was
                # issued on it; an event recorded on a stream OUTSIDE the capture must not be waited for from inside
                # (nothing of this step can be pending there in that case)
                with torch.cuda.stream(side):
                    side_in_capture = torch.cuda.is_current_stream_capturing()
                if side_in_capture or not torch.cuda.is_current_stream_capturing():
                    self.comm_stream.wait_stream(side)
            with torch.cuda.stream(self.comm_stream):
                dist.all_reduce(chunk, op=dist.ReduceOp.AVG, group=self.group)
        else:  # CPU / gloo (tests): no AVG reduction there
            dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group)
            chunk.div_(self.world)

    def finish(self):
        """Flush buckets that never completed (unused parameters) and make the compute stream wait for the comm."""
        for b in range(len(self.buckets)):
            self._launch(b)
        if self.comm_stream is not None:
            torch.cuda.current_stream(self.flat.device).wait_stream(self.comm_stream)


class DataParallel(nn.Module):
    """Drop-in for ``MultiGPUManager.setup_model_for_gpu`` (multi_gpu.py:20-37) under torchrun: wraps a hot-path
    model, keeps the ``.module`` access pattern (multi_gpu.py:16-18), broadcasts rank 0's parameters and buffers once,
    and averages gradients across ranks during backward."""

    def __init__(self, module: nn.Module, bucket_mb: float = 25.0, process_group=None):
        super().__init__()
        if not dist.is_initialized():
            raise RuntimeError("unet_zoo_b200.parallel.DataParallel needs torch.distributed to be initialised")
        self.module = module
        self.group = process_group
        self.bucket_bytes = int(bucket_mb * (1 << 20))
        p0 = next(module.parameters())
        self.comm_stream = torch.cuda.Stream(device=p0.device) if p0.is_cuda else None
        self.defer_sync = False      # True: no all-reduce during backward, the caller runs sync_gradients() after it
        with torch.no_grad():
            for t in list(module.parameters()) + list(module.buffers()):
                dist.broadcast(t, src=0, group=self.group)
        module._uz_make_param_grads = self._make_param_grads

    def _make_param_grads(self, params, device):
        if self.defer_sync:
            return ParamGrads(params, device)
        return BucketedGrads(params, device, self.group, self.bucket_bytes, self.comm_stream)

    def sync_gradients(self):
        """Average the flat gradient buffer of the last backward across ranks with one all-reduce on the current
        stream (``defer_sync`` mode).  The buffer is the one ``p.grad`` of every parameter is a slice of."""
        pg = getattr(self.module, "_last_param_grads", None)
        if pg is None:
            raise RuntimeError("DataParallel.sync_gradients: no backward has run on the wrapped model yet")
        if isinstance(pg, BucketedGrads):
            raise RuntimeError("DataParallel.sync_gradients: the last backward already all-reduced its buckets "
                               "(set defer_sync = True before the step)")
        if pg.flat.is_cuda:
            dist.all_reduce(pg.flat, op=dist.ReduceOp.AVG, group=self.group)
        else:  # CPU / gloo (tests): no AVG reduction there
            dist.all_reduce(pg.flat, op=dist.ReduceOp.SUM, group=self.group)
            pg.flat.div_(dist.get_world_size(self.group))

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def state_dict(self, *args, **kwargs):
        """Checkpoints keep the un-prefixed key format the reference writes (multi_gpu.py:39-42)."""
        return self.module.state_dict(*args, **kwargs)

    def load_state_dict(self, sd, *args, **kwargs):
        sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}
        return self.module.load_state_dict(sd, *args, **kwargs)
