"""The reference's training-step body captured into ONE CUDA graph.

``training_loop.py:112-121`` of the reference runs, per batch: ``zero_grad -> forward -> criterion -> backward ->
clip_grad_norm_(1.0) -> optimizer.step()`` as ~300 separate launches issued from Python.  On a B200 the whole step is
~50 ms of GPU work, so launch latency and host jitter between the many small kernels (BatchNorm finalizes, weight
packing, reductions) are visible.  :class:`GraphedTrainStep` records the step once (every kernel of the C-ABI library
is stream-ordered and allocation-free, the executor's buffers come from the graph's private memory pool) and replays
it with a single ``cudaGraphLaunch`` per batch: no tracing compiler, the same kernels in the same order.

Data-parallel (N > 1), two modes:

* ``capture_nccl=True`` (bench.py's default): the whole step INCLUDING the bucketed NCCL all-reduces of
  ``parallel.DataParallel`` is one graph.  The all-reduces are captured on the wrapper's communication stream (forked
  from / joined to the capturing stream by events, like the weight-gradient side stream), so on replay they overlap the
  remaining backward kernels exactly as in the eager bucketed mode, at one host call per step.  Needs the process
  group's watchdog not to poll CUDA events while this thread captures: set ``TORCH_NCCL_ASYNC_ERROR_HANDLING=0`` before
  ``init_process_group`` (PyTorch's own requirement for whole-network capture with NCCL) - bench.py does.
* ``grad_sync=`` (``parallel.DataParallel.sync_gradients`` on a wrapper in ``defer_sync`` mode): no NCCL call is
  captured; the step becomes graph A (zero_grad -> forward -> criterion -> backward into the flat gradient buffer), one
  eager all-reduce of that buffer (exposed), graph B (clip + optimizer step): three host calls per step.

Constraints: fixed batch shape; an optimizer that can be captured (``torch.optim.AdamW(..., fused=True,
capturable=True)``); the learning-rate schedule, if any, must live in a device tensor.
"""
from typing import Callable, Iterable, Optional

import torch


class GraphedTrainStep:
    def __init__(self, model: torch.nn.Module, criterion: Callable, optimizer: torch.optim.Optimizer,
                 example_x: torch.Tensor, example_target: torch.Tensor, max_norm: Optional[float] = 1.0,
                 params: Optional[Iterable[torch.nn.Parameter]] = None, warmup: int = 3,
                 grad_sync: Optional[Callable[[], None]] = None, capture_nccl: bool = False,
                 restore_after_warmup: bool = True):
        if not example_x.is_cuda:
            raise RuntimeError("GraphedTrainStep needs CUDA tensors (there is no CPU path)")
        self.model, self.criterion, self.optimizer, self.max_norm = model, criterion, optimizer, max_norm
        self.params = [p for p in (params if params is not None else model.parameters()) if p.requires_grad]
        self.x = example_x.clone()
        self.target = example_target.clone()
        self.loss: Optional[torch.Tensor] = None
        self.out = None
        self.grad_sync = grad_sync
        dev = example_x.device
        # the warm-up runs on a side stream by design (torch's capture recipe); autograd's note about it is noise here
        warn_off = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
        if warn_off is not None:
            warn_off(False)
        # the warm-up steps and the capture itself are REAL training steps on the example batch: weights, optimizer
        # moments / step counter and BatchNorm running statistics are put back afterwards (in place - the graph holds
        # their addresses), so the caller's first step starts from the state it handed in
        saved = self._snapshot() if restore_after_warmup else None
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                      # warm-up off the capture: lazy initialisation, autotuning
            for _ in range(max(1, warmup)):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        self.graph_tail = None
        if grad_sync is None and capture_nccl:
            # a process group is alive (its threads may touch the CUDA API meanwhile): thread-local capture mode
            with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
                self._body()
        elif grad_sync is None:
            with torch.cuda.graph(self.graph):
                self._body()
        else:
            # a process group is alive: its watchdog thread may touch the CUDA API while this thread captures
            with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
                self._fwd_bwd()
            self.graph_tail = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_tail, pool=self.graph.pool(), capture_error_mode="thread_local"):
                self._tail()
        if saved is not None:
            self._restore(saved)

    def _snapshot(self):
        opt = self.optimizer
        if hasattr(opt, "exp_avg_sq") and hasattr(opt, "flat"):        # unet_zoo_b200.optim.FlatAdamW
            ostate = ("flat", opt.exp_avg.clone(), opt.exp_avg_sq.clone(), opt.state.clone())
        else:
            ostate = ("torch", {id(p): {k: v.detach().clone() for k, v in st.items() if torch.is_tensor(v)}
                                for p, st in opt.state.items()})
        return {k: v.detach().clone() for k, v in self.model.state_dict().items()}, ostate

    @torch.no_grad()
    def _restore(self, saved):
        msd, ostate = saved
        for k, v in self.model.state_dict().items():
            v.copy_(msd[k])
        opt = self.optimizer
        if ostate[0] == "flat":
            opt.exp_avg.copy_(ostate[1])
            opt.exp_avg_sq.copy_(ostate[2])
            opt.state.copy_(ostate[3])
        else:
            for p, st in opt.state.items():
                old = ostate[1].get(id(p), {})
                for k, v in st.items():
                    if torch.is_tensor(v):                 # lazily created by the warm-up: back to its initial zero
                        v.copy_(old[k]) if k in old else v.zero_()
        torch.cuda.synchronize()

    def _body(self):
        self._fwd_bwd()
        if self.grad_sync is not None:
            self.grad_sync()
        self._tail()

    def _fwd_bwd(self):
        self.optimizer.zero_grad(set_to_none=True)
        self.out = self.model(self.x)
        if isinstance(self.out, dict):                     # U2NET: seven maps, loss weights 1.0 (training_loop.py:23-31)
            outs = list(self.out.values())
        else:
            outs = self.out if isinstance(self.out, (list, tuple)) else [self.out]
        loss = self.criterion(outs[0], self.target)
        for o in outs[1:]:                                 # deep supervision: the reference sums the per-head losses
            loss = loss + self.criterion(o, self.target)
        loss.backward()
        self.loss = loss.detach()

    def _tail(self):
        if self.max_norm is not None:
            torch.nn.utils.clip_grad_norm_(self.params, self.max_norm)
        self.optimizer.step()

    def __call__(self, x: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """Run one training step on (x, target); returns the (device) loss tensor of this step."""
        if x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x, non_blocking=True)
        if target.data_ptr() != self.target.data_ptr():
            self.target.copy_(target, non_blocking=True)
        self.graph.replay()
        if self.graph_tail is not None:
            self.grad_sync()               # eager: one all-reduce of the flat gradient buffer on the current stream
            self.graph_tail.replay()
        return self.loss


class GraphedInference:
    """Eval-mode forward of a hot-path model captured into one CUDA graph (fixed input shape): the ~56 launches of a
    resunet forward cost ~1.6 ms of host time in eager mode, which dominates below ~4 images of 512x512
    (evaluate.py / visualize.py style single-image calls, BASELINE config #5 small-batch corner)."""

    def __init__(self, model: torch.nn.Module, example_x: torch.Tensor, warmup: int = 2):
        if not example_x.is_cuda:
            raise RuntimeError("GraphedInference needs CUDA tensors (there is no CPU path)")
        if model.training:
            raise RuntimeError("GraphedInference captures the eval-mode forward: call model.eval() first")
        self.model = model
        self.x = example_x.clone()
        dev = example_x.device
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):
                model(self.x)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.out = model(self.x)

    def __call__(self, x: torch.Tensor):
        """Logits for x (same shape as the example); the returned tensor(s) are overwritten by the next call."""
        if x.shape != self.x.shape:
            raise RuntimeError(f"GraphedInference was captured for {tuple(self.x.shape)}, got {tuple(x.shape)}")
        if x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.out
