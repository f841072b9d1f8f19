"""Graph executor of the UNet hot path on the sm_100a kernel library.

The reference runs every layer as a separate ``torch.nn`` module call on NCHW fp32 tensors
(``unet_zoo/models/common_layers.py``).  Here a model's ``forward`` is described once as a sequence of fused
operators over NHWC bf16 activation views (:class:`Act`); each operator launches kernels of the C-ABI library
(``include/unet_zoo_b200.h``) and, in training mode, records a closure on a tape that computes its gradients.
``torch.cat`` never runs: producers write straight into channel slices of a pre-allocated wide buffer.

PyTorch is used for device memory (caching allocator), streams and the autograd boundary only.
"""
import contextlib
import ctypes
import os
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib as L

_c_void_p = ctypes.c_void_p

# Optional per-launch timing of the tensor-core kernels (bench.py's roofline leg): when set to a list, every
# uz_conv_igemm / uz_conv_wgrad / uz_bn_apply / uz_bn_bwd call is bracketed by CUDA events on the launch stream and
# appended as (kind, algorithmic FLOPs (GEMM kinds) or bytes (bn_* kinds), start_event, end_event, tag, algorithmic
# HBM bytes).  None (default) adds no overhead.
PROFILE: Optional[list] = None
_TAPS = {0: 9, 1: 9, 2: 1, 3: 9, 4: 9, 5: 1, 6: 1, 7: 1, 8: 4, 9: 9, 10: 36}


class _Timed:
    def __init__(self, kind, flops, tag="", nbytes=None):
        self.kind, self.flops, self.tag = kind, flops, tag
        self.nbytes = flops if nbytes is None else nbytes     # bandwidth kinds pass their bytes as `flops`

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *a):
        if PROFILE is not None:
            self.e1.record()
            PROFILE.append((self.kind, self.flops, self.e0, self.e1, self.tag, self.nbytes))


_MODE_NAMES = {L.CONV3X3: "conv3x3", L.CONV1X1: "conv1x1", L.CONVT2X2: "convT2x2", L.CONV3X3_DGRAD: "conv3x3.dgrad",
               L.CONVT2X2_DGRAD: "convT2x2.dgrad", L.CONV3X3_S2: "conv3x3s2", L.CONV1X1_S2: "conv1x1s2",
               L.UP2X_CONV3X3: "up2x+conv3x3", L.UP2X_CONV3X3_DGRAD: "up2x+conv3x3.dgrad"}
_SIDE_STREAMS: Dict[int, "torch.cuda.Stream"] = {}


def side_stream(device) -> "torch.cuda.Stream":
    """The per-device stream the weight-gradient GEMMs run on (see Ctx.on_side)."""
    idx = torch.device(device).index
    idx = torch.cuda.current_device() if idx is None else idx
    st = _SIDE_STREAMS.get(idx)
    if st is None:
        st = _SIDE_STREAMS[idx] = torch.cuda.Stream(device=idx)
    return st


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class Buf:
    """One NHWC device buffer [N, H, W, ld] (bf16; fp32 in the forward-only parity mode) plus (lazily) its gradient
    buffer of the same shape."""

    __slots__ = ("data", "grad", "written", "N", "H", "W", "ld", "want_gsum", "gsum", "esize")

    def __init__(self, N, H, W, ld, device, dtype=torch.bfloat16):
        self.data = torch.empty((N, H, W, ld), dtype=dtype, device=device)
        self.esize = self.data.element_size()
        self.grad = None
        self.written: List[Tuple[int, int]] = []  # channel intervals of `grad` that already hold a gradient
        self.N, self.H, self.W, self.ld = N, H, W, ld
        self.want_gsum = False   # a consumer of this buffer's gradient needs its per-channel sums (ConvT bias gradient)
        self.gsum: List[Tuple[int, int, torch.Tensor]] = []   # (c_lo, c_hi, stats[rows][2][c_hi-c_lo]) from a dgrad epilogue

    def grad_buf(self):
        if self.grad is None:
            self.grad = torch.empty_like(self.data)
        return self.grad


class Act:
    """Channel slice [c0, c0+C) of a :class:`Buf` - the unit every kernel call works on (pointer + pixel stride)."""

    __slots__ = ("buf", "c0", "C")

    def __init__(self, buf: Buf, c0: int = 0, C: Optional[int] = None):
        self.buf, self.c0 = buf, c0
        self.C = buf.ld - c0 if C is None else C
        if buf.esize == 2 and (self.c0 % 8 or self.C % 8):
            raise RuntimeError(f"unet_zoo_b200: channel slices must be multiples of 8 (c0={c0}, C={self.C})")

    N = property(lambda s: s.buf.N)
    H = property(lambda s: s.buf.H)
    W = property(lambda s: s.buf.W)
    ld = property(lambda s: s.buf.ld)
    ptr = property(lambda s: s.buf.data.data_ptr() + s.buf.esize * s.c0)
    npix = property(lambda s: s.buf.N * s.buf.H * s.buf.W)

    def slice(self, c0, C):
        return Act(self.buf, self.c0 + c0, C)

    def gptr(self):
        return self.buf.grad_buf().data_ptr() + 2 * self.c0

    def grad_writes(self) -> List[Tuple[int, int, bool]]:
        """Split this slice into maximal runs (c_off_in_slice, C, accumulate) and mark them written."""
        lo, hi = self.c0, self.c0 + self.C
        cuts = {lo, hi}
        for a, b in self.buf.written:
            if lo < a < hi:
                cuts.add(a)
            if lo < b < hi:
                cuts.add(b)
        cuts = sorted(cuts)
        runs = []
        for a, b in zip(cuts[:-1], cuts[1:]):
            acc = any(wa <= a and b <= wb for wa, wb in self.buf.written)
            if runs and runs[-1][2] == acc and runs[-1][0] + runs[-1][1] == a - lo:
                runs[-1] = (runs[-1][0], runs[-1][1] + (b - a), acc)
            else:
                runs.append((a - lo, b - a, acc))
        self.buf.written.append((lo, hi))
        for r0, rc, acc in runs:   # a second writer makes column sums taken by the first one stale
            if acc and self.buf.gsum:
                a, b = lo + r0, lo + r0 + rc
                self.buf.gsum = [e for e in self.buf.gsum if e[1] <= a or b <= e[0]]
        return runs

    def has_grad(self):
        lo, hi = self.c0, self.c0 + self.C
        return any(a < hi and lo < b for a, b in self.buf.written)

    def to_nchw(self) -> torch.Tensor:  # debugging / tests
        return self.buf.data[..., self.c0:self.c0 + self.C].permute(0, 3, 1, 2).float()


class GradAct(Act):
    """The gradient buffer of an :class:`Act` seen as an activation operand (dL/dy of a convolution without BatchNorm)."""

    __slots__ = ()

    def __init__(self, a: Act):
        super().__init__(a.buf, a.c0, a.C)

    ptr = property(lambda s: s.buf.grad_buf().data_ptr() + 2 * s.c0)


class ParamGrads:
    """Destination of parameter gradients: one flat fp32 buffer, parameters in registration order."""

    def __init__(self, params: Sequence[torch.nn.Parameter], device, on_ready: Optional[Callable] = None):
        self.offsets: Dict[int, Tuple[int, torch.Size]] = {}
        total = 0
        for p in params:
            self.offsets[id(p)] = (total, p.shape)
            total += (p.numel() + 3) // 4 * 4  # keep every gradient 16-byte aligned
        self.flat = torch.zeros(total, dtype=torch.float32, device=device)
        self.done = set()
        self.on_ready = on_ready
        # tensors that are NOT parameters but receive a gradient during backward: the zero-padded shadow operands of
        # models whose channel counts are not multiples of 8 (models/multiresunet.py); id -> (offset, shape, buffer)
        self.extern: Dict[int, Tuple[int, torch.Size, torch.Tensor]] = {}

    def add_extern(self, t: torch.Tensor, off: int, flat: torch.Tensor):
        self.extern[id(t)] = (off, t.shape, flat)

    def view(self, p) -> torch.Tensor:
        ext = self.extern.get(id(p))
        if ext is not None:
            return ext[2][ext[0]:ext[0] + p.numel()].view(ext[1])
        off, shape = self.offsets[id(p)]
        return self.flat[off:off + p.numel()].view(shape)

    def ready(self, p):
        """Called by the tape when the gradient of `p` is complete on the compute stream."""
        if id(p) in self.extern:     # the owner gathers these into the parameters' gradients and reports those
            return
        self.done.add(id(p))
        if self.on_ready is not None:
            self.on_ready(p)


class PackPlan:
    """All weight packs of a model's training step in ONE launch (uz_pack_weights_batched).  Learned from the pack calls
    of one complete step (Ctx.pack_log): only packs straight from a parameter's own fp32 storage qualify (no padded or
    scaled temporaries).  `run` re-packs every operand into persistent buffers at the start of a forward and returns
    the cache Ctx.pack consults; it declines (None) when a parameter has moved since the plan was made."""

    def __init__(self, log, params, device):
        own = {p.data_ptr(): p for p in params if p.dtype == torch.float32 and p.is_contiguous()}
        jobs, self.entries, first, total_out = [], [], 0, 0
        seen = set()
        for (w, kind, A, T, B, sa, sb) in log:
            key = (w.data_ptr(), kind)
            p = own.get(w.data_ptr())
            if p is None or key in seen or w.numel() != p.numel():
                continue
            seen.add(key)
            self.entries.append((p, kind, A, T, B, sa, sb, total_out))
            total_out += A * T * B
        self.device = device
        self.store = torch.empty(max(total_out, 1), dtype=torch.bfloat16, device=device)
        rows = []
        for (p, kind, A, T, B, sa, sb, off) in self.entries:
            rows.append([p.data_ptr(), self.store.data_ptr() + 2 * off, A, T, B, sa, sb, 1, T * B, first])
            first += A * B
        self.total_pairs = first
        self.jobs = torch.tensor(rows if rows else [[0] * 10], dtype=torch.int64).to(device)
        self.ptrs = [p.data_ptr() for (p, *_r) in self.entries]

    def run(self, ctx: "Ctx") -> Optional[Dict[tuple, torch.Tensor]]:
        if not self.entries or any(p.data_ptr() != q for (p, *_r), q in zip(self.entries, self.ptrs)):
            return None
        L.check(ctx.lib.uz_pack_weights_batched(self.jobs.data_ptr(), len(self.entries), self.total_pairs, ctx.stream),
                "uz_pack_weights_batched")
        ctx.launches += 1
        return {(p.data_ptr(), kind): self.store[off:off + A * T * B].view(A, T, B)
                for (p, kind, A, T, B, sa, sb, off) in self.entries}


class Ctx:
    """Per-forward execution context: stream, scratch memory, tape, parameter-gradient sink."""

    def __init__(self, device, training: bool, record: bool, precision: str = "bf16"):
        if precision not in ("bf16", "fp32"):
            raise ValueError(f"unet_zoo_b200: precision must be 'bf16' or 'fp32' (got {precision!r})")
        if precision == "fp32" and record:
            raise RuntimeError("unet_zoo_b200: the fp32 parity mode is forward-only (no tape)")
        self.lib = L.load()
        self.device = device
        self.training = training
        self.record = record
        self.fp32 = precision == "fp32"   # operators dispatch to engine_f32 (csrc/f32.cu)
        self.act_dtype = torch.float32 if self.fp32 else torch.bfloat16
        self.tape: List[Callable[["Ctx"], None]] = []
        self.stream = torch.cuda.current_stream(device).cuda_stream
        self.sms = self.lib.uz_device_sm_count()
        if self.sms < 0:
            L.check(self.sms, "device query")
        self.stats_rows = self.lib.uz_conv_stats_rows()
        self.reduce_rows = self.lib.uz_reduce_rows()
        self._ws = None
        # Weight gradients leave the critical path: they run on a side stream while the main stream continues with the
        # data gradient's consumers.  The HBM-bound BatchNorm backward of layer L-1 then shares the SMs with the
        # tensor-bound weight gradient of layer L (192-thread CTAs, 15 K registers: a BatchNorm block still fits).
        self.overlap_wgrad = record and os.environ.get("UZ_OVERLAP_WGRAD", "1") != "0"
        self._ws_main = None
        self._ws_side = None
        self._on_side = False
        self.pgrads: Optional[ParamGrads] = None
        self.input_norm = None                    # (mean, std) when the model is fed decoded uint8 images
        # Inference memo (HotPathModel hands its dict in for eval-mode forwards): packed / BatchNorm-folded weights and
        # running-statistics affines are constants there - re-deriving them cost ~130 small launches (0.6 of 12.4 ms) per
        # resunet forward.  Entries are validated against (storage pointer, version counter) of their source tensors.
        self.eval_cache: Optional[dict] = None
        self.pack_cache: Dict[tuple, torch.Tensor] = {}    # (weight data_ptr, kind) -> operand packed by a PackPlan
        self.pack_log: Optional[list] = None               # not None: record the packs a PackPlan could batch
        self.outputs: List[torch.Tensor] = []     # logits tensors, in the order the model returns them
        self.out_grads: List[Optional[torch.Tensor]] = []
        self.launches = 0

    # ------------------------------------------------------------------ memory
    def new_buf(self, N, H, W, C) -> Buf:
        return Buf(N, H, W, C, self.device, self.act_dtype)

    def new_act(self, N, H, W, C) -> Act:
        return Act(self.new_buf(N, H, W, C))

    def workspace(self, nbytes: int) -> torch.Tensor:
        if self._ws is None or self._ws.numel() < nbytes:   # one scratch buffer per stream (see on_side)
            self._ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=self.device)
        return self._ws

    @contextlib.contextmanager
    def on_side(self):
        """Enqueue the enclosed calls on the side stream, ordered after everything enqueued on the main stream so far.
        Allocations made inside belong to the side stream; the caller joins the streams before the tensors shared with
        the main stream can be released (`join_side`)."""
        main = torch.cuda.current_stream(self.device)
        side = side_stream(self.device)
        side.wait_stream(main)
        saved = (self.stream, self._ws)
        self._ws_main = self._ws
        self.stream, self._ws, self._on_side = side.cuda_stream, self._ws_side, True
        try:
            with torch.cuda.stream(side):
                yield
        finally:
            self._ws_side = self._ws
            self.stream, self._ws = saved
            self._on_side = False
            self.side_used = True

    def join_side(self):
        if getattr(self, "side_used", False):
            torch.cuda.current_stream(self.device).wait_stream(side_stream(self.device))
            self.side_used = False


    def f32(self, *shape) -> torch.Tensor:
        return torch.empty(shape, dtype=torch.float32, device=self.device)

    def cached(self, key, deps: Sequence[Optional[torch.Tensor]], make: Callable):
        """Inference memo: `make()` runs again only when one of `deps` moved or was modified in place."""
        c = self.eval_cache
        if c is None or self.training:
            return make()
        sig = tuple((t.data_ptr(), t._version) for t in deps if t is not None)
        hit = c.get(key)
        if hit is not None and hit[0] == sig:
            return hit[1]
        val = make()
        c[key] = (sig, val)
        return val

    # ------------------------------------------------------------------ kernel wrappers
    def pack(self, w: torch.Tensor, kind: str, row_scale: Optional[torch.Tensor] = None, kp: int = 0) -> torch.Tensor:
        """fp32 parameter -> bf16 GEMM operand [A][T][B] (see uz_pack_weight)."""
        w = w.detach()
        if not w.is_contiguous():
            w = w.contiguous()
        d0, d1 = w.shape[0], w.shape[1]
        T = w.shape[2] * w.shape[3]
        if kind == "conv_fwd":      # OIHW -> [Co][T][Ci]
            A, B, sa, sb = d0, d1, d1 * T, T
        elif kind == "conv_dgrad":  # OIHW -> [Ci][T][Co]
            A, B, sa, sb = d1, d0, T, d1 * T
        elif kind == "convt_fwd":   # IOHW -> [Co][4][Ci]
            A, B, sa, sb = d1, d0, T, d1 * T
        elif kind == "convt_dgrad":  # IOHW -> [Ci][4][Co]
            A, B, sa, sb = d0, d1, d1 * T, T
        elif kind == "stem":        # OIHW -> [Co][kp] with k = t*Ci + ci
            A, B, sa, sb = d0, d1, d1 * T, T
        else:
            raise ValueError(kind)
        pitch = T * B
        batchable = kind != "stem" and row_scale is None and w.dtype == torch.float32
        if batchable:
            hit = self.pack_cache.get((w.data_ptr(), kind))
            if hit is not None and hit.shape == (A, T, B):
                return hit
            if self.pack_log is not None:
                self.pack_log.append((w, kind, A, T, B, sa, sb))
        if kind == "stem":
            pitch = kp
            out = torch.zeros((A, kp), dtype=torch.bfloat16, device=w.device)
        else:
            out = torch.empty((A, T, B), dtype=torch.bfloat16, device=w.device)
        rc = self.lib.uz_pack_weight(w.data_ptr(), out.data_ptr(), A, T, B, sa, sb, 1, pitch, _ptr(row_scale), self.stream)
        L.check(rc, "uz_pack_weight")
        self.launches += 1
        return out

    def pack_up2x(self, w: torch.Tensor, transpose: bool) -> torch.Tensor:
        """Pre-summed phase weights of the nearest-x2 conv (uz_pack_up2x): [Co][16][Ci], or [Ci][16][Co] for the data
        gradient."""
        w = w.detach().contiguous()
        co, ci = w.shape[0], w.shape[1]
        out = torch.empty((ci, 16, co) if transpose else (co, 16, ci), dtype=torch.bfloat16, device=w.device)
        L.check(self.lib.uz_pack_up2x(w.data_ptr(), out.data_ptr(), co, ci, int(transpose), self.stream), "uz_pack_up2x")
        self.launches += 1
        return out

    def conv(self, mode: int, segs: Sequence[Act], wp: torch.Tensor, cout: int, bias: Optional[torch.Tensor],
             out_ptr: int, out_ld: int, out_h: int, out_w: int, stats: Optional[torch.Tensor] = None, flags: int = 0,
             in_grad: bool = False, dilation: int = 1):
        """uz_conv_igemm over a virtual concat of `segs` (activations, or their gradients when in_grad)."""
        n = len(segs)
        ptrs = (_c_void_p * n)(*[(s.gptr() if in_grad else s.ptr) for s in segs])
        cs = (ctypes.c_int * n)(*[s.C for s in segs])
        lds = (ctypes.c_int * n)(*[s.ld for s in segs])
        a = segs[0]
        # algorithmic FLOPs: 2 * output pixels * cout * cin * taps (per output pixel; convT has one tap per output)
        opix = a.N * out_h * out_w
        taps = {L.CONVT2X2: 1, L.CONV3X3_S2_DGRAD: 2.25}.get(mode, _TAPS[mode])
        if wp.dim() == 3 and wp.shape[1] == 16:   # nearest-x2 conv in phase form: EXECUTED FLOPs (4 of 9 / 16 of 36 taps)
            taps = 4 if mode == L.UP2X_CONV3X3 else 16
        cin = sum(s.C for s in segs)
        flops = 2.0 * opix * cout * cin * taps
        if flags & (L.FLAG_RES1X1 | L.FLAG_RES1X1_S2):    # last segment: 1x1 skip operand (one tap)
            flops = 2.0 * opix * cout * ((cin - segs[-1].C) * taps + segs[-1].C)
        # algorithmic HBM bytes: every operand once (input pixels, packed weights, output (read too when accumulating))
        nbytes = 2.0 * (a.npix * cin + opix * cout * (2 if flags & L.FLAG_ACCUMULATE else 1) + wp.numel())
        tag = f"{_MODE_NAMES.get(mode, mode)}{'' if dilation == 1 else f'.d{dilation}'} {cin}->{cout} @{out_h}x{out_w} n{a.N}"
        with _Timed("igemm_fwd", flops, tag, nbytes):
            rc = self.lib.uz_conv_igemm(mode | L.dilation(dilation), a.N, a.H, a.W, n, ptrs, cs, lds, wp.data_ptr(), cout, _ptr(bias), out_ptr,
                                        out_ld, out_h, out_w, _ptr(stats), flags, self.stream)
        L.check(rc, "uz_conv_igemm")
        self.launches += 1

    def wgrad(self, mode: int, x_ptr: int, N, H, W, cin: int, x_ld: int, dy_ptr: int, cout: int, dy_ld: int,
              dw: torch.Tensor, cin_total: int, c_off: int, accumulate: bool = False, dilation: int = 1):
        lmode = mode | L.dilation(dilation)
        need = self.lib.uz_conv_wgrad_workspace_bytes(lmode, N, H, W, cin, cout)
        ws = self.workspace(need)
        if mode == L.CONVT2X2:
            flops = 2.0 * N * H * W * cin * cout * 4
        elif mode in (L.CONV3X3_S2, L.CONV1X1_S2):
            flops = 2.0 * N * ((H - 1) // 2 + 1) * ((W - 1) // 2 + 1) * cin * cout * _TAPS[mode]
        elif mode == L.UP2X_CONV3X3:   # executed FLOPs: 16 tap-pixels per low-resolution pixel in phase form, else 36
            flops = 2.0 * N * H * W * cin * cout * (16 if self.lib.uz_up2x_phase_form(H, W, cout) else 36)
        else:
            flops = 2.0 * N * H * W * cin * cout * _TAPS[mode]
        if mode in (L.CONVT2X2, L.UP2X_CONV3X3):
            dy_pix = 4 * N * H * W
        elif mode in (L.CONV3X3_S2, L.CONV1X1_S2):
            dy_pix = N * ((H - 1) // 2 + 1) * ((W - 1) // 2 + 1)
        else:
            dy_pix = N * H * W
        # algorithmic HBM bytes: x and dy once, the fp32 gradient block once (read too when accumulating)
        nbytes = 2.0 * (N * H * W * cin + dy_pix * cout) + 4.0 * cin * cout * (4 if mode == L.CONVT2X2 else _TAPS[mode]) * (2 if accumulate else 1)
        with _Timed("igemm_wgrad", flops, f"{_MODE_NAMES.get(mode, mode)}.wgrad {cin}->{cout} @{H}x{W} n{N}", nbytes):
            rc = self.lib.uz_conv_wgrad(lmode, N, H, W, x_ptr, cin, x_ld, dy_ptr, cout, dy_ld, dw.data_ptr(), cin_total,
                                        c_off, int(accumulate), ws.data_ptr(), ws.numel(), self.stream)
        L.check(rc, "uz_conv_wgrad")
        self.launches += 2

    def bn_finalize(self, stats: torch.Tensor, rows: int, C: int, count: int, bn_mod, bnbuf: torch.Tensor):
        track = bn_mod.track_running_stats and bn_mod.running_mean is not None
        mom = bn_mod.momentum
        nbt = None
        if track:
            t = bn_mod.num_batches_tracked
            if mom is None or t is None or not t.is_cuda or t.dtype != torch.int64:
                bn_mod.num_batches_tracked += 1
                if mom is None:  # cumulative moving average, as nn.BatchNorm2d does for momentum=None
                    mom = 1.0 / float(bn_mod.num_batches_tracked)
            else:
                nbt = t.data_ptr()       # the counter is bumped inside the finalize kernel: no launch of its own
        rc = self.lib.uz_bn_finalize(stats.data_ptr(), rows, C, count, _ptr(bn_mod.weight), _ptr(bn_mod.bias),
                                     bn_mod.eps, mom if mom is not None else 0.0,
                                     _ptr(bn_mod.running_mean) if track else None,
                                     _ptr(bn_mod.running_var) if track else None, bnbuf.data_ptr(), nbt, self.stream)
        L.check(rc, "uz_bn_finalize")
        self.launches += 1

    def bn_apply(self, z: Act, scale_ptr, shift_ptr, relu: bool, y: Optional[Act], pooled: Optional[Act] = None,
                 residual: Optional[Act] = None):
        nbytes = 2.0 * z.npix * z.C * (1 + (1 if y else 0) + (0.25 if pooled else 0) + (1 if residual else 0))
        with _Timed("bn_apply", nbytes):
            rc = self.lib.uz_bn_apply(z.ptr, z.ld, z.N, z.H, z.W, z.C, scale_ptr, shift_ptr, int(relu),
                                      residual.ptr if residual else None, residual.ld if residual else 0,
                                      y.ptr if y else None, y.ld if y else 0,
                                      pooled.ptr if pooled else None, pooled.ld if pooled else 0, self.stream)
        L.check(rc, "uz_bn_apply")
        self.launches += 1

    def bn_bwd(self, z: Act, dy_ptr, dy_ld, dpool_ptr, dpool_ld, bnbuf: torch.Tensor, relu: bool, dz_ptr: int, dz_ld: int,
               dgamma, dbeta, dbias, dz_accumulate: bool = False):
        need = self.lib.uz_bn_bwd_workspace_bytes(z.C)
        ws = self.workspace(need)
        # algorithmic bytes: reduce pass reads dy, z (+ dpool/4); apply pass reads them again and writes dz
        per = (1 if dy_ptr else 0) + 1 + (0.25 if dpool_ptr else 0)
        nbytes = 2.0 * z.npix * z.C * (2 * per + 1 + (1 if dz_accumulate else 0))
        with _Timed("bn_bwd", nbytes):
            rc = self.lib.uz_bn_bwd(z.ptr, z.ld, dy_ptr, dy_ld, dpool_ptr, dpool_ld, z.N, z.H, z.W, z.C, bnbuf.data_ptr(),
                                    int(relu), dz_ptr, dz_ld, _ptr(dgamma), _ptr(dbeta), _ptr(dbias),
                                    2 if dz_accumulate else 0, ws.data_ptr(), ws.numel(), self.stream)
        L.check(rc, "uz_bn_bwd")
        self.launches += 4 if dbias is not None else 3

    def colsum(self, ptr: int, ld: int, npix: int, C: int, out: torch.Tensor, accumulate=False):
        need = self.lib.uz_colsum_workspace_bytes(C)
        ws = self.workspace(need)
        rc = self.lib.uz_colsum(ptr, ld, npix, C, out.data_ptr(), int(accumulate), ws.data_ptr(), ws.numel(), self.stream)
        L.check(rc, "uz_colsum")
        self.launches += 2

    # ------------------------------------------------------------------ parameter gradients
    def pgrad(self, p: torch.nn.Parameter) -> Optional[torch.Tensor]:
        if p is None or not p.requires_grad:
            return None
        return self.pgrads.view(p)

    def pready(self, *ps):
        for p in ps:
            if p is not None and p.requires_grad:
                self.pgrads.ready(p)


# ====================================================================================================== operators
def _fold_bn(conv_bias: Optional[torch.Tensor], bn) -> Tuple[torch.Tensor, torch.Tensor]:
    """Inference-mode BatchNorm folded into (row_scale, bias) of the preceding convolution."""
    invstd = torch.rsqrt(bn.running_var + bn.eps)
    g = bn.weight if bn.weight is not None else torch.ones_like(invstd)
    b = bn.bias if bn.bias is not None else torch.zeros_like(invstd)
    scale = (g * invstd).float().contiguous()
    cb = conv_bias if conv_bias is not None else torch.zeros_like(invstd)
    bias = ((cb - bn.running_mean) * scale + b).float().contiguous()
    return scale, bias


def stem_input(ctx: Ctx, x: torch.Tensor) -> Tuple[Act, str]:
    """NCHW fp32 images -> the operand of the first convolution.

    in_channels <= 7: 3x3 im2col into one 64-wide K block (the stem conv then runs as a 1x1 GEMM);
    otherwise a plain NHWC bf16 copy with channels padded to a multiple of 8.
    """
    if ctx.fp32:
        return _F32.stem_input(ctx, x)
    N, cin, H, W = x.shape
    if 9 * cin <= 64:
        kp = 16 if 9 * cin <= 16 else 32 if 9 * cin <= 32 else 64
        a = ctx.new_act(N, H, W, kp)
        if x.dtype == torch.uint8:
            # decoded bytes (NHWC uint8 behind an NCHW-shaped view): ToTensor + Normalize fused into the im2col pass
            mean, std = ctx.input_norm
            fm, fs = (ctypes.c_float * cin)(*mean), (ctypes.c_float * cin)(*std)
            L.check(ctx.lib.uz_stem_im2col_u8(x.data_ptr(), N, cin, H, W, fm, fs, a.ptr, kp, ctx.stream),
                    "uz_stem_im2col_u8")
            ctx.launches += 1
            return a, "im2col"
        L.check(ctx.lib.uz_stem_im2col(x.data_ptr(), N, cin, H, W, a.ptr, kp, ctx.stream), "uz_stem_im2col")
        ctx.launches += 1
        return a, "im2col"
    cpad = (cin + 7) // 8 * 8
    a = ctx.new_act(N, H, W, cpad)
    L.check(ctx.lib.uz_nchw_to_nhwc(x.data_ptr(), N, cin, H, W, a.ptr, cpad, ctx.stream), "uz_nchw_to_nhwc")
    ctx.launches += 1
    return a, "nhwc"


def _weight_for(conv, cin_total: int) -> torch.Tensor:
    """The convolution weight as the GEMM sees it: input channels zero-padded up to the operand's channel count (only
    the NHWC stem of models with in_channels > 7, whose image is padded to a multiple of 8 channels)."""
    w = conv.weight.detach()
    if cin_total == w.shape[1]:
        return w
    return torch.nn.functional.pad(w, (0, 0, 0, 0, 0, cin_total - w.shape[1]))


def conv_bn_act(ctx: Ctx, xs: Sequence[Act], conv, bn, out: Optional[Act] = None, pooled: Optional[Act] = None,
                relu: bool = True, mode: int = L.CONV3X3, stem: Optional[str] = None, need_dx: bool = True,
                dilation: int = 1) -> Act:
    """[ReLU](BatchNorm2d(Conv2d(cat(xs)))) with optional MaxPool2d(2,2) second output.
    `dilation` > 1 (mode CONV3X3 only): Conv2d(3, padding = dilation = d), REBNCONV of u2net.py:10.

    Reference: DoubleConv / ConvBlock / VGGBlock halves (common_layers.py:28-33,47-57; nested_unet.py:13-22),
    UpConvBlock with mode=UP2X_CONV3X3 (common_layers.py:70-76), DownSample pooling (common_layers.py:92-95).
    Training: conv epilogue emits raw bf16 z + per-channel sum/sum-sq partials -> finalize -> one apply pass.
    Inference: BatchNorm folded into the packed weights and bias, ReLU in the conv epilogue.
    """
    if ctx.fp32:
        if dilation != 1:
            raise RuntimeError("unet_zoo_b200: dilated convolutions have no fp32 parity path")
        return _F32.conv_bn_act(ctx, xs, conv, bn, out=out, pooled=pooled, relu=relu, mode=mode, stem=stem, need_dx=need_dx)
    if dilation != 1 and (mode != L.CONV3X3 or stem == "im2col"):
        raise ValueError("dilation needs a plain 3x3 convolution")
    a0 = xs[0]
    N, H, W = a0.N, a0.H, a0.W
    cout = conv.out_channels
    oscale = 2 if mode == L.UP2X_CONV3X3 else 1
    if mode in (L.CONV3X3_S2, L.CONV1X1_S2):
        Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    else:
        Ho, Wo = H * oscale, W * oscale
    if out is None:
        out = ctx.new_act(N, Ho, Wo, cout)
    is_stem = stem == "im2col"
    conv_mode = L.CONV1X1 if is_stem else mode
    cin_total = sum(a.C for a in xs)
    # nearest-x2 conv: phase form (2x2 windows of the low-resolution input, pre-summed weights) where the kernel takes it
    phase_form = mode == L.UP2X_CONV3X3 and len(xs) == 1 and bool(ctx.lib.uz_up2x_phase_form(H, W, cout))

    if not ctx.training:
        def folded():
            scale, bias = _fold_bn(conv.bias, bn) if bn is not None else (None, conv.bias)
            if phase_form:      # fold the BatchNorm scale into the fp32 weights, then pre-sum per phase
                w = conv.weight.detach() * scale.view(-1, 1, 1, 1) if scale is not None else conv.weight
                return ctx.pack_up2x(w, False), bias
            return ctx.pack(conv.weight if is_stem else _weight_for(conv, cin_total), "stem" if is_stem else "conv_fwd",
                            row_scale=scale, kp=a0.C if is_stem else 0), bias
        deps = [conv.weight, conv.bias] + ([bn.weight, bn.bias, bn.running_mean, bn.running_var] if bn is not None else [])
        wp, bias = ctx.cached(("conv_bn", id(conv), id(bn), cin_total, a0.C if is_stem else 0), deps, folded)
        ctx.conv(conv_mode, xs, wp, cout, bias, out.ptr, out.ld, Ho, Wo, None, L.FLAG_RELU if relu else 0, dilation=dilation)
        if pooled is not None:
            ctx.bn_apply(out, None, None, False, None, pooled)
        return out

    wp = ctx.pack_up2x(conv.weight, False) if phase_form else \
        ctx.pack(conv.weight if is_stem else _weight_for(conv, cin_total), "stem" if is_stem else "conv_fwd",
                 kp=a0.C if is_stem else 0)
    z = ctx.new_act(N, Ho, Wo, cout)
    stats = ctx.f32(ctx.stats_rows, 2, cout)
    ctx.conv(conv_mode, xs, wp, cout, conv.bias, z.ptr, z.ld, Ho, Wo, stats, L.FLAG_STATS, dilation=dilation)
    bnbuf = ctx.f32(4, cout)
    ctx.bn_finalize(stats, ctx.stats_rows, cout, N * Ho * Wo, bn, bnbuf)
    ctx.bn_apply(z, bnbuf[2].data_ptr(), bnbuf[3].data_ptr(), relu, out, pooled)

    if ctx.record:
        def backward(c: Ctx):
            has_dy = out.has_grad()
            has_dp = pooled is not None and pooled.has_grad()
            if not (has_dy or has_dp):
                raise RuntimeError("unet_zoo_b200: conv_bn_act output has no gradient (unused output?)")
            dgamma, dbeta = c.pgrad(bn.weight), c.pgrad(bn.bias)
            dbias = c.pgrad(conv.bias) if conv.bias is not None else None
            # dz overwrites z in place (same thread reads z[p] then writes dz[p])
            c.bn_bwd(z, out.gptr() if has_dy else None, out.ld, pooled.gptr() if has_dp else None,
                     pooled.ld if has_dp else 0, bnbuf, relu, z.ptr, z.ld, dgamma, dbeta, dbias)
            c.pready(bn.weight, bn.bias, conv.bias)
            _conv_backward(c, xs, conv, z, mode, is_stem, cin_total, need_dx, dilation)
        ctx.tape.append(backward)
    return out


def _conv_backward(c: Ctx, xs: Sequence[Act], conv, dz: Act, mode: int, is_stem: bool, cin_total: int, need_dx: bool,
                   dilation: int = 1):
    """Weight gradient (tcgen05 split-K) and data gradient (flipped-weight implicit GEMM) of one convolution;
    `dz` holds the gradient w.r.t. the convolution output (NHWC bf16)."""
    a0 = xs[0]
    cout = conv.out_channels
    dw = c.pgrad(conv.weight)

    def weight_gradient():
        if is_stem:
            kp = a0.C
            cin = conv.in_channels
            tmp = c.f32(cout, kp)
            c.wgrad(L.CONV1X1, a0.ptr, a0.N, a0.H, a0.W, kp, a0.ld, dz.ptr, cout, dz.ld, tmp, kp, 0)
            dw.copy_(tmp[:, :9 * cin].view(cout, 9, cin).permute(0, 2, 1).reshape(dw.shape))
        else:
            padded = cin_total != conv.in_channels      # NHWC stem with zero-padded input channels
            tgt = c.f32(cout, cin_total, *conv.weight.shape[2:]) if padded else dw
            off = 0
            for a in xs:
                if mode == L.UP2X_CONV3X3 and len(xs) == 1 and bool(c.lib.uz_up2x_phase_form(a.H, a.W, cout)) \
                        and os.environ.get("UZ_UP2X_WGRAD_MATERIALISE") is None:
                    # phase form on the halo-box kernel: the low-resolution operand against the four output phases of
                    # dZ (16 tap-pixels per input pixel instead of 36, no upsampled tensor)
                    c.wgrad(mode, a.ptr, a.N, a.H, a.W, a.C, a.ld, dz.ptr, cout, dz.ld, tgt, cin_total, off)
                elif mode == L.UP2X_CONV3X3 and os.environ.get("UZ_UP2X_WGRAD_GEN1") is None:
                    # materialise the nearest-x2 operand (one bandwidth pass) so that the weight gradient runs on the
                    # halo-box 3x3 kernel instead of the first-generation per-tap kernel (measured 2.5x faster)
                    up = c.new_act(a.N, 2 * a.H, 2 * a.W, a.C)
                    L.check(c.lib.uz_nearest2x_fwd(a.ptr, a.ld, a.N, a.H, a.W, a.C, up.ptr, up.ld, c.stream),
                            "uz_nearest2x_fwd")
                    c.launches += 1
                    c.wgrad(L.CONV3X3, up.ptr, up.N, up.H, up.W, up.C, up.ld, dz.ptr, cout, dz.ld, tgt, cin_total, off)
                else:
                    c.wgrad(mode, a.ptr, a.N, a.H, a.W, a.C, a.ld, dz.ptr, cout, dz.ld, tgt, cin_total, off,
                            dilation=dilation)
                off += a.C
            if padded:
                dw.copy_(tgt[:, :conv.in_channels])
        c.pready(conv.weight)

    has_dgrad = need_dx and not is_stem
    if dw is not None and not (has_dgrad and c.overlap_wgrad):
        weight_gradient()
    if has_dgrad:
        dgrad_mode = {L.CONV3X3: L.CONV3X3_DGRAD, L.CONV1X1: L.CONV1X1, L.CONV3X3_S2: L.CONV3X3_S2_DGRAD,
                      L.CONV1X1_S2: L.CONV1X1_S2_DGRAD, L.UP2X_CONV3X3: L.UP2X_CONV3X3_DGRAD}.get(mode)
        if dgrad_mode is None:
            raise NotImplementedError(f"data gradient for conv mode {mode}")
        up2x_phase = mode == L.UP2X_CONV3X3 and len(xs) == 1 and bool(c.lib.uz_up2x_phase_form(a0.H, a0.W, cout))
        wpd = c.pack_up2x(conv.weight, True) if up2x_phase else c.pack(conv.weight, "conv_dgrad")  # [Cin][T][Cout]
        off = 0
        for a in xs:
            for (r0, rc, acc) in a.grad_writes():
                if mode == L.CONV1X1_S2 and not acc:
                    # the stride-2 1x1 gradient only touches even pixels: clear the slice, then accumulate
                    a.buf.grad_buf()[..., a.c0 + r0:a.c0 + r0 + rc].zero_()
                    acc = True
                w_rows = wpd[off + r0: off + r0 + rc]
                stats = None
                if a.buf.want_gsum and not acc and mode == L.CONV3X3 and dilation == 1:
                    # the epilogue's per-channel sums of this gradient ARE the bias gradient of the ConvTranspose2d that
                    # produced (part of) the buffer: saves a full read of the gradient (uz_colsum)
                    stats = c.f32(c.stats_rows, 2, rc)
                    a.buf.gsum.append((a.c0 + r0, a.c0 + r0 + rc, stats))
                c.conv(dgrad_mode, [dz], w_rows, rc, None, a.gptr() + 2 * r0, a.ld, a.H, a.W, stats,
                       L.FLAG_STATS if stats is not None else (L.FLAG_ACCUMULATE if acc else 0), dilation=dilation)
            off += a.C
    if dw is not None and has_dgrad and c.overlap_wgrad:
        # after the data gradient in stream order: the data gradient is on the critical path, the weight gradient only
        # has to be finished by the end of backward (or by its gradient bucket's all-reduce)
        with c.on_side():
            weight_gradient()


def pad_into(ctx: Ctx, src: Act, dst: Act) -> Act:
    """F.pad(src, [0, dx, 0, dy]) written into `dst` (a channel slice of the concat buffer) - the reference's centre
    padding of the upsampled tensor when the skip tensor has an odd size (common_layers.py:110-113, :173-178: the pad is
    [d // 2, d - d // 2] = [0, 1] per axis).  Backward = crop of the gradient."""
    if src.N != dst.N or src.C != dst.C or not (0 <= dst.H - src.H <= 1 and 0 <= dst.W - src.W <= 1):
        raise RuntimeError(f"unet_zoo_b200: cannot pad {(src.N, src.C, src.H, src.W)} into {(dst.N, dst.C, dst.H, dst.W)}")
    es = src.buf.esize
    L.check(ctx.lib.uz_pad_copy(src.ptr, src.ld * es, src.H, src.W, dst.ptr, dst.ld * es, dst.H, dst.W, src.N,
                                src.C * es, ctx.stream), "uz_pad_copy")
    ctx.launches += 1
    if ctx.record:
        def backward(c: Ctx):
            if not dst.has_grad():
                raise RuntimeError("unet_zoo_b200: padded tensor has no gradient")
            runs = src.grad_writes()
            if any(acc for _, _, acc in runs):
                raise NotImplementedError("pad_into: source already has a gradient")
            L.check(c.lib.uz_pad_copy(dst.gptr(), dst.ld * 2, dst.H, dst.W, src.gptr(), src.ld * 2, src.H, src.W, src.N,
                                      src.C * 2, c.stream), "uz_pad_copy")
            c.launches += 1
        ctx.tape.append(backward)
    return dst


def conv_transpose2x2(ctx: Ctx, x: Act, up, out: Act) -> Act:
    """ConvTranspose2d(k=2, s=2) written straight into a channel slice of the concat buffer.

    Reference: UpSample_UNet.up (common_layers.py:104,108) / UpsampleResUnet (common_layers.py:201-207).
    Four 1x1 GEMMs (one per output phase) with a pixel-shuffle scatter epilogue.
    """
    if (out.H, out.W) != (2 * x.H, 2 * x.W):
        # odd skip size: the reference pads the upsampled tensor (common_layers.py:110-113) - one extra copy on this
        # (rare) path instead of a pitched scatter epilogue
        return pad_into(ctx, conv_transpose2x2(ctx, x, up, ctx.new_act(x.N, 2 * x.H, 2 * x.W, up.out_channels)), out)
    if ctx.fp32:
        return _F32.conv_transpose2x2(ctx, x, up, out)
    cout = up.out_channels
    wp = ctx.cached(("convt", id(up)), [up.weight], lambda: ctx.pack(up.weight, "convt_fwd"))
    ctx.conv(L.CONVT2X2, [x], wp, cout, up.bias, out.ptr, out.ld, 2 * x.H, 2 * x.W)
    if ctx.record:
        want_db = up.bias is not None and up.bias.requires_grad
        out.buf.want_gsum = out.buf.want_gsum or want_db

        def backward(c: Ctx):
            dw = c.pgrad(up.weight)

            def weight_gradient():
                c.wgrad(L.CONVT2X2, x.ptr, x.N, x.H, x.W, x.C, x.ld, out.gptr(), cout, out.ld, dw, x.C, 0)
                c.pready(up.weight)

            if dw is not None and not c.overlap_wgrad:
                weight_gradient()
            if want_db:
                lo, hi = out.c0, out.c0 + cout
                hit = next((e for e in out.buf.gsum if e[0] <= lo and hi <= e[1]), None)
                if hit is not None:
                    torch.sum(hit[2][:, 0, lo - hit[0]:hi - hit[0]], dim=0, out=c.pgrad(up.bias))
                else:
                    c.colsum(out.gptr(), out.ld, out.npix, cout, c.pgrad(up.bias))
            c.pready(up.bias)
            wpd = c.pack(up.weight, "convt_dgrad")  # [Cin][4][Cout]
            for (r0, rc, acc) in x.grad_writes():
                c.conv(L.CONVT2X2_DGRAD, [out], wpd[r0:r0 + rc], rc, None, x.gptr() + 2 * r0, x.ld, x.H, x.W, None,
                       L.FLAG_ACCUMULATE if acc else 0, in_grad=True)
            if dw is not None and c.overlap_wgrad:
                with c.on_side():
                    weight_gradient()
        ctx.tape.append(backward)
    return out


def head1x1(ctx: Ctx, y: Act, conv) -> torch.Tensor:
    """1x1 classifier head -> fp32 NCHW logits (OutConv, common_layers.py:125-128)."""
    if ctx.fp32:
        return _F32.head1x1(ctx, y, conv)
    K = conv.out_channels
    logits = torch.empty((y.N, K, y.H, y.W), dtype=torch.float32, device=ctx.device)
    w = conv.weight.detach().reshape(K, -1)
    rc = ctx.lib.uz_head_fwd(y.ptr, y.ld, y.N, y.H, y.W, y.C, K, w.data_ptr(), _ptr(conv.bias), logits.data_ptr(),
                             ctx.stream)
    L.check(rc, "uz_head_fwd")
    ctx.launches += (K + 3) // 4
    idx = len(ctx.outputs)
    ctx.outputs.append(logits)
    if ctx.record:
        def backward(c: Ctx):
            g = c.out_grads[idx] if idx < len(c.out_grads) else None
            if g is not None:
                head1x1_backward(c, y, conv, g)
        ctx.tape.append(backward)
    return logits


def head1x1_backward(c: Ctx, y: Act, conv, dlogits: torch.Tensor):
    K = conv.out_channels
    dl = dlogits.contiguous().float()
    w = conv.weight.detach().reshape(K, -1)
    dw = c.pgrad(conv.weight)
    db = c.pgrad(conv.bias) if conv.bias is not None else None
    if dw is None:
        dw = c.f32(K, y.C)
    if db is None:
        db = c.f32(K)
    need = c.lib.uz_head_bwd_workspace_bytes(y.C)
    ws = c.workspace(need)
    runs = y.grad_writes()
    if any(acc for _, _, acc in runs):
        raise NotImplementedError("classifier head input already has a gradient")
    rc = c.lib.uz_head_bwd(y.ptr, y.ld, dl.data_ptr(), y.N, y.H, y.W, y.C, K, w.data_ptr(), y.gptr(), y.ld,
                           dw.data_ptr(), db.data_ptr(), 0, ws.data_ptr(), ws.numel(), c.stream)
    L.check(rc, "uz_head_bwd")
    c.launches += 3 * ((K + 3) // 4)
    c.pready(conv.weight, conv.bias)


def conv_bn_act_head(ctx: Ctx, xs: Sequence[Act], conv, bn, head, relu: bool = True) -> torch.Tensor:
    """head(ReLU(BatchNorm2d(Conv2d(cat(xs))))) where the activation is consumed by the 1x1 head ONLY: the last
    DoubleConv / VGGBlock half followed by OutConv (common_layers.py:28-33,125-128; attention_unet.py:70,108;
    nested_unet.py:69,104).  Training: the full-resolution activation and its gradient are never stored
    (uz_bn_head_fwd / uz_bn_head_bwd).  Inference, wide heads or odd widths fall back to conv_bn_act + head1x1."""
    if ctx.fp32:
        return _F32.conv_bn_act_head(ctx, xs, conv, bn, head, relu=relu)
    cout, K = conv.out_channels, head.out_channels
    if not ctx.training or K > 4 or cout < 8 or cout > 256 or (cout & (cout - 1)):
        return head1x1(ctx, conv_bn_act(ctx, xs, conv, bn, relu=relu), head)
    a0 = xs[0]
    N, H, W = a0.N, a0.H, a0.W
    wp = ctx.pack(conv.weight, "conv_fwd")
    z = ctx.new_act(N, H, W, cout)
    stats = ctx.f32(ctx.stats_rows, 2, cout)
    ctx.conv(L.CONV3X3, xs, wp, cout, conv.bias, z.ptr, z.ld, H, W, stats, L.FLAG_STATS)
    bnbuf = ctx.f32(4, cout)
    ctx.bn_finalize(stats, ctx.stats_rows, cout, N * H * W, bn, bnbuf)
    logits = torch.empty((N, K, H, W), dtype=torch.float32, device=ctx.device)
    hw_ = head.weight.detach().reshape(K, -1)
    with _Timed("bn_apply", 2.0 * z.npix * cout):
        rc = ctx.lib.uz_bn_head_fwd(z.ptr, z.ld, N, H, W, cout, bnbuf[2].data_ptr(), bnbuf[3].data_ptr(), int(relu), K,
                                    hw_.data_ptr(), _ptr(head.bias), logits.data_ptr(), ctx.stream)
    L.check(rc, "uz_bn_head_fwd")
    ctx.launches += 1
    idx = len(ctx.outputs)
    ctx.outputs.append(logits)
    if ctx.record:
        def backward(c: Ctx):
            g = c.out_grads[idx] if idx < len(c.out_grads) else None
            if g is None:
                raise RuntimeError("unet_zoo_b200: the classifier head received no gradient")
            dl = g.contiguous().float()
            dw, db = c.pgrad(head.weight), c.pgrad(head.bias) if head.bias is not None else None
            ws = c.workspace(c.lib.uz_bn_head_workspace_bytes(cout, K))
            with _Timed("bn_bwd", 6.0 * z.npix * cout):
                rc = c.lib.uz_bn_head_bwd(z.ptr, z.ld, N, H, W, cout, bnbuf.data_ptr(), int(relu), K, hw_.data_ptr(),
                                          dl.data_ptr(), z.ptr, z.ld, _ptr(c.pgrad(bn.weight)), _ptr(c.pgrad(bn.bias)),
                                          _ptr(c.pgrad(conv.bias) if conv.bias is not None else None), _ptr(dw), _ptr(db),
                                          0, ws.data_ptr(), ws.numel(), c.stream)
            L.check(rc, "uz_bn_head_bwd")
            c.launches += 5
            c.pready(head.weight, head.bias, bn.weight, bn.bias, conv.bias)
            _conv_backward(c, xs, conv, z, L.CONV3X3, False, sum(a.C for a in xs), True)
        ctx.tape.append(backward)
    return logits


def _running_affine(bn, device) -> torch.Tensor:
    """[4][C] = {mean, invstd, scale, shift} of an inference-mode BatchNorm (running statistics)."""
    invstd = torch.rsqrt(bn.running_var.float() + bn.eps)
    g = bn.weight.float() if bn.weight is not None else torch.ones_like(invstd)
    b = bn.bias.float() if bn.bias is not None else torch.zeros_like(invstd)
    scale = g * invstd
    return torch.stack([bn.running_mean.float(), invstd, scale, b - bn.running_mean.float() * scale]).contiguous()


def conv_stats(ctx: Ctx, xs: Sequence[Act], conv, bn, mode: int = L.CONV3X3) -> Tuple[Act, torch.Tensor]:
    """Training-mode Conv2d + BatchNorm statistics without the apply pass: returns (raw bf16 z, bn[4][C])."""
    a0 = xs[0]
    cout = conv.out_channels
    z = ctx.new_act(a0.N, a0.H, a0.W, cout)
    wp = ctx.pack(conv.weight, "conv_fwd")
    stats = ctx.f32(ctx.stats_rows, 2, cout)
    ctx.conv(mode, xs, wp, cout, conv.bias, z.ptr, z.ld, a0.H, a0.W, stats, L.FLAG_STATS)
    bnbuf = ctx.f32(4, cout)
    ctx.bn_finalize(stats, ctx.stats_rows, cout, a0.npix, bn, bnbuf)
    return z, bnbuf


def bn_act(ctx: Ctx, x: Act, bn, relu: bool = True, out: Optional[Act] = None, pooled: Optional[Act] = None,
           stats: Optional[torch.Tensor] = None) -> Act:
    """[ReLU](BatchNorm2d(x)) on an existing activation - the pre-activation half of ResidualConv
    (common_layers.py:185-186,189-190) and the stand-alone BatchNorm2d(affine=False) layers of MultiResUnet
    (multiresunet.py:79,84,127,134); `pooled` receives MaxPool2d(2, 2) of the result (multiresunet.py:206).
    Training: one statistics pass + one apply pass; inference: apply only."""
    if ctx.fp32:
        if out is not None or pooled is not None:
            raise RuntimeError("unet_zoo_b200: bn_act into a given buffer / with pooling has no fp32 parity path")
        return _F32.bn_act(ctx, x, bn, relu=relu)
    if out is None:
        out = ctx.new_act(x.N, x.H, x.W, x.C)
    if not ctx.training:
        aff = ctx.cached(("affine", id(bn)), [bn.weight, bn.bias, bn.running_mean, bn.running_var],
                         lambda: _running_affine(bn, ctx.device))
        ctx.bn_apply(x, aff[2].data_ptr(), aff[3].data_ptr(), relu, out, pooled)
        return out
    rows = ctx.reduce_rows
    if stats is None:       # else: the producer of x already left its partial sums (dual_bn_add_relu)
        stats = ctx.f32(rows, 2, x.C)
        L.check(ctx.lib.uz_bn_stats(x.ptr, x.ld, x.npix, x.C, stats.data_ptr(), ctx.stream), "uz_bn_stats")
        ctx.launches += 1
    bnbuf = ctx.f32(4, x.C)
    ctx.bn_finalize(stats, rows, x.C, x.npix, bn, bnbuf)
    ctx.bn_apply(x, bnbuf[2].data_ptr(), bnbuf[3].data_ptr(), relu, out, pooled)
    if ctx.record:
        def backward(c: Ctx):
            has_dy = out.has_grad()
            has_dp = pooled is not None and pooled.has_grad()
            if not (has_dy or has_dp):
                raise RuntimeError("unet_zoo_b200: bn_act output has no gradient")
            for (r0, rc, acc) in x.grad_writes():
                xs_ = x.slice(r0, rc)
                sub = bnbuf if rc == x.C else torch.stack([bnbuf[k, r0:r0 + rc] for k in range(4)]).contiguous()
                dg, db = c.pgrad(bn.weight), c.pgrad(bn.bias)
                c.bn_bwd(xs_, out.gptr() + 2 * r0 if has_dy else None, out.ld, pooled.gptr() + 2 * r0 if has_dp else None,
                         pooled.ld if has_dp else 0, sub, relu, xs_.gptr(), x.ld,
                         dg[r0:r0 + rc] if dg is not None else None, db[r0:r0 + rc] if db is not None else None, None,
                         dz_accumulate=acc)
            c.pready(bn.weight, bn.bias)
        ctx.tape.append(backward)
    return out


def bn_add_relu_eval(ctx: Ctx, x: Act, bn, res: Act, out: Optional[Act] = None) -> Act:
    """Inference only: ReLU(BatchNorm_running(x) + res) in one pass (uz_bn_apply with its residual operand) - the
    `x = self.batch_norm1(x); x = x + temp; x = F.relu(x)` of a Multiresblock in eval mode (multiresunet.py:79-83)."""
    if ctx.training or ctx.fp32:
        raise RuntimeError("unet_zoo_b200: bn_add_relu_eval is an eval-mode bf16 operator")
    if out is None:
        out = ctx.new_act(x.N, x.H, x.W, x.C)
    aff = ctx.cached(("affine", id(bn)), [bn.weight, bn.bias, bn.running_mean, bn.running_var],
                     lambda: _running_affine(bn, ctx.device))
    ctx.bn_apply(x, aff[2].data_ptr(), aff[3].data_ptr(), True, out, None, residual=res)
    return out


def conv_plain(ctx: Ctx, xs: Sequence[Act], conv, out: Optional[Act] = None, mode: int = L.CONV3X3,
               accumulate: bool = False, stem: Optional[str] = None, need_dx: bool = True) -> Act:
    """Conv2d without normalisation or activation, optionally accumulating into `out` (out += conv(x)): the last
    convolution of ResidualConv.conv_block added onto the skip branch (common_layers.py:191,198) and the two stem
    convolutions of ResUnet summed in place (resunet.py:29-33,56)."""
    if ctx.fp32:
        return _F32.conv_plain(ctx, xs, conv, out=out, mode=mode, accumulate=accumulate, stem=stem, need_dx=need_dx)
    a0 = xs[0]
    cout = conv.out_channels
    if mode in (L.CONV3X3_S2, L.CONV1X1_S2):
        Ho, Wo = (a0.H - 1) // 2 + 1, (a0.W - 1) // 2 + 1
    else:
        Ho, Wo = a0.H, a0.W
    if out is None:
        if accumulate:
            raise ValueError("accumulate needs an existing output")
        out = ctx.new_act(a0.N, Ho, Wo, cout)
    is_stem = stem == "im2col"
    wp = ctx.cached(("conv_plain", id(conv), sum(a.C for a in xs), a0.C if is_stem else 0), [conv.weight],
                    lambda: ctx.pack(conv.weight if is_stem else _weight_for(conv, sum(a.C for a in xs)),
                                     "stem" if is_stem else "conv_fwd", kp=a0.C if is_stem else 0))
    ctx.conv(L.CONV1X1 if is_stem else mode, xs, wp, cout, conv.bias, out.ptr, out.ld, Ho, Wo, None,
             L.FLAG_ACCUMULATE if accumulate else 0)
    if ctx.record:
        cin_total = sum(a.C for a in xs)

        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: conv_plain output has no gradient")
            if conv.bias is not None and conv.bias.requires_grad:
                c.colsum(out.gptr(), out.ld, out.npix, cout, c.pgrad(conv.bias))
                c.pready(conv.bias)
            _conv_backward(c, xs, conv, GradAct(out), mode, is_stem, cin_total, need_dx)
        ctx.tape.append(backward)
    return out


def conv3x3_plus_1x1_fused(ctx: Ctx, h: Act, x: Act, conv2, skip: Callable, deps: Sequence, key, out: Optional[Act],
                           stride2: bool = False) -> Optional[Act]:
    """Inference only: ``conv2(h) + M x + bias`` as ONE implicit GEMM - the operand `x` of a 1x1 branch rides along as an
    extra K segment that only the centre tap reads (UZ_FLAG_RES1X1; stride-2 branch: the even-pixel lattice of `x`,
    UZ_FLAG_RES1X1_S2).  `skip()` returns the branch as (fp32 matrix M [Cout, x.C], fp32 bias [Cout] or None) with any
    BatchNorm already folded in.  Replaces the separate 1x1 launch and the read-modify-write of the output by the
    accumulating 3x3 launch.  Returns None when the halo-box kernel does not take the layer (tiny images): the caller
    falls back to the two-launch form."""
    if ctx.training or ctx.fp32 or h.H < 16 or h.W < 8 or os.environ.get("UZ_RESUNET_UNFUSED") is not None:
        return None
    if (stride2 and (x.H != 2 * h.H or x.W != 2 * h.W)) or (not stride2 and (x.H != h.H or x.W != h.W)):
        return None
    cout = conv2.out_channels
    if out is None:
        out = ctx.new_act(h.N, h.H, h.W, cout)

    def folded():
        m, bias = skip()
        w = torch.zeros((cout, h.C + x.C, 3, 3), dtype=torch.float32, device=ctx.device)
        w[:, :h.C] = conv2.weight.detach()
        w[:, h.C:, 1, 1] = m
        if conv2.bias is not None:
            bias = conv2.bias.detach() if bias is None else bias + conv2.bias.detach()
        return ctx.pack(w, "conv_fwd"), (bias.float().contiguous() if bias is not None else None)
    wp, bias = ctx.cached(key, [conv2.weight, conv2.bias] + list(deps), folded)
    ctx.conv(L.CONV3X3, [h, x], wp, cout, bias, out.ptr, out.ld, h.H, h.W, None,
             L.FLAG_RES1X1_S2 if stride2 else L.FLAG_RES1X1)
    return out


def residual_conv_tail_fused(ctx: Ctx, h: Act, x: Act, conv2, conv_s, bn_s, out: Optional[Act], stride2: bool) -> Optional[Act]:
    """The end of ResidualConv.forward (common_layers.py:196-199), ``conv2(h) + bn_s(conv_s(x))``, in inference mode:
    BatchNorm of the skip branch folded into its weight columns and the bias (conv3x3_plus_1x1_fused)."""
    def skip():
        scale, bias = _fold_bn(conv_s.bias, bn_s)
        return conv_s.weight.detach()[:, :, 0, 0] * scale.view(-1, 1), bias
    deps = [conv_s.weight, conv_s.bias, bn_s.weight, bn_s.bias, bn_s.running_mean, bn_s.running_var]
    return conv3x3_plus_1x1_fused(ctx, h, x, conv2, skip, deps, ("res_tail", id(conv2), id(conv_s), id(bn_s)), out, stride2)


def stem_pair_fused(ctx: Ctx, h: Act, conv2, im2col: Act, conv_stem, out: Optional[Act]) -> Optional[Act]:
    """ResUnet's input stage in inference mode (resunet.py:29-33,56): ``input_layer[3](h) + input_skip(image)`` - the second
    term is a 1x1 GEMM over the stem's im2col operand (k = tap * Cin + ci, zero-padded to `im2col.C` columns)."""
    def skip():
        w = conv_stem.weight.detach()
        co, ci = w.shape[0], w.shape[1]
        m = torch.zeros((co, im2col.C), dtype=torch.float32, device=ctx.device)
        m[:, :9 * ci] = w.permute(0, 2, 3, 1).reshape(co, 9 * ci)
        return m, (conv_stem.bias.detach() if conv_stem.bias is not None else None)
    return conv3x3_plus_1x1_fused(ctx, h, im2col, conv2, skip, [conv_stem.weight, conv_stem.bias],
                                  ("stem_pair", id(conv2), id(conv_stem)), out)


def bilinear_up2x(ctx: Ctx, x: Act, out: Act) -> Act:
    """nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True) written into a concat slice
    (nested_unet.py:33,81-93)."""
    if (out.H, out.W) != (2 * x.H, 2 * x.W):     # odd skip size: Up pads the upsampled tensor (common_layers.py:173-178)
        return pad_into(ctx, bilinear_up2x(ctx, x, ctx.new_act(x.N, 2 * x.H, 2 * x.W, x.C)), out)
    if ctx.fp32:
        return _F32.bilinear_up2x(ctx, x, out)
    L.check(ctx.lib.uz_bilinear2x_fwd(x.ptr, x.ld, x.N, x.H, x.W, x.C, out.ptr, out.ld, ctx.stream), "uz_bilinear2x_fwd")
    ctx.launches += 1
    if ctx.record:
        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: upsampled tensor has no gradient")
            for (r0, rc, acc) in x.grad_writes():
                L.check(c.lib.uz_bilinear2x_bwd(out.gptr() + 2 * r0, out.ld, x.N, x.H, x.W, rc, x.gptr() + 2 * r0, x.ld,
                                                int(acc), c.stream), "uz_bilinear2x_bwd")
                c.launches += 1
        ctx.tape.append(backward)
    return out


@contextlib.contextmanager
def _island_math(tf32: bool):
    """Pin the fp32 matmul / convolution precision of torch's own kernels for the duration of an island.

    torch >= 2.9 has two front ends for the same switches (`allow_tf32` and `fp32_precision`) and refuses to READ one
    after the other was written: use whichever the process is already on."""
    mm, cd = torch.backends.cuda.matmul, torch.backends.cudnn
    try:
        saved = (mm.allow_tf32, cd.allow_tf32)
        legacy = True
    except RuntimeError:            # the application set torch.backends.*.fp32_precision
        saved = (mm.fp32_precision, cd.conv.fp32_precision)
        legacy = False
    if legacy:
        mm.allow_tf32 = cd.allow_tf32 = tf32
    else:
        mm.fp32_precision = cd.conv.fp32_precision = "tf32" if tf32 else "ieee"
    try:
        yield
    finally:
        if legacy:
            mm.allow_tf32, cd.allow_tf32 = saved
        else:
            mm.fp32_precision, cd.conv.fp32_precision = saved


def add_posenc(ctx: Ctx, x: Act, pe: torch.Tensor, out: Optional[Act] = None, pooled: Optional[Act] = None):
    """x + positional encoding (`pe`: fixed fp32 [H, W, C] table, no gradient), unet_transformer.py:128,196,199;
    `pooled`: MaxPool2d(2) of the sum (Sconv_process[0], :153) - when only `pooled` is given the sum is never stored.
    Returns (out, pooled)."""
    if ctx.fp32:
        raise RuntimeError("unet_zoo_b200: add_posenc has no fp32 parity path")
    if out is None and pooled is None:
        out = ctx.new_act(x.N, x.H, x.W, x.C)
    pe = pe.contiguous()
    L.check(ctx.lib.uz_addpe_fwd(x.ptr, x.ld, pe.data_ptr(), x.N, x.H, x.W, x.C, out.ptr if out else None,
                                 out.ld if out else 0, pooled.ptr if pooled else None, pooled.ld if pooled else 0,
                                 ctx.stream), "uz_addpe_fwd")
    ctx.launches += 1
    if ctx.record:
        def backward(c: Ctx):
            has_dy = out is not None and out.has_grad()
            has_dp = pooled is not None and pooled.has_grad()
            if not (has_dy or has_dp):
                raise RuntimeError("unet_zoo_b200: add_posenc output has no gradient")
            runs = x.grad_writes()
            if len(runs) != 1:          # the table's row pitch is the full channel count
                raise NotImplementedError("add_posenc: input gradient partly written by another consumer")
            L.check(c.lib.uz_addpe_bwd(x.ptr, x.ld, pe.data_ptr(), x.N, x.H, x.W, x.C,
                                       out.gptr() if has_dy else None, out.ld if has_dy else 0,
                                       pooled.gptr() if has_dp else None, pooled.ld if has_dp else 0,
                                       x.gptr(), x.ld, int(runs[0][2]), c.stream), "uz_addpe_bwd")
            c.launches += 1
        ctx.tape.append(backward)
    return out, pooled


def adaptive_avgpool(ctx: Ctx, x: Act, OH: int, OW: int) -> Act:
    """F.adaptive_avg_pool2d(x, (OH, OW)) (unet_transformer.py:201-203)."""
    out = ctx.new_act(x.N, OH, OW, x.C)
    L.check(ctx.lib.uz_adaptive_avgpool_fwd(x.ptr, x.ld, x.N, x.H, x.W, x.C, out.ptr, out.ld, OH, OW, ctx.stream),
            "uz_adaptive_avgpool_fwd")
    ctx.launches += 1
    if ctx.record:
        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: adaptive_avgpool output has no gradient")
            for (r0, rc, acc) in x.grad_writes():
                L.check(c.lib.uz_adaptive_avgpool_bwd(out.gptr() + 2 * r0, out.ld, x.N, x.H, x.W, rc, OH, OW,
                                                      x.gptr() + 2 * r0, x.ld, int(acc), c.stream), "uz_adaptive_avgpool_bwd")
                c.launches += 1
        ctx.tape.append(backward)
    return out


def bilinear_resize(ctx: Ctx, x: Act, OH: int, OW: int, out: Optional[Act] = None, align_corners: bool = True) -> Act:
    """F.interpolate(x, size=(OH, OW), mode="bilinear", align_corners=...): True for unet_transformer.py:222, False for
    u2net.py:19-23 (`_upsample_like`)."""
    if out is None:
        out = ctx.new_act(x.N, OH, OW, x.C)
    L.check(ctx.lib.uz_bilinear_resize_fwd(x.ptr, x.ld, x.N, x.H, x.W, x.C, out.ptr, out.ld, OH, OW, int(align_corners),
                                           ctx.stream), "uz_bilinear_resize_fwd")
    ctx.launches += 1
    if ctx.record:
        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: bilinear_resize output has no gradient")
            for (r0, rc, acc) in x.grad_writes():
                need = c.lib.uz_bilinear_resize_bwd_workspace_bytes(x.N, x.H, x.W, rc, OH, OW)
                ws = c.workspace(need) if need else None
                L.check(c.lib.uz_bilinear_resize_bwd(out.gptr() + 2 * r0, out.ld, x.N, x.H, x.W, rc, OH, OW,
                                                     x.gptr() + 2 * r0, x.ld, int(acc), int(align_corners),
                                                     ws.data_ptr() if need else None, need, c.stream),
                        "uz_bilinear_resize_bwd")
                c.launches += 1
        ctx.tape.append(backward)
    return out


def conv_padded_out(ctx: Ctx, xs: Sequence[Act], conv, out: Act) -> Act:
    """Conv2d(C, K, 3, padding=1) with K < 8 output channels (the side convolutions of U^2-Net, u2net.py:238-243) on the
    tensor-core path: the filter bank is zero-padded to 8 output channels, `out` is an 8-channel slot whose channels
    K..7 come out as exact zeros.  Backward: weight / bias gradients of the K real rows, data gradient as usual."""
    if ctx.fp32:
        raise RuntimeError("unet_zoo_b200: conv_padded_out has no fp32 parity path")
    K = conv.out_channels
    if out.C != 8 or K > 8:
        raise ValueError("conv_padded_out: the output slot has 8 channels")
    a0 = xs[0]
    cin_total = sum(a.C for a in xs)

    def padded():
        w = torch.nn.functional.pad(conv.weight.detach(), (0, 0, 0, 0, 0, 0, 0, 8 - K))
        b = torch.nn.functional.pad(conv.bias.detach(), (0, 8 - K)) if conv.bias is not None else None
        return w, b
    if ctx.training:
        w8, b8 = padded()
        wp = ctx.pack(w8, "conv_fwd")
    else:
        wp, b8 = ctx.cached(("conv_padded", id(conv)), [conv.weight, conv.bias],
                            lambda: (lambda wb: (ctx.pack(wb[0], "conv_fwd"), wb[1]))(padded()))
    ctx.conv(L.CONV3X3, xs, wp, 8, b8, out.ptr, out.ld, a0.H, a0.W, None, 0)
    if ctx.record:
        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: conv_padded_out output has no gradient")
            dz = GradAct(out)
            db = c.pgrad(conv.bias) if conv.bias is not None else None
            if db is not None:
                tmp = c.f32(8)
                c.colsum(dz.ptr, dz.ld, out.npix, 8, tmp)
                db.copy_(tmp[:K])
            dw = c.pgrad(conv.weight)
            if dw is not None:
                tgt = c.f32(8, cin_total, 3, 3)
                off = 0
                for a in xs:
                    c.wgrad(L.CONV3X3, a.ptr, a.N, a.H, a.W, a.C, a.ld, dz.ptr, 8, dz.ld, tgt, cin_total, off)
                    off += a.C
                dw.copy_(tgt[:K])
            c.pready(conv.weight, conv.bias)
            wpd = c.pack(padded()[0], "conv_dgrad")                 # [Cin][9][8]
            off = 0
            for a in xs:
                for (r0, rc, acc) in a.grad_writes():
                    c.conv(L.CONV3X3_DGRAD, [dz], wpd[off + r0: off + r0 + rc], rc, None, a.gptr() + 2 * r0, a.ld, a.H, a.W,
                           None, L.FLAG_ACCUMULATE if acc else 0)
                off += a.C
        ctx.tape.append(backward)
    return out


def fused_side_head(ctx: Ctx, fuse: Act, outconv, K: int, nsides: int) -> torch.Tensor:
    """The output stage of U^2-Net in ONE pass over the buffer of upsampled side maps (side i in channels [8 i, 8 i + K)):
    logits[:, :K] = outconv(cat(sides)) (u2net.py:285), logits[:, K (1 + i) : K (2 + i)] = side i itself (fp32 NCHW, the
    tensors the reference returns, :287-295).  Backward: one pass gives the gradient of every side map (its own logit
    gradient + outconv's share) and outconv's weight / bias gradients (uz_side_head_fwd / _bwd)."""
    rows = K * (1 + nsides)
    dev = ctx.device
    if fuse.C < 8 * nsides:
        raise ValueError("fused_side_head: the slot buffer is too narrow")
    w = outconv.weight.detach().reshape(K, nsides * K).contiguous()
    logits = torch.empty((fuse.N, rows, fuse.H, fuse.W), dtype=torch.float32, device=dev)
    L.check(ctx.lib.uz_side_head_fwd(fuse.ptr, fuse.ld, fuse.N, fuse.H, fuse.W, nsides, K, w.data_ptr(), _ptr(outconv.bias),
                                     logits.data_ptr(), ctx.stream), "uz_side_head_fwd")
    ctx.launches += 1
    idx = len(ctx.outputs)
    ctx.outputs.append(logits)
    if ctx.record:
        def backward(c: Ctx):
            g = c.out_grads[idx] if idx < len(c.out_grads) else None
            if g is None:
                raise RuntimeError("unet_zoo_b200: fused_side_head output has no gradient")
            dl = g.contiguous().float()
            dw = c.pgrad(outconv.weight)
            db = c.pgrad(outconv.bias) if outconv.bias is not None else None
            dw = dw if dw is not None else c.f32(K, nsides * K)
            db = db if db is not None else c.f32(K)
            ws = c.workspace(c.lib.uz_side_head_bwd_workspace_bytes(nsides, K))
            if any(acc for _, _, acc in fuse.grad_writes()):
                raise NotImplementedError("fused_side_head input already has a gradient")
            L.check(c.lib.uz_side_head_bwd(fuse.ptr, fuse.ld, dl.data_ptr(), fuse.N, fuse.H, fuse.W, nsides, K, w.data_ptr(),
                                           fuse.gptr(), fuse.ld, dw.data_ptr(), db.data_ptr(), ws.data_ptr(), ws.numel(),
                                           c.stream), "uz_side_head_bwd")
            c.launches += 3
            c.pready(outconv.weight, outconv.bias)
        ctx.tape.append(backward)
    return logits


def residual_add(ctx: Ctx, a: Act, b: Act, out: Optional[Act] = None, pooled: Optional[Act] = None,
                 relu: bool = False) -> Act:
    """out = [ReLU](a + b) with optional MaxPool2d(2, 2) of the result as a second output: `return hx1d + hxin` of the
    RSU blocks (u2net.py:78,127,171,210,237) followed by the stage pooling (u2net.py:222-232); with `relu` the
    `x = x + temp; x = F.relu(x)` of Multiresblock / Respath (multiresunet.py:82-83,125-126,132-133)."""
    if ctx.fp32:
        raise RuntimeError("unet_zoo_b200: residual_add has no fp32 parity path")
    if out is None:
        out = ctx.new_act(a.N, a.H, a.W, a.C)
    L.check(ctx.lib.uz_residual_add_fwd(a.ptr, a.ld, b.ptr, b.ld, a.N, a.H, a.W, a.C, out.ptr, out.ld,
                                        pooled.ptr if pooled else None, pooled.ld if pooled else 0, int(relu), ctx.stream),
            "uz_residual_add_fwd")
    ctx.launches += 1
    if ctx.record:
        def backward(c: Ctx):
            has_dy = out.has_grad()
            has_dp = pooled is not None and pooled.has_grad()
            if not (has_dy or has_dp):
                raise RuntimeError("unet_zoo_b200: residual_add output has no gradient")
            ra, rb = a.grad_writes(), b.grad_writes()
            if len(ra) != 1 or len(rb) != 1:
                raise NotImplementedError("residual_add: input gradient partly written by another consumer")
            L.check(c.lib.uz_residual_add_bwd(out.ptr, out.ld, a.N, a.H, a.W, a.C, out.gptr() if has_dy else None,
                                              out.ld if has_dy else 0, pooled.gptr() if has_dp else None,
                                              pooled.ld if has_dp else 0, a.gptr(), a.ld, int(ra[0][2]), b.gptr(), b.ld,
                                              int(rb[0][2]), int(relu), c.stream), "uz_residual_add_bwd")
            c.launches += 1
        ctx.tape.append(backward)
    return out


class BnBranch:
    """One operand of :func:`dual_bn_add_relu`: a raw (pre-BatchNorm) tensor, its batch statistics, whether a ReLU follows
    the BatchNorm, and how its gradient travels on (into a convolution's backward, or into the tensor's gradient buffer)."""
    __slots__ = ("z", "bnbuf", "relu", "conv", "xs", "mode", "need_dx", "cin_total")


def bn_branch_conv(ctx: Ctx, xs: Sequence[Act], conv, bn, relu: bool, mode: int = L.CONV3X3, need_dx: bool = True) -> BnBranch:
    """Convolution + train-mode BatchNorm statistics, the normalised tensor left to the consumer (no apply pass)."""
    a0 = xs[0]
    br = BnBranch()
    br.conv, br.xs, br.mode, br.need_dx, br.relu = conv, list(xs), mode, need_dx, relu
    br.cin_total = sum(a.C for a in xs)
    cout = conv.out_channels
    br.z = ctx.new_act(a0.N, a0.H, a0.W, cout)
    wp = ctx.pack(_weight_for(conv, br.cin_total), "conv_fwd")
    stats = ctx.f32(ctx.stats_rows, 2, cout)
    ctx.conv(mode, xs, wp, cout, conv.bias, br.z.ptr, br.z.ld, a0.H, a0.W, stats, L.FLAG_STATS)
    br.bnbuf = ctx.f32(4, cout)
    ctx.bn_finalize(stats, ctx.stats_rows, cout, a0.npix, bn, br.bnbuf)
    return br


def bn_branch_act(ctx: Ctx, x: Act, bn) -> BnBranch:
    """Train-mode BatchNorm statistics of an existing activation (no ReLU behind it), normalised tensor left to the consumer."""
    br = BnBranch()
    br.conv, br.xs, br.mode, br.need_dx, br.relu, br.cin_total = None, None, 0, True, False, 0
    br.z = x
    stats = ctx.f32(ctx.reduce_rows, 2, x.C)
    L.check(ctx.lib.uz_bn_stats(x.ptr, x.ld, x.npix, x.C, stats.data_ptr(), ctx.stream), "uz_bn_stats")
    ctx.launches += 1
    br.bnbuf = ctx.f32(4, x.C)
    ctx.bn_finalize(stats, ctx.reduce_rows, x.C, x.npix, bn, br.bnbuf)
    return br


def _dual_backward(c: Ctx, A: BnBranch, B: BnBranch, dy_ptr: int, dy_ld: int, outer=None):
    """Backward of relu([relu](BN(A.z)) + BN(B.z)): both BatchNorm backwards in one reduce + one apply pass, then the
    branches' convolution backwards.  `outer` = (bnbuf, coef) of a BatchNorm that follows (see dual_bn_add_relu_bn)."""
    a, b = A.z, B.z
    dst = []
    for br in (A, B):
        if br.conv is not None:          # dz overwrites z in place, then the convolution's backward consumes it
            dst.append((br.z.ptr, br.z.ld, 0, c.pgrad(br.conv.bias) if br.conv.bias is not None else None))
        else:
            runs = br.z.grad_writes()
            if len(runs) != 1:
                raise NotImplementedError("dual_bn_add_relu: operand gradient partly written by another consumer")
            dst.append((br.z.gptr(), br.z.ld, int(runs[0][2]), None))
    ws = c.workspace(c.lib.uz_dual_bn_bwd_workspace_bytes(a.C))
    with _Timed("bn_bwd", 2.0 * a.npix * a.C * 8):
        L.check(c.lib.uz_dual_bn_add_relu_bwd(a.ptr, a.ld, A.bnbuf.data_ptr(), int(A.relu), b.ptr, b.ld,
                                              B.bnbuf.data_ptr(), a.N, a.H, a.W, a.C, dy_ptr, dy_ld,
                                              outer[0].data_ptr() if outer else None, outer[1].data_ptr() if outer else None,
                                              dst[0][0], dst[0][1], dst[0][2], _ptr(dst[0][3]),
                                              dst[1][0], dst[1][1], dst[1][2], _ptr(dst[1][3]),
                                              ws.data_ptr(), ws.numel(), c.stream), "uz_dual_bn_add_relu_bwd")
    c.launches += 4
    for br in (A, B):
        if br.conv is not None:
            c.pready(br.conv.bias)
            _conv_backward(c, br.xs, br.conv, br.z, br.mode, False, br.cin_total, br.need_dx)


def _dual_forward(ctx: Ctx, A: BnBranch, B: BnBranch, out: Optional[Act], stats: Optional[torch.Tensor]) -> Act:
    if ctx.fp32 or not ctx.training:
        raise RuntimeError("unet_zoo_b200: dual_bn_add_relu is a train-mode bf16 operator")
    a, b = A.z, B.z
    if (a.N, a.H, a.W, a.C) != (b.N, b.H, b.W, b.C) or B.relu:
        raise ValueError("dual_bn_add_relu: operands must have one shape and only the first may carry a ReLU")
    if out is None:
        out = ctx.new_act(a.N, a.H, a.W, a.C)
    with _Timed("bn_apply", 2.0 * a.npix * a.C * 3):
        L.check(ctx.lib.uz_dual_bn_add_relu_fwd(a.ptr, a.ld, A.bnbuf.data_ptr(), int(A.relu), b.ptr, b.ld, B.bnbuf.data_ptr(),
                                                a.N, a.H, a.W, a.C, out.ptr, out.ld, _ptr(stats), ctx.stream),
                "uz_dual_bn_add_relu_fwd")
    ctx.launches += 1
    return out


def dual_bn_add_relu(ctx: Ctx, A: BnBranch, B: BnBranch, out: Optional[Act] = None,
                     stats: Optional[torch.Tensor] = None) -> Act:
    """out = ReLU([ReLU](BN(A.z)) + BN(B.z)) in one pass over the two raw tensors (uz_dual_bn_add_relu_fwd / _bwd): the
    tails of Multiresblock and Respath (multiresunet.py:79-83,122-126).  Training mode only (batch statistics).
    `stats` ([reduce_rows][2][C] fp32): receives the sum / sum-of-squares partials of `out` for the BatchNorm that follows
    (hand it to bn_act(stats=...))."""
    out = _dual_forward(ctx, A, B, out, stats)
    if ctx.record:
        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: dual_bn_add_relu output has no gradient")
            _dual_backward(c, A, B, out.gptr(), out.ld)
        ctx.tape.append(backward)
    return out


def dual_bn_add_relu_bn(ctx: Ctx, A: BnBranch, B: BnBranch, bn, out: Optional[Act] = None,
                        pooled: Optional[Act] = None) -> Act:
    """out = BatchNorm(ReLU([ReLU](BN(A.z)) + BN(B.z))) [+ MaxPool2d(2, 2) of it]: the whole tail of a Multiresblock /
    Respath step (multiresunet.py:79-84,122-127).  Forward: the fused add emits the sum s and its statistics, one apply
    pass normalises it.  Backward without a pooled gradient: the BatchNorm's backward is reduced to its coefficients
    (uz_bn_bwd_coef, one pass over s and d(out)) and applied on the fly inside the fused add's backward, which recomputes
    s from the raw tensors - d(s) is never written or read.  With a pooled gradient (window routing needs the stored
    output) the BatchNorm backward runs on its own first."""
    a = A.z
    rows = ctx.reduce_rows
    stats = ctx.f32(rows, 2, a.C)
    s = _dual_forward(ctx, A, B, None, stats)
    bnbuf = ctx.f32(4, a.C)
    ctx.bn_finalize(stats, rows, a.C, a.npix, bn, bnbuf)
    if out is None:
        out = ctx.new_act(a.N, a.H, a.W, a.C)
    ctx.bn_apply(s, bnbuf[2].data_ptr(), bnbuf[3].data_ptr(), False, out, pooled)
    if ctx.record:
        def backward(c: Ctx):
            has_dy = out.has_grad()
            has_dp = pooled is not None and pooled.has_grad()
            if not (has_dy or has_dp):
                raise RuntimeError("unet_zoo_b200: dual_bn_add_relu_bn output has no gradient")
            if has_dp or os.environ.get("UZ_DUAL_BN_SPLIT") == "1":
                c.bn_bwd(s, out.gptr() if has_dy else None, out.ld, pooled.gptr() if has_dp else None,
                         pooled.ld if has_dp else 0, bnbuf, False, s.gptr(), s.ld, None, None, None)
                _dual_backward(c, A, B, s.gptr(), s.ld)
                return
            coef = c.f32(2, a.C)
            ws = c.workspace(c.lib.uz_bn_bwd_workspace_bytes(a.C))
            with _Timed("bn_bwd", 2.0 * a.npix * a.C * 2):
                L.check(c.lib.uz_bn_bwd_coef(s.ptr, s.ld, out.gptr(), out.ld, a.N, a.H, a.W, a.C, bnbuf.data_ptr(), 0,
                                             coef.data_ptr(), None, None, None, 0, ws.data_ptr(), ws.numel(), c.stream),
                        "uz_bn_bwd_coef")
            c.launches += 2
            _dual_backward(c, A, B, out.gptr(), out.ld, outer=(bnbuf, coef))
        ctx.tape.append(backward)
    return out


def nhwc_logits(ctx: Ctx, y: Act, K: int) -> torch.Tensor:
    """Channels [0, K) of `y` as the model's fp32 NCHW output: MultiResUnet returns the BatchNorm2d(affine=False) of its
    1x1 conv_final directly (multiresunet.py:238-240); that conv-BN pair runs with the classes padded to 8 channels.
    Backward: the incoming fp32 gradient becomes the (zero-padded) bf16 gradient of `y`."""
    if ctx.fp32:
        raise RuntimeError("unet_zoo_b200: nhwc_logits has no fp32 parity path")
    if y.C != 8 or not 1 <= K <= 8:
        raise ValueError("nhwc_logits: an 8-channel activation and 1 <= K <= 8 classes")
    logits = torch.empty((y.N, K, y.H, y.W), dtype=torch.float32, device=ctx.device)
    L.check(ctx.lib.uz_nhwc_to_nchw_f32(y.ptr, y.ld, y.N, y.H, y.W, K, logits.data_ptr(), ctx.stream), "uz_nhwc_to_nchw_f32")
    ctx.launches += 1
    idx = len(ctx.outputs)
    ctx.outputs.append(logits)
    if ctx.record:
        def backward(c: Ctx):
            g = c.out_grads[idx] if idx < len(c.out_grads) else None
            if g is None:
                raise RuntimeError("unet_zoo_b200: the model output received no gradient")
            runs = y.grad_writes()
            if len(runs) != 1 or runs[0][2]:
                raise NotImplementedError("nhwc_logits: the activation already has a gradient")
            dl = g.contiguous().float()
            L.check(c.lib.uz_nchw_to_nhwc(dl.data_ptr(), y.N, K, y.H, y.W, y.gptr(), y.ld, c.stream), "uz_nchw_to_nhwc")
            c.launches += 1
        ctx.tape.append(backward)
    return logits


def token_island(ctx: Ctx, x: Act, fn: Callable[[torch.Tensor], torch.Tensor], params: Sequence[torch.nn.Parameter],
                 out: Optional[Act] = None) -> Act:
    """A sub-graph that works on the activation as bf16 TOKENS: ``fn`` maps ``[N, H*W, C]`` (the NHWC buffer seen flat, no
    layout change) to ``[N, H*W, C']``.  Used for the attention bottlenecks, whose matrix products / softmaxes are the
    library's own kernels wrapped as autograd Functions (unet_zoo_b200/attention.py); torch's autograd only chains them
    and supplies the elementwise glue.  Gradients of `params` land in the flat gradient buffer."""
    xs = list(x) if isinstance(x, (list, tuple)) else [x]     # several inputs of the same spatial size (cross-attention)
    x = xs[0]
    xins = []
    for a in xs:
        t = a.buf.data[..., a.c0:a.c0 + a.C].reshape(a.N, a.H * a.W, a.C).detach()
        xins.append(t if t.is_contiguous() else t.contiguous())
    if ctx.record:
        for t in xins:
            t.requires_grad_(True)
        with torch.enable_grad():
            y = fn(*xins)
    else:
        y = fn(*xins)
    if y.dim() != 3 or y.shape[0] != x.N or y.shape[1] != x.H * x.W or y.dtype != torch.bfloat16:
        raise RuntimeError(f"unet_zoo_b200: token island returned {tuple(y.shape)} {y.dtype} for {(x.N, x.H * x.W, x.C)}")
    if out is None:
        out = ctx.new_act(x.N, x.H, x.W, y.shape[2])
    out.buf.data[..., out.c0:out.c0 + out.C].copy_(y.detach().reshape(x.N, x.H, x.W, -1))
    if ctx.record:
        params = [p for p in params if p.requires_grad]

        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: token island output has no gradient")
            dy = out.buf.grad_buf()[..., out.c0:out.c0 + out.C].reshape(x.N, x.H * x.W, -1)
            grads = torch.autograd.grad([y], xins + params, [dy.contiguous()], allow_unused=True)
            for a, g in zip(xs, grads[:len(xs)]):
                ga = a.buf.grad_buf()
                dx = (g if g is not None else torch.zeros_like(xins[0][..., :a.C])).reshape(a.N, a.H, a.W, a.C)
                for (r0, rc, acc) in a.grad_writes():
                    dst = ga[..., a.c0 + r0:a.c0 + r0 + rc]
                    if acc:
                        dst.add_(dx[..., r0:r0 + rc])
                    else:
                        dst.copy_(dx[..., r0:r0 + rc])
            for p, g in zip(params, grads[len(xs):]):
                if g is not None:
                    c.pgrad(p).copy_(g)
            c.pready(*params)
        ctx.tape.append(backward)
    return out


def torch_island(ctx: Ctx, x: Act, fn: Callable[[torch.Tensor], torch.Tensor], params: Sequence[torch.nn.Parameter],
                 out: Optional[Act] = None) -> Act:
    """A sub-graph of stock torch ops between two fused operators: ``fn`` maps the NCHW fp32 image of `x` to an NCHW
    tensor of the same spatial size (the attention bottleneck of transatt_unet, transatt_unet.py:143-154 - small
    batched matmuls / softmaxes at 1/16 resolution that are not on the path BASELINE.json names).  The island runs on
    the executor's stream; in training its local autograd graph is kept and differentiated from the tape, the gradients
    of `params` (the island's own nn.Parameters) land in the flat gradient buffer like every other parameter's."""
    xin = x.buf.data[..., x.c0:x.c0 + x.C].permute(0, 3, 1, 2).float().contiguous()
    # fp32 parity mode: strict fp32 (1e-4 bar).  bf16 mode: the island's operands are bf16-rounded activations anyway, so
    # its matmuls / 1x1 convolutions may use the TF32 tensor-core path of cuBLAS / cuDNN (fp32 accumulate)
    tf32 = (not ctx.fp32) and os.environ.get("UZ_ISLAND_TF32", "1") != "0"
    with _island_math(tf32):
        if ctx.record:
            xin.requires_grad_(True)
            with torch.enable_grad():
                y = fn(xin)
        else:
            y = fn(xin)
    if y.dim() != 4 or y.shape[0] != x.N or y.shape[2:] != (x.H, x.W):
        raise RuntimeError(f"unet_zoo_b200: torch island returned {tuple(y.shape)} for an input of "
                           f"{(x.N, x.C, x.H, x.W)}")
    if out is None:
        out = ctx.new_act(x.N, x.H, x.W, y.shape[1])
    out.buf.data[..., out.c0:out.c0 + out.C].copy_(y.detach().permute(0, 2, 3, 1))
    if ctx.record:
        params = [p for p in params if p.requires_grad]

        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: torch island output has no gradient")
            dy = out.buf.grad_buf()[..., out.c0:out.c0 + out.C].permute(0, 3, 1, 2).to(y.dtype)
            with _island_math(tf32):
                grads = torch.autograd.grad([y], [xin] + params, [dy], allow_unused=True)
            gx = x.buf.grad_buf()
            dx = grads[0].permute(0, 2, 3, 1)
            for (r0, rc, acc) in x.grad_writes():
                dst = gx[..., x.c0 + r0:x.c0 + r0 + rc]
                if acc:
                    dst.add_(dx[..., r0:r0 + rc])
                else:
                    dst.copy_(dx[..., r0:r0 + rc])
            for p, g in zip(params, grads[1:]):
                if g is not None:
                    c.pgrad(p).copy_(g)
            c.pready(*params)
        ctx.tape.append(backward)
    return out


def attention_gate(ctx: Ctx, g: Act, x: Act, att, out: Act) -> Act:
    """AttentionBlock.forward (attention_unet.py:34-40): out = x * sigmoid(BN(w_psi . ReLU(BN(W_g g) + BN(W_x x)))),
    written into a channel slice of the decoder's concat buffer.  The 1x1 convolutions are tcgen05 GEMMs with the
    BatchNorm statistics in their epilogue; everything else is three bandwidth kernels (csrc/gate.cu)."""
    if ctx.fp32:
        return _F32.attention_gate(ctx, g, x, att, out)
    conv_g, bn_g = att.w_g[0], att.w_g[1]
    conv_x, bn_x = att.w_x[0], att.w_x[1]
    conv_p, bn_p = att.psi[0], att.psi[1]
    F = conv_g.out_channels
    npix = x.npix
    lib = ctx.lib
    if ctx.training:
        zg, bg = conv_stats(ctx, [g], conv_g, bn_g, L.CONV1X1)
        zx, bx = conv_stats(ctx, [x], conv_x, bn_x, L.CONV1X1)
    else:
        zg, zx = ctx.new_act(g.N, g.H, g.W, F), ctx.new_act(x.N, x.H, x.W, F)
        ctx.conv(L.CONV1X1, [g], ctx.cached(("gate_w", id(conv_g)), [conv_g.weight], lambda: ctx.pack(conv_g.weight, "conv_fwd")),
                 F, conv_g.bias, zg.ptr, zg.ld, g.H, g.W)
        ctx.conv(L.CONV1X1, [x], ctx.cached(("gate_w", id(conv_x)), [conv_x.weight], lambda: ctx.pack(conv_x.weight, "conv_fwd")),
                 F, conv_x.bias, zx.ptr, zx.ld, x.H, x.W)
        bg, bx = (ctx.cached(("affine", id(b)), [b.weight, b.bias, b.running_mean, b.running_var],
                             lambda b=b: _running_affine(b, ctx.device)) for b in (bn_g, bn_x))
    w_psi = conv_p.weight.detach().reshape(-1).contiguous()
    psi_raw = ctx.f32(npix)
    partial = ctx.f32(ctx.reduce_rows, 2, 1)
    L.check(lib.uz_gate_psi_fwd(zg.ptr, zg.ld, zx.ptr, zx.ld, npix, F, bg.data_ptr(), bx.data_ptr(), w_psi.data_ptr(),
                                _ptr(conv_p.bias), psi_raw.data_ptr(), partial.data_ptr(), ctx.stream), "uz_gate_psi_fwd")
    ctx.launches += 1
    if ctx.training:
        bp = ctx.f32(4, 1)
        ctx.bn_finalize(partial, ctx.reduce_rows, 1, npix, bn_p, bp)
    else:
        bp = ctx.cached(("affine", id(bn_p)), [bn_p.weight, bn_p.bias, bn_p.running_mean, bn_p.running_var],
                        lambda: _running_affine(bn_p, ctx.device))
    L.check(lib.uz_gate_apply(x.ptr, x.ld, psi_raw.data_ptr(), bp.data_ptr(), npix, x.C, out.ptr, out.ld, ctx.stream),
            "uz_gate_apply")
    ctx.launches += 1

    if ctx.record:
        def backward(c: Ctx):
            if not out.has_grad():
                raise RuntimeError("unet_zoo_b200: attention gate output has no gradient")
            runs = x.grad_writes()
            if all(acc for _, _, acc in runs):
                acc_dx = 1
            elif not any(acc for _, _, acc in runs):
                acc_dx = 0
            else:  # partly written: clear the untouched channel runs, then accumulate everywhere
                for (r0, rc, acc) in runs:
                    if not acc:
                        x.buf.grad_buf()[..., x.c0 + r0:x.c0 + r0 + rc].zero_()
                acc_dx = 1
            dr = c.new_act(x.N, x.H, x.W, F)
            dw = c.pgrad(conv_p.weight)
            dwv = dw.view(-1) if dw is not None else c.f32(F)
            db = c.pgrad(conv_p.bias) if conv_p.bias is not None else None
            need = c.lib.uz_gate_bwd_workspace_bytes(npix, F)
            ws = c.workspace(need)
            rc_ = c.lib.uz_gate_bwd(x.ptr, x.ld, out.gptr(), out.ld, psi_raw.data_ptr(), bp.data_ptr(), npix, x.C,
                                    x.gptr(), x.ld, acc_dx, zg.ptr, zg.ld, zx.ptr, zx.ld, F, bg.data_ptr(), bx.data_ptr(),
                                    w_psi.data_ptr(), dr.ptr, dr.ld, dwv.data_ptr(), _ptr(db), _ptr(c.pgrad(bn_p.weight)),
                                    _ptr(c.pgrad(bn_p.bias)), ws.data_ptr(), ws.numel(), c.stream)
            L.check(rc_, "uz_gate_bwd")
            c.launches += 5
            c.pready(conv_p.weight, conv_p.bias, bn_p.weight, bn_p.bias)
            for (z, bnb, conv, bn, src) in ((zg, bg, conv_g, bn_g, g), (zx, bx, conv_x, bn_x, x)):
                c.bn_bwd(z, dr.ptr, dr.ld, None, 0, bnb, False, z.ptr, z.ld, c.pgrad(bn.weight), c.pgrad(bn.bias),
                         c.pgrad(conv.bias) if conv.bias is not None else None)
                c.pready(bn.weight, bn.bias, conv.bias)
                _conv_backward(c, [src], conv, z, L.CONV1X1, False, src.C, True)
        ctx.tape.append(backward)
    return out


from . import engine_f32 as _F32  # noqa: E402  (fp32 parity mode; imports this module back)
