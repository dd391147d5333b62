"""Data-parallel gradient sync, overlapped with backward.

Replaces what DeepSpeed ZeRO does for the reference (finetuner/ds_config.json:27-45: bucketed gradient reduction,
``overlap_comm``, enabled at FT.py:911-923): gradients live in flat buckets (each ``param.grad`` is a view), buckets are
filled in reverse parameter order as backward runs, and a full bucket is all-reduced at once on a separate
communication stream while backward continues.  Under gradient accumulation only the last micro-batch reduces.
Plain DP: every rank holds a full replica (no ZeRO partitioning).

Gradient precision.  The reference reduces in fp32 (ds_config.json:45 ``"communication_data_type": "fp32"``) and
DeepSpeed accumulates micro-batch gradients of 16-bit parameters in fp32.  ``main_grads=True`` (the default whenever
a parameter is bf16) does the same: every bucket is an fp32 buffer, each parameter gets a ``main_grad`` view of it,
the bf16 ``param.grad`` autograd produces for a micro-batch is added into ``main_grad`` by the grad-ready hook and
dropped, the fused lm_head loss adds its weight gradient into ``main_grad`` straight from the dW GEMM, and the
all-reduce and the optimizer consume the fp32 buckets.  ``main_grads=False`` keeps gradients, accumulation and the wire
in the parameter dtype (half the NVLink bytes, bf16 rounding on every add).

Backends
  "ftk"   the C ABI's NCCL communicator (ftk_dp_*): one ncclAllReduce per bucket on a high-priority stream over
          NVLink/NVSwitch, pre-scaled by 1/world.  GPU only.
  "gloo"  torch.distributed on CPU tensors -- exists so the bucketing / hook / accumulation logic is covered by
          world_size-2 tests on machines without GPUs.  Never used on a GPU.
"""
from __future__ import annotations

import contextlib
import ctypes
from typing import Iterable, List, Optional

import torch
import torch.distributed as dist

from . import _lib


class _Bucket:
    __slots__ = ("flat", "params", "ready", "launched", "work")

    def __init__(self, flat: torch.Tensor, params: List[torch.nn.Parameter]):
        self.flat, self.params, self.work = flat, params, None
        self.ready = set()      # ids of the parameters whose gradient is final in this optimizer step
        self.launched = False   # the bucket's all-reduce has been issued in this optimizer step


class GradSync:
    def __init__(self, params: Iterable[torch.nn.Parameter], rank: int, world: int, bucket_bytes: int = 128 << 20,
                 backend: str = "ftk", grad_dtype: Optional[torch.dtype] = None, average: bool = True,
                 main_grads: Optional[bool] = None):
        self.rank, self.world, self.backend, self.average = rank, world, backend, average
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        if main_grads is None:
            main_grads = grad_dtype is None and any(p.dtype == torch.bfloat16 for p in self.params)
        self.main_grads = bool(main_grads)
        if self.main_grads:
            grad_dtype = torch.float32
        self.device = self.params[0].device
        if backend == "ftk" and self.device.type != "cuda":
            raise _lib.FtkError('backend "ftk" needs CUDA parameters (use "gloo" only for CPU tests)')
        if backend not in ("ftk", "gloo"):
            raise ValueError(backend)
        self.sync_now = True
        self.launched = 0
        self._handle = None
        self.buckets: List[_Bucket] = []
        self._bucket_of = {}
        # reverse registration order ~ the order backward produces gradients
        cur: List[torch.nn.Parameter] = []
        cur_bytes = 0
        groups: List[List[torch.nn.Parameter]] = []
        for p in reversed(self.params):
            dt = grad_dtype or p.dtype
            nbytes = p.numel() * torch.empty((), dtype=dt).element_size()
            if cur and (cur_bytes + nbytes > bucket_bytes or (grad_dtype or cur[0].dtype) != dt):
                groups.append(cur)
                cur, cur_bytes = [], 0
            cur.append(p)
            cur_bytes += nbytes
        if cur:
            groups.append(cur)
        for g in groups:
            dt = grad_dtype or g[0].dtype
            n = sum((p.numel() + 7) // 8 * 8 for p in g)  # keep every view 16-byte aligned
            flat = torch.zeros(n, dtype=dt, device=self.device)
            off = 0
            for p in g:
                view = flat[off:off + p.numel()].view_as(p)
                if self.main_grads:
                    p.main_grad = view  # the fused lm_head loss looks for this attribute (ops._wgrad_target)
                    p.grad = None
                else:
                    p.grad = view
                off += (p.numel() + 7) // 8 * 8
            b = _Bucket(flat, g)
            self.buckets.append(b)
            for p in g:
                self._bucket_of[p] = b
                p.register_post_accumulate_grad_hook(self._on_ready)
        if backend == "ftk":
            self._init_ftk()

    # ------------------------------------------------------------------ communicator bootstrap
    def _init_ftk(self) -> None:
        L = _lib.lib()
        buf = ctypes.create_string_buffer(128)
        if self.world > 1:
            if not dist.is_initialized():
                raise _lib.FtkError("torch.distributed must be initialised to exchange the NCCL unique id")
            box = [None]
            if self.rank == 0:
                _lib.check(L.ftk_dp_unique_id(buf), "ftk_dp_unique_id")
                box = [bytes(buf.raw)]
            dist.broadcast_object_list(box, src=0)
            buf = ctypes.create_string_buffer(box[0], 128)
        else:
            _lib.check(L.ftk_dp_unique_id(buf), "ftk_dp_unique_id")
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(L.ftk_dp_init(buf, self.rank, self.world, ctypes.byref(handle)), "ftk_dp_init")
        self._handle = handle

    # ------------------------------------------------------------------ hooks
    def _on_ready(self, p: torch.nn.Parameter) -> None:
        if self.main_grads and p.grad is not None:
            p.main_grad.add_(p.grad)  # bf16 micro-batch gradient into the fp32 main gradient
            p.grad = None
        self._note_ready(p, None)

    def _note_ready(self, p: torch.nn.Parameter, after: Optional[torch.cuda.Event]) -> None:
        if not self.sync_now:
            return
        b = self._bucket_of[p]
        if b.launched or id(p) in b.ready:
            return  # e.g. mark_ready() earlier in the step, then the autograd hook (it fires even for a None gradient)
        b.ready.add(id(p))
        if len(b.ready) == len(b.params):
            self._launch(b, after)

    def mark_ready(self, p: torch.nn.Parameter, after: Optional[torch.cuda.Event] = None) -> None:
        """Tell the sync that ``p``'s gradient is final although no autograd hook will fire for it: the fused lm_head
        loss writes an untied ``lm_head.weight`` gradient into ``main_grad`` (or ``grad``) itself.  If that completes
        ``p``'s bucket, its all-reduce is launched now, and with ``after`` (the event given to
        ``ops.set_dw_done_event``) it waits for that event only, not for what the compute stream was given since:
        the transfer then runs beside the dH GEMM.  Ignored inside ``no_sync()``.  Only for parameters that receive
        no further gradient in this step (not a tied ``wte``: the embedding's backward still adds to it)."""
        self._note_ready(p, after)

    def _launch(self, b: _Bucket, after: Optional[torch.cuda.Event] = None) -> None:
        scale = (1.0 / self.world) if self.average else 1.0
        self.launched += 1
        b.launched = True
        if self.backend == "ftk":
            dt = _lib.FTK_F32 if b.flat.dtype == torch.float32 else _lib.FTK_BF16
            if b.flat.dtype not in (torch.float32, torch.bfloat16):
                raise _lib.FtkError(f"unsupported gradient dtype {b.flat.dtype}")
            stream = torch.cuda.current_stream(self.device).cuda_stream
            with torch.cuda.device(self.device):
                if after is not None:
                    _lib.check(_lib.lib().ftk_dp_allreduce_after(self._handle, b.flat.data_ptr(), b.flat.numel(), dt,
                                                                 scale, int(after.cuda_event)),
                               "ftk_dp_allreduce_after")
                else:
                    _lib.check(_lib.lib().ftk_dp_allreduce(self._handle, b.flat.data_ptr(), b.flat.numel(), dt, scale,
                                                           stream), "ftk_dp_allreduce")
        else:
            if scale != 1.0:
                b.flat.mul_(scale)
            b.work = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, async_op=True) if self.world > 1 else None

    # ------------------------------------------------------------------ public API
    @contextlib.contextmanager
    def no_sync(self):
        """Gradient accumulation: micro-batches inside this context only accumulate locally."""
        old, self.sync_now = self.sync_now, False
        try:
            yield
        finally:
            self.sync_now = old

    def finish(self, reduce: bool = True) -> None:
        """Call after the last backward of an optimizer step: reduces any bucket whose hooks did not all fire
        (parameters without gradient this step) and makes the current stream wait for the communication stream.
        ``reduce=False`` only re-arms the buckets (a step measured without its communication)."""
        for b in self.buckets:
            if not b.launched and reduce:
                self._launch(b)
        if self.backend == "ftk":
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(_lib.lib().ftk_dp_join(self._handle, stream), "ftk_dp_join")
        else:
            for b in self.buckets:
                if b.work is not None:
                    b.work.wait()
                    b.work = None
        for b in self.buckets:
            b.ready.clear()
            b.launched = False

    def zero_grad(self, overwritten_by_fused_loss: Optional[torch.nn.Parameter] = None) -> None:
        """Zero the flat buckets (keeps every ``param.grad`` a view; do not use ``set_to_none=True``).
        ``overwritten_by_fused_loss``: the lm_head weight whose gradient the fused loss writes first in every step
        (before the embedding's backward of a tied ``wte``, before anything else touches it).  Its part of the bucket
        is not zeroed; the gradient tensor is marked so that the next fused call overwrites it instead of adding
        (412 MB less to write and 412 MB less to read per step for a 103 M-element fp32 main gradient)."""
        skip = overwritten_by_fused_loss
        for b in self.buckets:
            if skip is not None and any(p is skip for p in b.params):
                tgt = skip.main_grad if self.main_grads else skip.grad
                lo = tgt.data_ptr() - b.flat.data_ptr()
                lo //= b.flat.element_size()
                hi = lo + tgt.numel()
                if lo > 0:
                    b.flat[:lo].zero_()
                if hi < b.flat.numel():
                    b.flat[hi:].zero_()
                tgt._ftk_overwrite = True
            else:
                b.flat.zero_()

    def close(self) -> None:
        if self._handle is not None:
            _lib.lib().ftk_dp_destroy(self._handle)
            self._handle = None
