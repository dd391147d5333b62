"""Device-side optimizer step over the gradient buckets.

Stands where DeepSpeed's AdamW + WarmupLR + gradient clipping stood for the reference (finetuner/ds_config.json:10-26,
:44; hyper-parameters from TrainingArguments at FT.py:987-998: lr = --lr, weight_decay 0.01, warmup_ratio).  The moments
(and, for bf16 parameters, an fp32 master copy) live in flat buffers parallel to GradSync's gradient buckets, so one
kernel launch updates a whole bucket right after its all-reduce; nothing crosses PCIe.
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch

from . import _lib
from .dp import GradSync


class WarmupLR:
    """DeepSpeed's WarmupLR (the scheduler named at ds_config.json:19-26): lr = min + (max - min) * gamma with
    gamma = log(it + 1) / log(warmup_steps) ("log", DeepSpeed's default) or it / warmup_steps ("linear") while
    it < warmup_steps, then 1.  ``it`` is the number of optimizer steps taken so far."""

    def __init__(self, max_lr: float, warmup_steps: int, min_lr: float = 0.0, warmup_type: str = "log"):
        if warmup_type not in ("log", "linear"):
            raise ValueError(warmup_type)
        self.max_lr, self.min_lr, self.warmup_type = max_lr, min_lr, warmup_type
        self.warmup_steps = max(2, int(warmup_steps))
        self._inv_log = 1.0 / math.log(self.warmup_steps)

    def lr_at(self, it: int) -> float:
        if it < self.warmup_steps:
            gamma = self._inv_log * math.log(it + 1) if self.warmup_type == "log" else it / self.warmup_steps
        else:
            gamma = 1.0
        return self.min_lr + (self.max_lr - self.min_lr) * gamma


class FusedAdamW:
    """AdamW over ``GradSync`` buckets.  ``step()`` = [global grad-norm -> clip coefficient] -> one fused kernel per
    bucket.  Parameters of a bucket are packed into a flat buffer once (each ``param.data`` becomes a view)."""

    def __init__(self, sync: GradSync, lr: float = 5e-5, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, max_grad_norm: Optional[float] = 1.0,
                 schedule: Optional[WarmupLR] = None):
        if sync.device.type != "cuda":
            raise _lib.FtkError("FusedAdamW runs on CUDA buckets only")
        self.sync, self.lr, self.betas, self.eps, self.wd = sync, lr, betas, eps, weight_decay
        self.max_grad_norm, self.schedule = max_grad_norm, schedule
        self.steps = 0
        dev = sync.device
        self.flat_params: List[torch.Tensor] = []
        self.master: List[Optional[torch.Tensor]] = []
        self.m: List[torch.Tensor] = []
        self.v: List[torch.Tensor] = []
        for b in sync.buckets:
            n = b.flat.numel()
            pdt = b.params[0].dtype
            if pdt not in (torch.float32, torch.bfloat16) or any(p.dtype != pdt for p in b.params):
                raise _lib.FtkError("parameters of a bucket must all be fp32 or all bf16")
            if b.flat.dtype not in (torch.float32, torch.bfloat16) or (pdt == torch.float32
                                                                        and b.flat.dtype != torch.float32):
                raise _lib.FtkError("gradient buckets must be fp32, or bf16 for bf16 parameters")
            flat = torch.zeros(n, dtype=pdt, device=dev)
            off = 0
            for p in b.params:  # same packing as GradSync: 8-element aligned views
                flat[off:off + p.numel()].copy_(p.data.reshape(-1))
                p.data = flat[off:off + p.numel()].view_as(p)
                off += (p.numel() + 7) // 8 * 8
            self.flat_params.append(flat)
            self.master.append(flat.float() if pdt == torch.bfloat16 else None)
            self.m.append(torch.zeros(n, dtype=torch.float32, device=dev))
            self.v.append(torch.zeros(n, dtype=torch.float32, device=dev))
        L = _lib.lib()
        self._ws = torch.empty(L.ftk_grad_sqnorm_workspace(), dtype=torch.uint8, device=dev)
        self._sqnorm = torch.zeros(1, dtype=torch.float32, device=dev)
        self._coef = torch.ones(1, dtype=torch.float32, device=dev)

    def current_lr(self) -> float:
        return self.schedule.lr_at(self.steps) if self.schedule is not None else self.lr

    def step(self) -> None:
        """Call after ``sync.finish()``: the buckets then hold the averaged gradients on the current stream."""
        L = _lib.lib()
        dev = self.sync.device
        s = torch.cuda.current_stream(dev).cuda_stream
        code = lambda dt: _lib.FTK_F32 if dt == torch.float32 else _lib.FTK_BF16  # noqa: E731
        lr = self.current_lr()
        self.steps += 1
        with torch.cuda.device(dev):
            clip_ptr = 0
            if self.max_grad_norm is not None:
                for i, b in enumerate(self.sync.buckets):
                    _lib.check(L.ftk_grad_sqnorm(b.flat.data_ptr(), b.flat.numel(), code(b.flat.dtype),
                                                 self._sqnorm.data_ptr(), int(i > 0), self._ws.data_ptr(),
                                                 self._ws.numel(), s), "ftk_grad_sqnorm")
                _lib.check(L.ftk_clip_coef(self._sqnorm.data_ptr(), float(self.max_grad_norm), self._coef.data_ptr(), s),
                           "ftk_clip_coef")
                clip_ptr = self._coef.data_ptr()
            for i, b in enumerate(self.sync.buckets):
                fp = self.flat_params[i]
                _lib.check(L.ftk_adamw_step(fp.data_ptr(), b.flat.data_ptr(),
                                            0 if self.master[i] is None else self.master[i].data_ptr(),
                                            self.m[i].data_ptr(), self.v[i].data_ptr(), fp.numel(), code(fp.dtype),
                                            code(b.flat.dtype), lr, self.betas[0], self.betas[1], self.eps, self.wd,
                                            self.steps, 1.0, clip_ptr, s), "ftk_adamw_step")

    def grad_norm(self) -> torch.Tensor:
        """The global gradient norm measured by the last ``step()`` (device scalar; reading it synchronises)."""
        return self._sqnorm.sqrt()
