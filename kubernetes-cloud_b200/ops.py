"""PyTorch custom ops over the C ABI (include/ftk.h).  PyTorch is plumbing here: it owns device memory and the
current stream; all arithmetic happens inside lib/libftk.so.  No op has a CPU or eager fallback.

  torch.ops.ftk.gather_windows   -> ftk_gather_windows   (replaces FT.py:662-692, :1030-1035, :476-477)
  torch.ops.ftk.lmhead_ce_fwd    -> ftk_lmhead_ce_fwd    (replaces HF lm_head + ForCausalLMLoss, FT.py:479)
  torch.ops.ftk.lmhead_ce_bwd    -> ftk_lmhead_ce_bwd    (their autograd)
  fused_lmhead_ce(hidden, weight, bias, targets) -> mean loss, differentiable w.r.t. hidden / weight / bias
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib

IGNORE = -100
_launches = 0  # kernels launched through this module (bench.py reports it as gpu_launches)


def launch_count() -> int:
    return _launches


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _need_cuda(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.FtkError("ftk ops run on CUDA tensors only (there is no CPU fallback)")


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


@torch.library.custom_op("ftk::gather_windows", mutates_args=())
def gather_windows(tokens: torch.Tensor, ctx_idx: torch.Tensor, T: int, pad_id: int, pad_is_eos: bool,
                   shift_labels: bool) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """tokens: flat uint16 (or int16 view) token stream on the GPU; ctx_idx: int64 [B] on the GPU.
    Returns input_ids int64 [B,T], labels int64 [B,T] (-100 where masked; shifted if asked), attention_mask
    bool [B,T], n_valid int32 [B]."""
    global _launches
    _need_cuda(tokens, ctx_idx)
    if tokens.element_size() != 2 or not tokens.is_contiguous():
        raise _lib.FtkError("tokens must be a contiguous 16-bit tensor")
    if ctx_idx.dtype != torch.int64 or not ctx_idx.is_contiguous():
        raise _lib.FtkError("ctx_idx must be a contiguous int64 tensor")
    n_ctx = tokens.numel() // T
    B = ctx_idx.numel()
    dev = tokens.device
    ids = torch.empty((B, T), dtype=torch.int64, device=dev)
    labels = torch.empty((B, T), dtype=torch.int64, device=dev)
    mask = torch.empty((B, T), dtype=torch.bool, device=dev)
    n_valid = torch.empty((B,), dtype=torch.int32, device=dev)
    if B == 0:
        return ids, labels, mask, n_valid
    with torch.cuda.device(dev):
        rc = _lib.lib().ftk_gather_windows(tokens.data_ptr(), n_ctx, T, ctx_idx.data_ptr(), B, pad_id,
                                           int(pad_is_eos), int(shift_labels), ids.data_ptr(), labels.data_ptr(),
                                           mask.data_ptr(), n_valid.data_ptr(), _stream(tokens))
    _lib.check(rc, "ftk_gather_windows")
    _launches += 1
    return ids, labels, mask, n_valid


@gather_windows.register_fake
def _(tokens, ctx_idx, T, pad_id, pad_is_eos, shift_labels):
    B = ctx_idx.numel()
    return (tokens.new_empty((B, T), dtype=torch.int64), tokens.new_empty((B, T), dtype=torch.int64),
            tokens.new_empty((B, T), dtype=torch.bool), tokens.new_empty((B,), dtype=torch.int32))


def _check_hw(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], targets: torch.Tensor) -> None:
    _need_cuda(H, W, bias, targets)
    if H.dtype != torch.bfloat16 or W.dtype != torch.bfloat16:
        raise _lib.FtkError("hidden and weight must be bfloat16")
    if bias is not None and bias.dtype != torch.bfloat16:
        raise _lib.FtkError("bias must be bfloat16")
    if H.dim() != 2 or W.dim() != 2 or H.shape[1] != W.shape[1]:
        raise _lib.FtkError(f"bad shapes hidden {tuple(H.shape)} weight {tuple(W.shape)}")
    if not (H.is_contiguous() and W.is_contiguous() and targets.is_contiguous()):
        raise _lib.FtkError("hidden, weight and targets must be contiguous")
    if targets.dtype != torch.int64 or targets.numel() != H.shape[0]:
        raise _lib.FtkError("targets must be int64 with one entry per hidden row")


@torch.library.custom_op("ftk::lmhead_ce_fwd", mutates_args=())
def lmhead_ce_fwd(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor],
                  targets: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Returns (loss_sum fp32 [1], n_valid int32 [1], lse fp32 [M]) for logits = H @ W.T (+ bias) without ever
    writing the logits to HBM."""
    global _launches
    _check_hw(H, W, bias, targets)
    M, d = H.shape
    V = W.shape[0]
    dev = H.device
    lse = torch.empty((M,), dtype=torch.float32, device=dev)
    loss_sum = torch.empty((1,), dtype=torch.float32, device=dev)
    n_valid = torch.empty((1,), dtype=torch.int32, device=dev)
    L = _lib.lib()
    ws_bytes = L.ftk_lmhead_ce_fwd_workspace(M, d, V)
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = L.ftk_lmhead_ce_fwd(H.data_ptr(), W.data_ptr(), _ptr(bias), targets.data_ptr(), M, d, V, lse.data_ptr(),
                                 loss_sum.data_ptr(), n_valid.data_ptr(), ws.data_ptr(), ws_bytes, _stream(H))
    _lib.check(rc, "ftk_lmhead_ce_fwd")
    _launches += 2  # fused GEMM+softmax kernel, finalize kernel
    return loss_sum, n_valid, lse


@lmhead_ce_fwd.register_fake
def _(H, W, bias, targets):
    return (H.new_empty((1,), dtype=torch.float32), H.new_empty((1,), dtype=torch.int32),
            H.new_empty((H.shape[0],), dtype=torch.float32))


_slab_rows = 0  # 0 = library default; tests shrink it to exercise the multi-slab path


def set_bwd_slab_rows(rows: int) -> None:
    global _slab_rows
    _slab_rows = int(rows)


def _dw_target(W: torch.Tensor, dw_fp32: bool, dW_into: Optional[torch.Tensor]):
    """(tensor the dW GEMM writes, dtype code, accumulate flag).  A target marked ``_ftk_overwrite`` (GradSync's lazy
    ``zero_grad``) is overwritten instead of added to, once: the zeroing pass and the first read of it are saved."""
    if dW_into is None:
        dW = torch.empty(W.shape, dtype=torch.float32 if dw_fp32 else torch.bfloat16, device=W.device)
        return dW, (_lib.FTK_F32 if dw_fp32 else _lib.FTK_BF16), 0
    _need_cuda(dW_into)
    if dW_into.shape != W.shape or not dW_into.is_contiguous() or dW_into.dtype not in (torch.float32, torch.bfloat16):
        raise _lib.FtkError("dW_into must be a contiguous bf16 or fp32 tensor of the weight's shape")
    acc = 1
    if getattr(dW_into, "_ftk_overwrite", False):
        dW_into._ftk_overwrite = False
        acc = 0
    return dW_into, (_lib.FTK_F32 if dW_into.dtype == torch.float32 else _lib.FTK_BF16), acc


def _bwd_impl(H, W, bias, targets, lse, grad_out, n_valid, grad_scale, dw_fp32, dW_into):
    global _launches
    _check_hw(H, W, bias, targets)
    _need_cuda(lse, grad_out, n_valid)
    M, d = H.shape
    V = W.shape[0]
    dev = H.device
    dH = torch.empty((M, d), dtype=torch.bfloat16, device=dev)
    dW, dw_code, acc = _dw_target(W, dw_fp32, dW_into)
    dbias = torch.empty((V if bias is not None else 0,), dtype=torch.float32, device=dev)
    if grad_out is not None:
        grad_out = grad_out.to(torch.float32).contiguous()
    L = _lib.lib()
    ws_bytes = L.ftk_lmhead_ce_bwd_workspace(M, d, V, _slab_rows)
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = L.ftk_lmhead_ce_bwd(H.data_ptr(), W.data_ptr(), _ptr(bias), targets.data_ptr(), lse.data_ptr(), M, d, V,
                                 grad_scale, _ptr(grad_out), _ptr(n_valid), dH.data_ptr(), dW.data_ptr(), dw_code, acc,
                                 dbias.data_ptr() if bias is not None else 0, ws.data_ptr(), ws_bytes, _stream(H))
    _lib.check(rc, "ftk_lmhead_ce_bwd")
    slab = _slab_rows if _slab_rows > 0 else 16384
    n_slabs = (M + slab - 1) // slab
    _launches += n_slabs * (3 + (1 if bias is not None else 0))
    return dH, dW, dbias


@torch.library.custom_op("ftk::lmhead_ce_bwd", mutates_args=())
def lmhead_ce_bwd(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], targets: torch.Tensor,
                  lse: torch.Tensor, grad_out: Optional[torch.Tensor], n_valid: Optional[torch.Tensor],
                  grad_scale: float, dw_fp32: bool) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Returns (dH bf16 [M,d], dW [V,d] bf16 or fp32, dbias fp32 [V] or empty).
    dZ = (softmax - onehot) * grad_scale * grad_out / n_valid on valid rows."""
    return _bwd_impl(H, W, bias, targets, lse, grad_out, n_valid, grad_scale, dw_fp32, None)


@lmhead_ce_bwd.register_fake
def _(H, W, bias, targets, lse, grad_out, n_valid, grad_scale, dw_fp32):
    return (torch.empty_like(H), W.new_empty(W.shape, dtype=torch.float32 if dw_fp32 else torch.bfloat16),
            H.new_empty((W.shape[0] if bias is not None else 0,), dtype=torch.float32))


@torch.library.custom_op("ftk::lmhead_ce_bwd_into", mutates_args=("dW_into",))
def lmhead_ce_bwd_into(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], targets: torch.Tensor,
                       lse: torch.Tensor, grad_out: Optional[torch.Tensor], n_valid: Optional[torch.Tensor],
                       grad_scale: float, dW_into: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """The same backward with the dW GEMM adding into ``dW_into`` (bf16 or fp32 [V,d]) inside its epilogue: tied
    embeddings, gradient accumulation, an fp32 main gradient.  Returns (dH, dbias)."""
    dH, _, dbias = _bwd_impl(H, W, bias, targets, lse, grad_out, n_valid, grad_scale, False, dW_into)
    return dH, dbias


@lmhead_ce_bwd_into.register_fake
def _(H, W, bias, targets, lse, grad_out, n_valid, grad_scale, dW_into):
    return torch.empty_like(H), H.new_empty((W.shape[0] if bias is not None else 0,), dtype=torch.float32)


@torch.library.custom_op("ftk::lmhead_ce_fwd_keep", mutates_args=())
def lmhead_ce_fwd_keep(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], targets: torch.Tensor
                       ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """The forward that also keeps the softmax numerators: returns (loss_sum, n_valid, lse, probs bf16 [M, ld],
    cref fp32 [n_tiles, Mpad]) for ``lmhead_ce_bwd_kept``."""
    global _launches
    _check_hw(H, W, bias, targets)
    M, d = H.shape
    V = W.shape[0]
    dev = H.device
    L = _lib.lib()
    lse = torch.empty((M,), dtype=torch.float32, device=dev)
    loss_sum = torch.empty((1,), dtype=torch.float32, device=dev)
    n_valid = torch.empty((1,), dtype=torch.int32, device=dev)
    ld = (V + 63) // 64 * 64
    probs = torch.empty((M, ld), dtype=torch.bfloat16, device=dev)
    cref = torch.empty((L.ftk_lmhead_ce_cref_bytes(M, V) // 4,), dtype=torch.float32, device=dev)
    ws_bytes = L.ftk_lmhead_ce_fwd_workspace(M, d, V)
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = L.ftk_lmhead_ce_fwd_keep(H.data_ptr(), W.data_ptr(), _ptr(bias), targets.data_ptr(), M, d, V,
                                      lse.data_ptr(), loss_sum.data_ptr(), n_valid.data_ptr(), probs.data_ptr(),
                                      cref.data_ptr(), ws.data_ptr(), ws_bytes, _stream(H))
    _lib.check(rc, "ftk_lmhead_ce_fwd_keep")
    _launches += 2
    return loss_sum, n_valid, lse, probs, cref


@lmhead_ce_fwd_keep.register_fake
def _(H, W, bias, targets):
    M, V = H.shape[0], W.shape[0]
    return (H.new_empty((1,), dtype=torch.float32), H.new_empty((1,), dtype=torch.int32),
            H.new_empty((M,), dtype=torch.float32), H.new_empty((M, (V + 63) // 64 * 64)),
            H.new_empty(((V + 255) // 256 * ((M + 255) // 256 * 256),), dtype=torch.float32))


def _bwd_kept_impl(H, W, has_bias, targets, lse, probs, cref, grad_out, n_valid, grad_scale, dw_fp32, dW_into):
    global _launches
    _check_hw(H, W, None, targets)
    _need_cuda(lse, probs, cref, grad_out, n_valid)
    M, d = H.shape
    V = W.shape[0]
    dev = H.device
    dH = torch.empty((M, d), dtype=torch.bfloat16, device=dev)
    dW, dw_code, acc = _dw_target(W, dw_fp32, dW_into)
    dbias = torch.empty((V if has_bias else 0,), dtype=torch.float32, device=dev)
    if grad_out is not None:
        grad_out = grad_out.to(torch.float32).contiguous()
    with torch.cuda.device(dev):
        rc = _lib.lib().ftk_lmhead_ce_bwd_kept(H.data_ptr(), W.data_ptr(), targets.data_ptr(), lse.data_ptr(),
                                               probs.data_ptr(), cref.data_ptr(), M, d, V, grad_scale, _ptr(grad_out),
                                               _ptr(n_valid), dH.data_ptr(), dW.data_ptr(), dw_code, acc,
                                               dbias.data_ptr() if has_bias else 0, _stream(H))
    _lib.check(rc, "ftk_lmhead_ce_bwd_kept")
    _launches += 3 + (1 if has_bias else 0)
    return dH, dW, dbias


@torch.library.custom_op("ftk::lmhead_ce_bwd_kept", mutates_args=("probs",))
def lmhead_ce_bwd_kept(H: torch.Tensor, W: torch.Tensor, has_bias: bool, targets: torch.Tensor, lse: torch.Tensor,
                       probs: torch.Tensor, cref: torch.Tensor, grad_out: Optional[torch.Tensor],
                       n_valid: Optional[torch.Tensor], grad_scale: float, dw_fp32: bool
                       ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Backward from kept numerators (no logit recompute).  ``probs`` is overwritten with dZ."""
    return _bwd_kept_impl(H, W, has_bias, targets, lse, probs, cref, grad_out, n_valid, grad_scale, dw_fp32, None)


@lmhead_ce_bwd_kept.register_fake
def _(H, W, has_bias, targets, lse, probs, cref, grad_out, n_valid, grad_scale, dw_fp32):
    return (torch.empty_like(H), W.new_empty(W.shape, dtype=torch.float32 if dw_fp32 else torch.bfloat16),
            H.new_empty((W.shape[0] if has_bias else 0,), dtype=torch.float32))


@torch.library.custom_op("ftk::lmhead_ce_bwd_kept_into", mutates_args=("probs", "dW_into"))
def lmhead_ce_bwd_kept_into(H: torch.Tensor, W: torch.Tensor, has_bias: bool, targets: torch.Tensor,
                            lse: torch.Tensor, probs: torch.Tensor, cref: torch.Tensor,
                            grad_out: Optional[torch.Tensor], n_valid: Optional[torch.Tensor], grad_scale: float,
                            dW_into: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``lmhead_ce_bwd_kept`` with the dW GEMM adding into ``dW_into``.  Returns (dH, dbias)."""
    dH, _, dbias = _bwd_kept_impl(H, W, has_bias, targets, lse, probs, cref, grad_out, n_valid, grad_scale, False,
                                  dW_into)
    return dH, dbias


@lmhead_ce_bwd_kept_into.register_fake
def _(H, W, has_bias, targets, lse, probs, cref, grad_out, n_valid, grad_scale, dW_into):
    return torch.empty_like(H), H.new_empty((W.shape[0] if has_bias else 0,), dtype=torch.float32)


_fused_slab_rows = 0  # 0 = the library's choice; tests shrink it to exercise several slabs


def set_fused_slab_rows(rows: int) -> None:
    global _fused_slab_rows
    _fused_slab_rows = int(rows)


def _fused_impl(H, W, bias, targets, grad_scale, dw_fp32, dw_event, dW_into):
    global _launches
    _check_hw(H, W, bias, targets)
    M, d = H.shape
    V = W.shape[0]
    dev = H.device
    L = _lib.lib()
    lse = torch.empty((M,), dtype=torch.float32, device=dev)
    loss_sum = torch.empty((1,), dtype=torch.float32, device=dev)
    n_valid = torch.empty((1,), dtype=torch.int32, device=dev)
    dH = torch.empty((M, d), dtype=torch.bfloat16, device=dev)
    dW, dw_code, acc = _dw_target(W, dw_fp32, dW_into)
    dbias = torch.empty((V if bias is not None else 0,), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        ws_bytes = L.ftk_lmhead_ce_fused_workspace(M, d, V, _fused_slab_rows)
        slab = L.ftk_lmhead_ce_fused_slab_rows(M, d, V, _fused_slab_rows)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        rc = L.ftk_lmhead_ce_fused(H.data_ptr(), W.data_ptr(), _ptr(bias), targets.data_ptr(), M, d, V, grad_scale, 0,
                                   lse.data_ptr(), loss_sum.data_ptr(), n_valid.data_ptr(), dH.data_ptr(),
                                   dW.data_ptr(), dw_code, acc, dbias.data_ptr() if bias is not None else 0,
                                   int(dw_event), ws.data_ptr(), ws_bytes, _fused_slab_rows, _stream(H))
    _lib.check(rc, "ftk_lmhead_ce_fused")
    n_slabs = (M + slab - 1) // slab
    _launches += 1 + n_slabs * (5 + (1 if bias is not None else 0))  # count; per slab: sweep, finalize, dZ, dH, dW
    return loss_sum, n_valid, lse, dH, dW, dbias


@torch.library.custom_op("ftk::lmhead_ce_fused", mutates_args=())
def lmhead_ce_fused(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], targets: torch.Tensor,
                    grad_scale: float, dw_fp32: bool, dw_event: int
                    ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """Forward and backward in one call (ftk_lmhead_ce_fused): returns (loss_sum, n_valid, lse, dH, dW, dbias) with
    the gradients of ``grad_scale * mean loss``.  ``dw_event``: 0, or the handle of a CUDA event to record once dW
    is complete (the last slab then computes dW before dH)."""
    return _fused_impl(H, W, bias, targets, grad_scale, dw_fp32, dw_event, None)


@lmhead_ce_fused.register_fake
def _(H, W, bias, targets, grad_scale, dw_fp32, dw_event):
    return (H.new_empty((1,), dtype=torch.float32), H.new_empty((1,), dtype=torch.int32),
            H.new_empty((H.shape[0],), dtype=torch.float32), torch.empty_like(H),
            W.new_empty(W.shape, dtype=torch.float32 if dw_fp32 else torch.bfloat16),
            H.new_empty((W.shape[0] if bias is not None else 0,), dtype=torch.float32))


@torch.library.custom_op("ftk::lmhead_ce_fused_into", mutates_args=("dW_into",))
def lmhead_ce_fused_into(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], targets: torch.Tensor,
                         grad_scale: float, dw_event: int, dW_into: torch.Tensor
                         ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """``lmhead_ce_fused`` with the dW GEMM adding into ``dW_into``.  Returns (loss_sum, n_valid, lse, dH, dbias)."""
    loss_sum, n_valid, lse, dH, _, dbias = _fused_impl(H, W, bias, targets, grad_scale, False, dw_event, dW_into)
    return loss_sum, n_valid, lse, dH, dbias


@lmhead_ce_fused_into.register_fake
def _(H, W, bias, targets, grad_scale, dw_event, dW_into):
    return (H.new_empty((1,), dtype=torch.float32), H.new_empty((1,), dtype=torch.int32),
            H.new_empty((H.shape[0],), dtype=torch.float32), torch.empty_like(H),
            H.new_empty((W.shape[0] if bias is not None else 0,), dtype=torch.float32))


def _scale_unless_one(t: torch.Tensor, factor: torch.Tensor, poison: bool = False) -> None:
    """t *= factor (a device scalar) unless factor == 1, decided on the device: no host synchronisation."""
    global _launches
    if t.numel() == 0:
        return
    code = _lib.FTK_F32 if t.dtype == torch.float32 else _lib.FTK_BF16
    per = 4 if t.dtype == torch.float32 else 8
    flat = t.view(-1)
    n = flat.numel() // per * per
    with torch.cuda.device(t.device):
        _lib.check(_lib.lib().ftk_scale_unless_one(flat.data_ptr(), n, code, factor.data_ptr(), int(poison), _stream(t)),
                   "ftk_scale_unless_one")
    _launches += 1
    if n != flat.numel():  # a tail shorter than one 16-byte vector (odd vocabulary sizes of dbias)
        flat[n:].mul_(torch.where(factor == 1, factor, factor * (float("nan") if poison else 1.0)))


_backward_mode = "fused"
_keep_budget_bytes = 8 << 30
_dw_done_event: Optional[torch.cuda.Event] = None


def set_backward_mode(mode: str, keep_budget_bytes: int = 8 << 30) -> None:
    """How ``fused_lmhead_ce`` obtains its gradients.
    "fused" (default): when both hidden and weight need gradients, the forward call computes them as well
    (ftk_lmhead_ce_fused), slab by slab, so the logits are formed once and nothing of size [tokens, vocab] outlives a
    slab; ``backward`` only rescales by the upstream gradient.  dH and dW exist from forward to backward.
    "recompute": forward keeps only lse; the backward recomputes the logits tile by tile (a third more tensor work).
    "keep": the forward stores ALL bf16 softmax numerators (tokens x vocab x 2 bytes, used only while they fit
    ``keep_budget_bytes``) and the backward skips the recompute GEMM."""
    global _backward_mode, _keep_budget_bytes
    if mode not in ("fused", "recompute", "keep"):
        raise ValueError(mode)
    _backward_mode, _keep_budget_bytes = mode, int(keep_budget_bytes)


def set_dw_done_event(event: Optional[torch.cuda.Event]) -> None:
    """Data-parallel callers ("fused" mode): an event ``fused_lmhead_ce`` records as soon as the weight gradient is
    complete.  With one set, the last slab computes dW before dH and runs that dH GEMM on the SMs the
    ``sm_reserve`` option leaves to others, so an all-reduce waiting on the event (``GradSync.mark_ready``) overlaps
    it.  None switches this off.  The event must have been recorded once already (its handle exists only then)."""
    global _dw_done_event
    _dw_done_event = event


def _dw_event_handle() -> int:
    return 0 if _dw_done_event is None else int(_dw_done_event.cuda_event)


def _wgrad_target(weight: torch.Tensor) -> Optional[torch.Tensor]:
    """Where the dW GEMM can add its result directly instead of handing autograd a fresh [V, d] tensor to add
    (206 MB read + add + write per micro-batch at the GPT-Neo-1.3B head):
      * ``weight.main_grad`` -- an fp32 (or bf16) main-gradient buffer the caller attached to the parameter
        (``GradSync`` does; the convention of Megatron's fused weight-gradient accumulation).  ``weight.grad`` is then
        left alone and no post-accumulate hook fires for this contribution;
      * otherwise an existing, contiguous ``weight.grad`` of the parameter's shape in bf16 or fp32 (the second and
        later micro-batches of a gradient-accumulation step, or a tied ``wte`` whose gradient already exists).
    None: the first contribution, returned to autograd as usual."""
    mg = getattr(weight, "main_grad", None)
    if mg is not None:
        return mg
    g = weight.grad
    if (g is not None and g.is_cuda and g.shape == weight.shape and g.is_contiguous()
            and g.dtype in (torch.bfloat16, torch.float32)):
        return g
    return None


class _FusedLmHeadCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hidden, weight, bias, targets, loss_scale):
        M, V = hidden.shape[0], weight.shape[0]
        want_h, want_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]  # grad mode is off inside forward
        ctx.mode = "none"
        ctx.loss_scale = float(loss_scale)
        ctx.consumed = False
        if _backward_mode == "fused" and want_h and want_w:
            ctx.mode = "fused"
            into = _wgrad_target(weight)
            if into is not None:
                loss_sum, n_valid, lse, dH, dbias = torch.ops.ftk.lmhead_ce_fused_into(hidden, weight, bias, targets,
                                                                                       ctx.loss_scale, _dw_event_handle(), into)
                dW = None
            else:
                loss_sum, n_valid, lse, dH, dW, dbias = torch.ops.ftk.lmhead_ce_fused(hidden, weight, bias, targets,
                                                                                      ctx.loss_scale, False, _dw_event_handle())
            ctx.grads = (dH, dW, dbias if bias is not None else None, into is not None)
            ctx.bias_dtype = bias.dtype if bias is not None else None
        elif (_backward_mode == "keep" and (want_h or want_w)
              and _lib.lib().ftk_lmhead_ce_probs_bytes(M, V) <= _keep_budget_bytes):
            ctx.mode = "keep"
            loss_sum, n_valid, lse, probs, cref = torch.ops.ftk.lmhead_ce_fwd_keep(hidden, weight, bias, targets)
            ctx.save_for_backward(hidden, weight, bias, targets, lse, n_valid, probs, cref)
        else:
            if want_h or want_w:
                ctx.mode = "recompute"
            loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(hidden, weight, bias, targets)
            ctx.save_for_backward(hidden, weight, bias, targets, lse, n_valid)
        # mean over valid rows (transformers fixed_cross_entropy, reduction="mean"); 0/0 -> NaN as in the reference
        loss = (loss_sum / n_valid.to(torch.float32)).reshape(())
        return loss * ctx.loss_scale if ctx.loss_scale != 1.0 else loss

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.consumed:
            raise _lib.FtkError("fused_lmhead_ce: this loss was already backpropagated (gradients / kept numerators "
                                "are consumed by the first backward; use set_backward_mode('recompute') to do it twice)")
        go = grad_out.reshape(1).to(torch.float32)
        if ctx.mode == "fused":
            # the gradients of loss_scale * mean-loss exist already; all that is left is the upstream factor, which
            # is 1 for ``loss.backward()`` -- checked on the device, so the common case costs three empty launches
            dH, dW, dbias, accumulated = ctx.grads
            ctx.grads = None
            ctx.consumed = True
            if accumulated:
                # dW already sits inside the caller's gradient buffer and cannot be rescaled on its own: make a wrong
                # assumption visible (NaN in dH) rather than return a silently mis-scaled weight gradient
                _scale_unless_one(dH, go, poison=True)
            else:
                _scale_unless_one(dH, go)
                _scale_unless_one(dW, go)
            if dbias is not None:
                _scale_unless_one(dbias, go)
            return dH, dW, (dbias.to(ctx.bias_dtype) if dbias is not None else None), None, None
        weight = ctx.saved_tensors[1]
        into = _wgrad_target(weight) if ctx.needs_input_grad[1] else None
        dW = None
        if ctx.mode == "keep":
            hidden, weight, bias, targets, lse, n_valid, probs, cref = ctx.saved_tensors
            if into is not None:
                dH, dbias = torch.ops.ftk.lmhead_ce_bwd_kept_into(hidden, weight, bias is not None, targets, lse, probs,
                                                                  cref, go, n_valid, ctx.loss_scale, into)
            else:
                dH, dW, dbias = torch.ops.ftk.lmhead_ce_bwd_kept(hidden, weight, bias is not None, targets, lse, probs,
                                                                 cref, go, n_valid, ctx.loss_scale, False)
            ctx.consumed = True
        else:
            hidden, weight, bias, targets, lse, n_valid = ctx.saved_tensors
            if into is not None:
                dH, dbias = torch.ops.ftk.lmhead_ce_bwd_into(hidden, weight, bias, targets, lse, go, n_valid,
                                                             ctx.loss_scale, into)
            else:
                dH, dW, dbias = torch.ops.ftk.lmhead_ce_bwd(hidden, weight, bias, targets, lse, go, n_valid,
                                                            ctx.loss_scale, False)
        return dH, dW, (dbias.to(bias.dtype) if bias is not None else None), None, None


def fused_lmhead_ce(hidden: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor],
                    targets: torch.Tensor, loss_scale: float = 1.0) -> torch.Tensor:
    """``loss_scale`` x mean causal-LM cross-entropy of logits = hidden @ weight.T (+ bias) against already shifted
    targets (-100 = ignore).  hidden [M,d] bf16, weight [V,d] bf16.  Differentiable; the full logits never reach HBM
    (``set_backward_mode`` describes the three ways the gradients are obtained).

    ``loss_scale`` is where a training loop puts its 1 / gradient_accumulation_steps: in the default mode the
    gradients are produced during this call, already carrying that factor, and ``loss.backward()`` then finds an
    upstream gradient of exactly 1 and has nothing left to do.  Any other upstream gradient is still honoured by a
    rescaling pass, except for a weight gradient that was accumulated straight into ``weight.main_grad`` /
    ``weight.grad``: there it poisons dH with NaN instead of returning a wrong result."""
    return _FusedLmHeadCE.apply(hidden, weight, bias, targets, loss_scale)
