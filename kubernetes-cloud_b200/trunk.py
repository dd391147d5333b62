"""The transformer trunk on the B200 kernels: linear layers, MLP, attention, LayerNorm, whole blocks.

The reference trains whatever HF causal LM ``--model`` names through ``model(**inputs)`` (FT.py:479) with gradient
checkpointing on (FT.py:884).  The trunk is third-party code; this module swaps the pieces of it that the existing
tcgen05 kernel template covers, leaving the module tree, parameter names and checkpoints untouched:

  ``linear(x, weight, bias)``                  torch.nn.functional.linear            -> ftk_linear_fwd / ftk_linear_bwd
  ``mlp_gelu_new(x, w_fc, b_fc, w_proj, b_proj)``   c_proj(gelu_new(c_fc(x)))       -> two GEMMs; bias + GELU (and its saved
                                                derivative) live in the first one's epilogue, the multiplication by that
                                                derivative in the epilogue of the second
                                                one's dX GEMM, so no elementwise kernel ever touches the [tokens, 4d]
                                                activations
  ``patch_gpt_neo(model)``                     rebinds the forward of every GPTNeoMLP and of the attention projections

  ``attention(q, k, v, key_mask, heads, window, scale)``   HF's eager fp32 softmax(Q.K^T).V with its causal / sliding-
                                                window bias                           -> ftk_attention_fwd / _bwd (the [T, T]
                                                scores never leave TMEM)
  ``layer_norm(x, weight, bias, eps)``         torch.nn.functional.layer_norm          -> ftk_layernorm_fwd / _bwd
  ``patch_gpt_neo(model)`` also rebinds every GPTNeoBlock's forward to ``_block_forward`` below: the two residual adds
                                                of a block move into the epilogues of out_proj's and c_proj's GEMMs, the
                                                three dX GEMMs of q/k/v_proj chain into one sum, and the LayerNorm
                                                backward adds the residual branch's gradient itself, so a block runs
                                                no elementwise kernel at all, forward or backward.
No CPU or eager fallback: CPU tensors raise.
"""
from __future__ import annotations

import types
from typing import Optional, Tuple

import torch

from . import _lib
from .ops import _need_cuda, _ptr, _stream, _wgrad_target

ACT_NONE, ACT_GELU_NEW = 0, 1  # FTK_ACT_* in include/ftk.h
_launches = 0


def launch_count() -> int:
    return _launches


def _check_like(t: Optional[torch.Tensor], rows: int, cols: int, what: str) -> None:
    if t is not None and (t.dtype != torch.bfloat16 or t.shape != (rows, cols) or not t.is_contiguous() or not t.is_cuda):
        raise _lib.FtkError(f"{what} must be a contiguous bfloat16 CUDA [{rows}, {cols}] tensor")


def _check(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor]) -> Tuple[int, int, int]:
    _need_cuda(x, w, b)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or (b is not None and b.dtype != torch.bfloat16):
        raise _lib.FtkError("ftk linear layers take bfloat16 activations, weights and biases")
    if x.dim() != 2 or w.dim() != 2 or x.shape[1] != w.shape[1] or not x.is_contiguous() or not w.is_contiguous():
        raise _lib.FtkError(f"bad operands: x {tuple(x.shape)} weight {tuple(w.shape)} (contiguous 2-D, shared K)")
    M, K = x.shape
    N = w.shape[0]
    if K % 64 or N % 64:
        raise _lib.FtkError(f"K={K} and N={N} must be multiples of 64")
    return M, K, N


@torch.library.custom_op("ftk::linear_fwd", mutates_args=())
def linear_fwd(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], residual: Optional[torch.Tensor], act: int,
               save_pre: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """(y, dact): y = act(x @ w.T + b) + residual; dact = gelu_new'(x @ w.T + b) when ``save_pre`` (else empty)."""
    global _launches
    M, K, N = _check(x, w, b)
    _check_like(residual, M, N, "residual")
    y = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    pre = torch.empty((M, N) if save_pre else (0,), dtype=torch.bfloat16, device=x.device)
    with torch.cuda.device(x.device):
        rc = _lib.lib().ftk_linear_fwd(x.data_ptr(), w.data_ptr(), _ptr(b), _ptr(residual), M, K, N, act, y.data_ptr(),
                                       pre.data_ptr() if save_pre else 0, _stream(x))
    _lib.check(rc, "ftk_linear_fwd")
    _launches += 1
    return y, pre


@linear_fwd.register_fake
def _(x, w, b, residual, act, save_pre):
    return x.new_empty((x.shape[0], w.shape[0])), x.new_empty((x.shape[0], w.shape[0]) if save_pre else (0,))


def _bwd_impl(x, w, dy, mul_pre, dx_add, need_dx, dw_mode, dW_into, need_db):
    global _launches
    M, K, N = _check(x, w, None)
    _need_cuda(dy, mul_pre)
    _check_like(dx_add, M, K, "dx_add")
    _check_like(mul_pre, M, K, "mul_pre")
    if dy.shape != (M, N) or dy.dtype != torch.bfloat16 or not dy.is_contiguous():
        raise _lib.FtkError("dy must be a contiguous bfloat16 [M, N] tensor")
    dev = x.device
    dx = torch.empty((M, K) if need_dx else (0,), dtype=torch.bfloat16, device=dev)
    acc = 0
    if dw_mode == 0:      # no weight gradient
        dw, code = x.new_empty((0,)), _lib.FTK_BF16
    elif dW_into is not None:
        dw, code, acc = dW_into, (_lib.FTK_F32 if dW_into.dtype == torch.float32 else _lib.FTK_BF16), 1
        if getattr(dW_into, "_ftk_overwrite", False):
            dW_into._ftk_overwrite, acc = False, 0
    else:
        dw, code = torch.empty((N, K), dtype=torch.bfloat16, device=dev), _lib.FTK_BF16
    db = torch.empty((N,) if need_db else (0,), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().ftk_linear_bwd(x.data_ptr(), w.data_ptr(), dy.data_ptr(), _ptr(mul_pre), _ptr(dx_add), M, K, N,
                                       dx.data_ptr() if need_dx else 0, dw.data_ptr() if dw_mode else 0, code, acc,
                                       db.data_ptr() if need_db else 0, _stream(x))
    _lib.check(rc, "ftk_linear_bwd")
    _launches += int(need_dx) + int(dw_mode != 0) + int(need_db)
    return dx, dw, db


@torch.library.custom_op("ftk::linear_bwd", mutates_args=())
def linear_bwd(x: torch.Tensor, w: torch.Tensor, dy: torch.Tensor, mul_pre: Optional[torch.Tensor],
               dx_add: Optional[torch.Tensor], need_dx: bool, need_dw: bool, need_db: bool
               ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(dx, dw, db) of y = x @ w.T + b; dx is multiplied by ``mul_pre`` (the saved gelu_new' of the layer that produced
    x) when given, then ``dx_add`` is added.
    Unwanted ones are empty."""
    return _bwd_impl(x, w, dy, mul_pre, dx_add, need_dx, int(need_dw), None, need_db)


@linear_bwd.register_fake
def _(x, w, dy, mul_pre, dx_add, need_dx, need_dw, need_db):
    return (x.new_empty(x.shape if need_dx else (0,)), w.new_empty(w.shape if need_dw else (0,)),
            x.new_empty((w.shape[0],) if need_db else (0,), dtype=torch.float32))


@torch.library.custom_op("ftk::linear_bwd_into", mutates_args=("dW_into",))
def linear_bwd_into(x: torch.Tensor, w: torch.Tensor, dy: torch.Tensor, mul_pre: Optional[torch.Tensor],
                    dx_add: Optional[torch.Tensor], need_dx: bool, need_db: bool, dW_into: torch.Tensor
                    ) -> Tuple[torch.Tensor, torch.Tensor]:
    """``linear_bwd`` with the dW GEMM adding into ``dW_into`` (bf16 or fp32 [N, K]).  Returns (dx, db)."""
    dx, _, db = _bwd_impl(x, w, dy, mul_pre, dx_add, need_dx, 2, dW_into, need_db)
    return dx, db


@linear_bwd_into.register_fake
def _(x, w, dy, mul_pre, dx_add, need_dx, need_db, dW_into):
    return x.new_empty(x.shape if need_dx else (0,)), x.new_empty((w.shape[0],) if need_db else (0,), dtype=torch.float32)


def _backward(x, w, b, dy, mul_pre, need_dx, need_dw, need_db, dx_add=None):
    """dx, dw (None when it was added into the parameter's gradient buffer in the kernel), db."""
    into = _wgrad_target(w) if need_dw else None
    if into is not None:
        dx, db = torch.ops.ftk.linear_bwd_into(x, w, dy, mul_pre, dx_add, need_dx, need_db, into)
        dw = None
    else:
        dx, dw, db = torch.ops.ftk.linear_bwd(x, w, dy, mul_pre, dx_add, need_dx, need_dw, need_db)
        dw = dw if need_dw else None
    return (dx if need_dx else None), dw, (db.to(b.dtype) if need_db else None)


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, residual):
        y, _ = torch.ops.ftk.linear_fwd(x, w, b, residual, ACT_NONE, False)
        ctx.save_for_backward(x, w, b)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, b = ctx.saved_tensors
        dy = dy.contiguous()
        dx, dw, db = _backward(x, w, b, dy, None, ctx.needs_input_grad[0], ctx.needs_input_grad[1],
                               b is not None and ctx.needs_input_grad[2])
        return dx, dw, db, (dy if ctx.needs_input_grad[3] else None)  # the residual's gradient is dy itself


class _MlpGeluNew(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w_fc, b_fc, w_proj, b_proj, residual):
        a, dact = torch.ops.ftk.linear_fwd(x, w_fc, b_fc, None, ACT_GELU_NEW, True)  # dact = gelu_new'(c_fc(x))
        y, _ = torch.ops.ftk.linear_fwd(a, w_proj, b_proj, residual, ACT_NONE, False)
        ctx.save_for_backward(x, w_fc, b_fc, w_proj, b_proj, dact, a)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w_fc, b_fc, w_proj, b_proj, dact, a = ctx.saved_tensors
        ng = ctx.needs_input_grad
        dy = dy.contiguous()
        # c_proj: its dX GEMM hands back d(pre) directly (its epilogue multiplies by the saved gelu_new'(pre))
        dpre, dw_proj, db_proj = _backward(a, w_proj, b_proj, dy, dact, True, ng[3], b_proj is not None and ng[4])
        dx, dw_fc, db_fc = _backward(x, w_fc, b_fc, dpre, None, ng[0], ng[1], b_fc is not None and ng[2])
        return dx, dw_fc, db_fc, dw_proj, db_proj, (dy if ng[5] else None)


class _Qkv(torch.autograd.Function):
    """Three projections of one input.  Backward: the three dX GEMMs chain (each adds the previous one's result in
    its epilogue), so autograd has nothing left to sum."""

    @staticmethod
    def forward(ctx, x, wq, bq, wk, bk, wv, bv):
        q, _ = torch.ops.ftk.linear_fwd(x, wq, bq, None, ACT_NONE, False)
        k, _ = torch.ops.ftk.linear_fwd(x, wk, bk, None, ACT_NONE, False)
        v, _ = torch.ops.ftk.linear_fwd(x, wv, bv, None, ACT_NONE, False)
        ctx.save_for_backward(x, wq, bq, wk, bk, wv, bv)
        return q, k, v

    @staticmethod
    def backward(ctx, dq, dk, dv):
        x, wq, bq, wk, bk, wv, bv = ctx.saved_tensors
        ng = ctx.needs_input_grad
        dx, grads = None, []
        for w, b, dy, iw in ((wq, bq, dq, 1), (wk, bk, dk, 3), (wv, bv, dv, 5)):
            dx, dw, db = _backward(x, w, b, dy.contiguous(), None, ng[0], ng[iw], b is not None and ng[iw + 1], dx_add=dx)
            grads += [dw, db]
        return (dx, *grads)


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None,
           residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``torch.nn.functional.linear`` (+ ``residual``, shaped like the result) for bf16 CUDA tensors on the tcgen05
    kernels; any leading dimensions."""
    lead = x.shape[:-1]
    res = None if residual is None else residual.reshape(-1, weight.shape[0]).contiguous()
    y = _Linear.apply(x.reshape(-1, x.shape[-1]).contiguous(), weight, bias, res)
    return y.reshape(*lead, weight.shape[0])


def mlp_gelu_new(x: torch.Tensor, w_fc: torch.Tensor, b_fc: Optional[torch.Tensor], w_proj: torch.Tensor,
                 b_proj: Optional[torch.Tensor], residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``c_proj(gelu_new(c_fc(x)))`` (+ ``residual``) -- HF GPTNeoMLP / GPTJMLP without their dropout (resid_dropout = 0
    in both configs) -- as two GEMMs with the bias, the activation, its derivative and the residual add inside their
    epilogues."""
    lead = x.shape[:-1]
    res = None if residual is None else residual.reshape(-1, w_proj.shape[0]).contiguous()
    y = _MlpGeluNew.apply(x.reshape(-1, x.shape[-1]).contiguous(), w_fc, b_fc, w_proj, b_proj, res)
    return y.reshape(*lead, w_proj.shape[0])


def qkv(x: torch.Tensor, wq, bq, wk, bk, wv, bv) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``(linear(x, wq, bq), linear(x, wk, bk), linear(x, wv, bv))`` with one chained input gradient."""
    lead = x.shape[:-1]
    q, k, v = _Qkv.apply(x.reshape(-1, x.shape[-1]).contiguous(), wq, bq, wk, bk, wv, bv)
    return q.reshape(*lead, -1), k.reshape(*lead, -1), v.reshape(*lead, -1)


# ------------------------------------------------------------------------------------------------ LayerNorm
def _check_ln(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor]) -> Tuple[int, int]:
    _need_cuda(x, w, b)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or (b is not None and b.dtype != torch.bfloat16):
        raise _lib.FtkError("ftk layer_norm takes bfloat16 activations, weight and bias")
    if x.dim() != 2 or not x.is_contiguous() or w.shape != (x.shape[1],) or (b is not None and b.shape != w.shape):
        raise _lib.FtkError(f"layer_norm: x {tuple(x.shape)} (contiguous 2-D), weight {tuple(w.shape)}")
    return x.shape[0], x.shape[1]


@torch.library.custom_op("ftk::layernorm_fwd", mutates_args=())
def layernorm_fwd(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], eps: float
                  ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(y, mean, rstd) of ``layer_norm(x, (D,), w, b, eps)`` over the last dimension of a bf16 [M, D] matrix."""
    global _launches
    M, D = _check_ln(x, w, b)
    y = torch.empty_like(x)
    mean = torch.empty((M,), dtype=torch.float32, device=x.device)
    rstd = torch.empty((M,), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        rc = _lib.lib().ftk_layernorm_fwd(x.data_ptr(), w.data_ptr(), _ptr(b), M, D, eps, y.data_ptr(), mean.data_ptr(),
                                          rstd.data_ptr(), _stream(x))
    _lib.check(rc, "ftk_layernorm_fwd")
    _launches += 1
    return y, mean, rstd


@layernorm_fwd.register_fake
def _(x, w, b, eps):
    return torch.empty_like(x), x.new_empty((x.shape[0],), dtype=torch.float32), x.new_empty((x.shape[0],), dtype=torch.float32)


@torch.library.custom_op("ftk::layernorm_bwd", mutates_args=())
def layernorm_bwd(dy: torch.Tensor, x: torch.Tensor, w: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor,
                  dx_add: Optional[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(dx, dweight, dbias) of ``layernorm_fwd``, the last two in fp32; ``dx_add`` (shaped like x) is added to dx."""
    global _launches
    M, D = _check_ln(x, w, None)
    _need_cuda(dy, mean, rstd, dx_add)
    _check_like(dy, M, D, "dy")
    _check_like(dx_add, M, D, "dx_add")
    if mean.dtype != torch.float32 or rstd.dtype != torch.float32 or mean.numel() != M or rstd.numel() != M:
        raise _lib.FtkError("layernorm_bwd: mean and rstd must be float32 [M] tensors")
    dx = torch.empty_like(x)
    dg = torch.empty((D,), dtype=torch.float32, device=x.device)
    db = torch.empty((D,), dtype=torch.float32, device=x.device)
    L = _lib.lib()
    with torch.cuda.device(x.device):
        ws_bytes = L.ftk_layernorm_bwd_workspace(M, D)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=x.device)
        rc = L.ftk_layernorm_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), mean.data_ptr(), rstd.data_ptr(), _ptr(dx_add),
                                 M, D, dx.data_ptr(), dg.data_ptr(), db.data_ptr(), ws.data_ptr(), ws_bytes, _stream(x))
    _lib.check(rc, "ftk_layernorm_bwd")
    _launches += 2
    return dx, dg, db


@layernorm_bwd.register_fake
def _(dy, x, w, mean, rstd, dx_add):
    return torch.empty_like(x), x.new_empty(w.shape, dtype=torch.float32), x.new_empty(w.shape, dtype=torch.float32)


class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, eps):
        y, mean, rstd = torch.ops.ftk.layernorm_fwd(x, w, b, eps)
        ctx.save_for_backward(x, w, mean, rstd)
        ctx.has_bias = b is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, mean, rstd = ctx.saved_tensors
        dx, dg, db = torch.ops.ftk.layernorm_bwd(dy.contiguous(), x, w, mean, rstd, None)
        return dx, dg.to(w.dtype), (db.to(w.dtype) if ctx.has_bias else None), None


class _LayerNormFork(torch.autograd.Function):
    """(x, layer_norm(x)): the point of a pre-LN block where the residual branch and the normalised branch part.
    Backward: one kernel returns d(residual branch) + d(LayerNorm input)."""

    @staticmethod
    def forward(ctx, x, w, b, eps):
        y, mean, rstd = torch.ops.ftk.layernorm_fwd(x, w, b, eps)
        ctx.save_for_backward(x, w, mean, rstd)
        ctx.has_bias = b is not None
        return x, y  # autograd hands the caller a view of x for the first output

    @staticmethod
    def backward(ctx, d_pass, dy):
        x, w, mean, rstd = ctx.saved_tensors
        if dy is None:
            return d_pass, None, None, None
        dx, dg, db = torch.ops.ftk.layernorm_bwd(dy.contiguous(), x, w, mean, rstd,
                                                 None if d_pass is None else d_pass.contiguous())
        return dx, dg.to(w.dtype), (db.to(w.dtype) if ctx.has_bias else None), None


def layer_norm(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, eps: float = 1e-5) -> torch.Tensor:
    """``torch.nn.functional.layer_norm(x, (D,), weight, bias, eps)`` for bf16 CUDA tensors; any leading dimensions;
    D a multiple of 256, at most 8192."""
    lead = x.shape[:-1]
    y = _LayerNorm.apply(x.reshape(-1, x.shape[-1]).contiguous(), weight, bias, float(eps))
    return y.reshape(*lead, x.shape[-1])


def layer_norm_fork(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, eps: float = 1e-5
                    ) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(x, layer_norm(x))`` for a contiguous bf16 [M, D] matrix, with the sum of the two branches' gradients formed
    inside the LayerNorm's backward kernel."""
    return _LayerNormFork.apply(x, weight, bias, float(eps))


@torch.library.custom_op("ftk::attention_fwd", mutates_args=())
def attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, key_mask: Optional[torch.Tensor], B: int, T: int,
                  H: int, window: int, scale: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """q, k, v: bf16 [B * T, H * head_dim] (the projections' outputs).  Returns (o, lse fp32 [B * H * T])."""
    global _launches
    _need_cuda(q, k, v, key_mask)
    for t in (q, k, v):
        if t.dtype != torch.bfloat16 or t.dim() != 2 or t.shape != q.shape or not t.is_contiguous():
            raise _lib.FtkError("attention operands must be equal-shaped contiguous bfloat16 [B * T, H * head_dim] matrices")
    if q.shape[0] != B * T or q.shape[1] % H:
        raise _lib.FtkError(f"attention: operand shape {tuple(q.shape)} does not match B={B} T={T} H={H}")
    if key_mask is not None and (key_mask.dtype != torch.uint8 or key_mask.numel() != B * T or not key_mask.is_contiguous()):
        raise _lib.FtkError("attention: key_mask must be a contiguous uint8 tensor with B * T entries")
    dh = q.shape[1] // H
    o = torch.empty_like(q)
    lse = torch.empty((B * H * T,), dtype=torch.float32, device=q.device)
    with torch.cuda.device(q.device):
        rc = _lib.lib().ftk_attention_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), _ptr(key_mask), B, T, H, dh, window,
                                          scale, o.data_ptr(), lse.data_ptr(), _stream(q))
    _lib.check(rc, "ftk_attention_fwd")
    _launches += 1
    return o, lse


@attention_fwd.register_fake
def _(q, k, v, key_mask, B, T, H, window, scale):
    return torch.empty_like(q), q.new_empty((B * H * T,), dtype=torch.float32)


@torch.library.custom_op("ftk::attention_bwd", mutates_args=())
def attention_bwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor, d_o: torch.Tensor,
                  lse: torch.Tensor, key_mask: Optional[torch.Tensor], B: int, T: int, H: int, window: int, scale: float
                  ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(dq, dk, dv) of ``attention_fwd``."""
    global _launches
    _need_cuda(q, k, v, o, d_o, lse, key_mask)
    if d_o.shape != q.shape or d_o.dtype != torch.bfloat16 or not d_o.is_contiguous():
        raise _lib.FtkError("attention_bwd: d_o must be a contiguous bfloat16 tensor shaped like q")
    dh = q.shape[1] // H
    dq, dk, dv = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
    L = _lib.lib()
    ws_bytes = L.ftk_attention_bwd_workspace(B, T, H)
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=q.device)
    with torch.cuda.device(q.device):
        rc = L.ftk_attention_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), d_o.data_ptr(), lse.data_ptr(),
                                 _ptr(key_mask), B, T, H, dh, window, scale, dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
                                 ws.data_ptr(), ws_bytes, _stream(q))
    _lib.check(rc, "ftk_attention_bwd")
    _launches += 3
    return dq, dk, dv


@attention_bwd.register_fake
def _(q, k, v, o, d_o, lse, key_mask, B, T, H, window, scale):
    return torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)


class _Attention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, key_mask, B, T, H, window, scale):
        o, lse = torch.ops.ftk.attention_fwd(q, k, v, key_mask, B, T, H, window, scale)
        ctx.save_for_backward(q, k, v, o, lse, key_mask)
        ctx.cfg = (B, T, H, window, scale)
        return o

    @staticmethod
    def backward(ctx, d_o):
        q, k, v, o, lse, key_mask = ctx.saved_tensors
        B, T, H, window, scale = ctx.cfg
        dq, dk, dv = torch.ops.ftk.attention_bwd(q, k, v, o, d_o.contiguous(), lse, key_mask, B, T, H, window, scale)
        return dq, dk, dv, None, None, None, None, None, None


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, key_mask: Optional[torch.Tensor], heads: int,
              window: int = 0, scale: float = 1.0) -> torch.Tensor:
    """Causal self-attention softmax(scale * Q.K^T + bias).V per head, as HF GPT-Neo's eager ``_attn`` computes it (fp32
    scores, no 1/sqrt(d) factor there: pass ``scale``).  ``q, k, v``: bf16 [B, T, heads * head_dim], heads side by side
    as the projections produce them (no split / merge copies); ``key_mask``: bool / uint8 [B, T], False = never
    attended (padding), or None; ``window`` > 0: sliding window (key j visible from i iff i - window < j <= i).
    T must be a multiple of 128, head_dim 64 or 128.  Returns [B, T, heads * head_dim]."""
    B, T, D = q.shape
    km = None
    if key_mask is not None:
        km = key_mask.reshape(B * T).to(torch.uint8).contiguous()
    o = _Attention.apply(q.reshape(B * T, D).contiguous(), k.reshape(B * T, D).contiguous(),
                         v.reshape(B * T, D).contiguous(), km, B, T, heads, int(window), float(scale))
    return o.reshape(B, T, D)


def _key_mask_of(attention_mask: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """Which keys may be attended, [B, T] bool, from the mask HF hands its attention layers: a 4-D additive float mask
    (0 = keep, large negative = drop) or a 4-D bool mask (True = keep), already combined with the causal pattern.  The
    last query row sees every earlier key, so its row is exactly the padding mask."""
    if attention_mask is None:
        return None
    if attention_mask.dim() == 2:
        return attention_mask.bool()
    row = attention_mask[:, 0, -1, :]
    return row if row.dtype == torch.bool else row >= 0


def _block_forward(self, hidden_states, layer_past=None, attention_mask=None, use_cache=False, output_attentions=False,
                   **kwargs):
    """HF GPTNeoBlock.forward (ln_1 -> attention -> + residual -> ln_2 -> MLP -> + residual) on this library's kernels
    only: LayerNorm forks, chained q/k/v projections, attention, and the two residual adds inside the epilogues of
    out_proj's and c_proj's GEMMs.  Same parameters, same result up to bf16 rounding (the adds happen in fp32 before
    the single rounding); no elementwise kernel runs, forward or backward."""
    if layer_past is not None or use_cache or output_attentions or kwargs.get("head_mask") is not None:
        raise _lib.FtkError("the fused block serves training steps only (no KV cache, no attention maps, no head mask)")
    B, T, D = hidden_states.shape
    att = self.attn.attention
    x = hidden_states.reshape(B * T, D)
    if not x.is_contiguous():
        x = x.contiguous()
    res, h = layer_norm_fork(x, self.ln_1.weight, self.ln_1.bias, self.ln_1.eps)
    q, k, v = _Qkv.apply(h, att.q_proj.weight, att.q_proj.bias, att.k_proj.weight, att.k_proj.bias, att.v_proj.weight,
                         att.v_proj.bias)
    km = _key_mask_of(attention_mask)
    if km is not None:
        km = km.reshape(B * T).to(torch.uint8).contiguous()
    o = _Attention.apply(q, k, v, km, B, T, att.num_heads, att._ftk_window, 1.0)  # GPT-Neo: no 1/sqrt(d) factor
    s = _Linear.apply(o, att.out_proj.weight, att.out_proj.bias, res)            # attn_output + residual
    res2, h2 = layer_norm_fork(s, self.ln_2.weight, self.ln_2.bias, self.ln_2.eps)
    mlp = self.mlp
    y = _MlpGeluNew.apply(h2, mlp.c_fc.weight, mlp.c_fc.bias, mlp.c_proj.weight, mlp.c_proj.bias, res2)  # residual + ff
    return y.reshape(B, T, D), None


def patch_gpt_neo(model, attention_kernels: Optional[bool] = None, fuse_blocks: Optional[bool] = None) -> int:
    """Rebind, in place, the forward of every GPT-Neo MLP (``c_fc`` -> gelu_new -> ``c_proj``), of the attention's
    ``q_proj`` / ``k_proj`` / ``v_proj`` / ``out_proj``, (``attention_kernels``) of the self-attention itself and
    (``fuse_blocks``, needs the attention kernels and a hidden size that is a multiple of 256) of every LayerNorm and
    of the block as a whole (``_block_forward``) to the kernels above.  Parameters, module names and state dict are
    unchanged.  Returns the number of modules rebound; raises if the model is not a bf16 CUDA GPT-Neo with the
    gelu_new activation.  The attention kernels need the sequence length to be a multiple of 128 and no attention
    dropout; other lengths raise at call time."""
    cfg = model.config
    if getattr(cfg, "model_type", "") != "gpt_neo" or getattr(cfg, "activation_function", "") != "gelu_new":
        raise _lib.FtkError("patch_gpt_neo: expected a GPT-Neo model with activation_function='gelu_new'")
    if cfg.hidden_size % 64 or getattr(cfg, "intermediate_size", None) not in (None, 4 * cfg.hidden_size):
        raise _lib.FtkError("patch_gpt_neo: hidden_size must be a multiple of 64 (and the MLP 4x as wide)")
    head_dim = cfg.hidden_size // cfg.num_heads
    if attention_kernels is None:  # the attention kernels exist for head_dim 64 and 128; other sizes keep HF's attention
        attention_kernels = head_dim in (64, 128)
    elif attention_kernels and head_dim not in (64, 128):
        raise _lib.FtkError(f"patch_gpt_neo: the attention kernels need head_dim 64 or 128, not {head_dim}")
    ln_ok = cfg.hidden_size % 256 == 0 and cfg.hidden_size <= 8192
    if fuse_blocks is None:
        fuse_blocks = bool(attention_kernels) and ln_ok
    elif fuse_blocks and not (attention_kernels and ln_ok):
        raise _lib.FtkError("patch_gpt_neo: fuse_blocks needs the attention kernels and hidden_size % 256 == 0 (<= 8192)")
    n = 0
    if attention_kernels:
        cfg.use_cache = False  # as the reference sets it for training (FT.py:888); the fused attention keeps no KV cache

    def ln_forward(self, x):
        return layer_norm(x, self.weight, self.bias, self.eps)

    for block in model.base_model.h:
        mlp = block.mlp
        if float(getattr(mlp.dropout, "p", 0.0)) != 0.0:
            raise _lib.FtkError("patch_gpt_neo: resid_dropout must be 0 (the fused MLP has no dropout)")

        def mlp_forward(self, hidden_states):
            return mlp_gelu_new(hidden_states, self.c_fc.weight, self.c_fc.bias, self.c_proj.weight, self.c_proj.bias)

        mlp.forward = types.MethodType(mlp_forward, mlp)
        n += 1
        attn = block.attn.attention
        for name in ("q_proj", "k_proj", "v_proj", "out_proj"):
            lin = getattr(attn, name)

            def lin_forward(self, x):
                return linear(x, self.weight, self.bias)

            lin.forward = types.MethodType(lin_forward, lin)
            n += 1
        if attention_kernels:
            if float(getattr(attn.attn_dropout, "p", 0.0)) != 0.0 or float(getattr(attn.resid_dropout, "p", 0.0)) != 0.0:
                raise _lib.FtkError("patch_gpt_neo: attention_dropout and resid_dropout must be 0")
            window = int(cfg.window_size) if attn.attention_type == "local" else 0
            attn._ftk_window = window

            def attn_forward(self, hidden_states, attention_mask=None, layer_past=None, use_cache=False,
                             output_attentions=False, _window=window, **kwargs):
                if layer_past is not None or use_cache or output_attentions or kwargs.get("head_mask") is not None:
                    raise _lib.FtkError("the fused attention serves training steps only (no KV cache, no attention maps, "
                                        "no head mask)")
                q, k, v = qkv(hidden_states, self.q_proj.weight, self.q_proj.bias, self.k_proj.weight, self.k_proj.bias,
                              self.v_proj.weight, self.v_proj.bias)
                o = attention(q, k, v, _key_mask_of(attention_mask), self.num_heads, _window, 1.0)  # GPT-Neo: no scale
                return self.out_proj(o), None

            attn.forward = types.MethodType(attn_forward, attn)
            n += 1
        if fuse_blocks:
            for ln in (block.ln_1, block.ln_2):
                if not ln.elementwise_affine or ln.weight is None:
                    raise _lib.FtkError("patch_gpt_neo: LayerNorm without an affine weight")
                ln.forward = types.MethodType(ln_forward, ln)
                n += 1
            block.forward = types.MethodType(_block_forward, block)
            n += 1
    if fuse_blocks:
        ln_f = model.base_model.ln_f
        ln_f.forward = types.MethodType(ln_forward, ln_f)
        n += 1
    return n
