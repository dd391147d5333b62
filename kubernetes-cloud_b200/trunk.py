"""Linear layers and MLP of the transformer trunk on the B200 kernels (first pieces).

The reference trains whatever HF causal LM ``--model`` names through ``model(**inputs)`` (FT.py:479) with gradient
checkpointing on (FT.py:884).  The trunk is third-party code; this module swaps the pieces of it that the existing
tcgen05 kernel template covers, leaving the module tree, parameter names and checkpoints untouched:

  ``linear(x, weight, bias)``                  torch.nn.functional.linear            -> ftk_linear_fwd / ftk_linear_bwd
  ``mlp_gelu_new(x, w_fc, b_fc, w_proj, b_proj)``   c_proj(gelu_new(c_fc(x)))       -> two GEMMs; bias + GELU live in the first
                                                one's epilogue, and the GELU's derivative in the epilogue of the second
                                                one's dX GEMM, so no elementwise kernel ever touches the [tokens, 4d]
                                                activations
  ``patch_gpt_neo(model)``                     rebinds the forward of every GPTNeoMLP and of the attention projections

  ``attention(q, k, v, key_mask, heads, window, scale)``   HF's eager fp32 softmax(Q.K^T).V with its causal / sliding-
                                                window bias                           -> ftk_attention_fwd / _bwd (the [T, T]
                                                scores never leave TMEM)
LayerNorm and the residual adds are still torch's; they are the next pieces of this row.
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
def linear_fwd(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], act: int, save_pre: bool
               ) -> Tuple[torch.Tensor, torch.Tensor]:
    """(y, pre): y = act(x @ w.T + b); pre = x @ w.T + b when ``save_pre`` (else empty)."""
    global _launches
    M, K, N = _check(x, w, b)
    y = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    pre = torch.empty((M, N) if save_pre else (0,), dtype=torch.bfloat16, device=x.device)
    with torch.cuda.device(x.device):
        rc = _lib.lib().ftk_linear_fwd(x.data_ptr(), w.data_ptr(), _ptr(b), M, K, N, act, y.data_ptr(),
                                       pre.data_ptr() if save_pre else 0, _stream(x))
    _lib.check(rc, "ftk_linear_fwd")
    _launches += 1
    return y, pre


@linear_fwd.register_fake
def _(x, w, b, act, save_pre):
    return x.new_empty((x.shape[0], w.shape[0])), x.new_empty((x.shape[0], w.shape[0]) if save_pre else (0,))


def _bwd_impl(x, w, dy, mul_pre, need_dx, dw_mode, dW_into, need_db):
    global _launches
    M, K, N = _check(x, w, None)
    _need_cuda(dy, mul_pre)
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
        rc = _lib.lib().ftk_linear_bwd(x.data_ptr(), w.data_ptr(), dy.data_ptr(), _ptr(mul_pre), M, K, N,
                                       dx.data_ptr() if need_dx else 0, dw.data_ptr() if dw_mode else 0, code, acc,
                                       db.data_ptr() if need_db else 0, _stream(x))
    _lib.check(rc, "ftk_linear_bwd")
    _launches += int(need_dx) + int(dw_mode != 0) + int(need_db)
    return dx, dw, db


@torch.library.custom_op("ftk::linear_bwd", mutates_args=())
def linear_bwd(x: torch.Tensor, w: torch.Tensor, dy: torch.Tensor, mul_pre: Optional[torch.Tensor], need_dx: bool,
               need_dw: bool, need_db: bool) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(dx, dw, db) of y = x @ w.T + b; dx is multiplied by gelu_new'(mul_pre) when given.  Unwanted ones are empty."""
    return _bwd_impl(x, w, dy, mul_pre, need_dx, int(need_dw), None, need_db)


@linear_bwd.register_fake
def _(x, w, dy, mul_pre, need_dx, need_dw, need_db):
    return (x.new_empty(x.shape if need_dx else (0,)), w.new_empty(w.shape if need_dw else (0,)),
            x.new_empty((w.shape[0],) if need_db else (0,), dtype=torch.float32))


@torch.library.custom_op("ftk::linear_bwd_into", mutates_args=("dW_into",))
def linear_bwd_into(x: torch.Tensor, w: torch.Tensor, dy: torch.Tensor, mul_pre: Optional[torch.Tensor], need_dx: bool,
                    need_db: bool, dW_into: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``linear_bwd`` with the dW GEMM adding into ``dW_into`` (bf16 or fp32 [N, K]).  Returns (dx, db)."""
    dx, _, db = _bwd_impl(x, w, dy, mul_pre, need_dx, 2, dW_into, need_db)
    return dx, db


@linear_bwd_into.register_fake
def _(x, w, dy, mul_pre, need_dx, need_db, dW_into):
    return x.new_empty(x.shape if need_dx else (0,)), x.new_empty((w.shape[0],) if need_db else (0,), dtype=torch.float32)


def _backward(x, w, b, dy, mul_pre, need_dx, need_dw, need_db):
    """dx, dw (None when it was added into the parameter's gradient buffer in the kernel), db."""
    into = _wgrad_target(w) if need_dw else None
    if into is not None:
        dx, db = torch.ops.ftk.linear_bwd_into(x, w, dy, mul_pre, need_dx, need_db, into)
        dw = None
    else:
        dx, dw, db = torch.ops.ftk.linear_bwd(x, w, dy, mul_pre, need_dx, need_dw, need_db)
        dw = dw if need_dw else None
    return (dx if need_dx else None), dw, (db.to(b.dtype) if need_db else None)


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        y, _ = torch.ops.ftk.linear_fwd(x, w, b, ACT_NONE, False)
        ctx.save_for_backward(x, w, b)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, b = ctx.saved_tensors
        dx, dw, db = _backward(x, w, b, dy.contiguous(), None, ctx.needs_input_grad[0], ctx.needs_input_grad[1],
                               b is not None and ctx.needs_input_grad[2])
        return dx, dw, db


class _MlpGeluNew(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w_fc, b_fc, w_proj, b_proj):
        a, pre = torch.ops.ftk.linear_fwd(x, w_fc, b_fc, ACT_GELU_NEW, True)
        y, _ = torch.ops.ftk.linear_fwd(a, w_proj, b_proj, ACT_NONE, False)
        ctx.save_for_backward(x, w_fc, b_fc, w_proj, b_proj, pre, a)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w_fc, b_fc, w_proj, b_proj, pre, a = ctx.saved_tensors
        ng = ctx.needs_input_grad
        # c_proj: its dX GEMM hands back d(pre) directly (epilogue multiplies by gelu_new'(pre))
        dpre, dw_proj, db_proj = _backward(a, w_proj, b_proj, dy.contiguous(), pre, True, ng[3],
                                           b_proj is not None and ng[4])
        dx, dw_fc, db_fc = _backward(x, w_fc, b_fc, dpre, None, ng[0], ng[1], b_fc is not None and ng[2])
        return dx, dw_fc, db_fc, dw_proj, db_proj


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``torch.nn.functional.linear`` for bf16 CUDA tensors on the tcgen05 kernels; any leading dimensions."""
    lead = x.shape[:-1]
    y = _Linear.apply(x.reshape(-1, x.shape[-1]).contiguous(), weight, bias)
    return y.reshape(*lead, weight.shape[0])


def mlp_gelu_new(x: torch.Tensor, w_fc: torch.Tensor, b_fc: Optional[torch.Tensor], w_proj: torch.Tensor,
                 b_proj: Optional[torch.Tensor]) -> torch.Tensor:
    """``c_proj(gelu_new(c_fc(x)))`` -- HF GPTNeoMLP / GPTJMLP without their dropout (resid_dropout = 0 in both
    configs) -- as two GEMMs with the bias, the activation and its derivative inside their epilogues."""
    lead = x.shape[:-1]
    y = _MlpGeluNew.apply(x.reshape(-1, x.shape[-1]).contiguous(), w_fc, b_fc, w_proj, b_proj)
    return y.reshape(*lead, w_proj.shape[0])


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


def patch_gpt_neo(model, attention_kernels: Optional[bool] = None) -> int:
    """Rebind, in place, the forward of every GPT-Neo MLP (``c_fc`` -> gelu_new -> ``c_proj``), of the attention's
    ``q_proj`` / ``k_proj`` / ``v_proj`` / ``out_proj`` and (``attention_kernels``) of the self-attention itself to the
    kernels above.  Parameters, module names and state dict are unchanged.  Returns the number of modules rebound;
    raises if the model is not a bf16 CUDA GPT-Neo with the gelu_new activation.  The attention kernels need the
    sequence length to be a multiple of 128 and no attention dropout; other lengths raise at call time."""
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
    n = 0
    if attention_kernels:
        cfg.use_cache = False  # as the reference sets it for training (FT.py:888); the fused attention keeps no KV cache
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

            def attn_forward(self, hidden_states, attention_mask=None, layer_past=None, use_cache=False,
                             output_attentions=False, _window=window, **kwargs):
                if layer_past is not None or use_cache or output_attentions:
                    raise _lib.FtkError("the fused attention serves training steps only (no KV cache, no attention maps)")
                q, k, v = self.q_proj(hidden_states), self.k_proj(hidden_states), self.v_proj(hidden_states)
                o = attention(q, k, v, _key_mask_of(attention_mask), self.num_heads, _window, 1.0)  # GPT-Neo: no scale
                return self.out_proj(o), None

            attn.forward = types.MethodType(attn_forward, attn)
            n += 1
    return n
