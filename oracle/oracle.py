"""TEST INFRASTRUCTURE -- not part of the product.

CPU oracle for the training hot path of finetuner-workflow/finetuner. Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this
module; the product package (``kubernetes-cloud_b200``) never does and fails loudly without its CUDA library.

What it restates (paths relative to the reference root, FT.py = finetuner-workflow/finetuner/finetuner.py):

* batch assembly -- ``TokenizedDataset.load`` FT.py:662-692, ``collector`` FT.py:1030-1035 and the label fill
  of ``ModifiedTrainer.compute_loss`` FT.py:476-477.  Two forms: numpy (``gather_numpy``) and plain C
  (``oracle/gather_oracle.c`` through ctypes, ``gather_c``).
* ordering -- FT.py:739, :757-772 (``torch.manual_seed`` + ``random_split`` permutation, or file order) and
  the rank-strided slice the HF Trainer's DistributedSampler applies (torch/utils/data/distributed.py).
* lm_head + loss -- the arithmetic lives in a third-party dependency that is NOT in /root/reference:
  ``transformers~=4.28.1`` (finetuner/requirements.txt:1), call site FT.py:479.  Its published algorithm:
  ``logits = hidden @ W.T (+ b)``; ``logits.float()``; ``shift``; ``cross_entropy(ignore_index=-100, mean)``
  (transformers ``loss_utils.ForCausalLMLoss``).  Restated here with torch CPU fp32 ops.

Parity pin: the batch-assembly restatement is checked against fixtures produced by executing the reference's
own code (tests/golden/make_golden.py lifts the class/function bodies from the reference file).  The loss
restatement is checked against fixtures produced by the installed HF transformers (5.5.0) GPT-Neo / GPT-J
models on CPU -- the reference's pinned 4.28.1 is absent, so for the loss the pin is "installed stand-in".
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional, Tuple

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
IGNORE = -100


# ---------------------------------------------------------------------------------------------- batch assembly
def gather_numpy(tokens: np.ndarray, T: int, ctx_idx, pad_id: int, pad_is_eos: bool, shift_labels: bool = False):
    """numpy restatement of FT.py:662-692 + FT.py:1030-1035 + FT.py:476-477 (+ HF's causal shift).

    Returns (input_ids int64 [B,T], labels int64 [B,T], mask bool [B,T], n_valid int32 [B]).
    """
    tokens = np.asarray(tokens, dtype=np.uint16)
    n_ctx = tokens.size // T  # FT.py:647: a trailing partial context is dropped
    idx = np.asarray(ctx_idx, dtype=np.int64).reshape(-1)
    B = idx.size
    ids = np.empty((B, T), dtype=np.int64)
    mask = np.empty((B, T), dtype=bool)
    for b in range(B):
        i = int(idx[b])
        if not (0 <= i < n_ctx):
            raise IndexError(f"context index {i} out of range [0, {n_ctx})")
        arr = tokens[i * T:(i + 1) * T].astype(np.int64)  # FT.py:666-672
        ids[b] = arr
        if pad_is_eos and i != n_ctx - 1:  # FT.py:674
            mask[b] = True
        else:
            mask[b] = arr != pad_id  # FT.py:691
    labels = np.where(mask, ids, IGNORE)  # FT.py:1034 then FT.py:476-477
    if shift_labels:
        shifted = np.full_like(labels, IGNORE)
        shifted[:, :-1] = labels[:, 1:]
        labels = shifted
    n_valid = (labels != IGNORE).sum(axis=1).astype(np.int32)
    return ids, labels, mask, n_valid


_c_lib = None


def build_c(force: bool = False) -> str:
    """Compile oracle/gather_oracle.c -> oracle/_build/libftk_oracle.so (gcc only)."""
    out = os.path.join(_HERE, "_build", "libftk_oracle.so")
    src = os.path.join(_HERE, "gather_oracle.c")
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-o", out, src])
    return out


def _load_c():
    global _c_lib
    if _c_lib is None:
        lib = ctypes.CDLL(build_c())
        lib.ftk_oracle_gather.restype = None
        lib.ftk_oracle_gather.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_void_p,
                                          ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                          ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _c_lib = lib
    return _c_lib


def gather_c(tokens: np.ndarray, T: int, ctx_idx, pad_id: int, pad_is_eos: bool, shift_labels: bool = False):
    """Same contract as gather_numpy, computed by the plain-C restatement."""
    lib = _load_c()
    tokens = np.ascontiguousarray(tokens, dtype=np.uint16)
    n_ctx = tokens.size // T
    idx = np.ascontiguousarray(np.asarray(ctx_idx, dtype=np.int64).reshape(-1))
    if idx.size and (idx.min() < 0 or idx.max() >= n_ctx):
        raise IndexError("context index out of range")
    B = idx.size
    ids = np.empty((B, T), dtype=np.int64)
    labels = np.empty((B, T), dtype=np.int64)
    mask = np.empty((B, T), dtype=np.uint8)
    n_valid = np.empty((B,), dtype=np.int32)
    lib.ftk_oracle_gather(tokens.ctypes.data, n_ctx, T, idx.ctypes.data, B, pad_id, int(bool(pad_is_eos)),
                          int(bool(shift_labels)), ids.ctypes.data, labels.ctypes.data, mask.ctypes.data,
                          n_valid.ctypes.data)
    return ids, labels, mask.astype(bool), n_valid


# ---------------------------------------------------------------------------------------------- ordering
def train_order(n_ctx: int, seed: int, shuffle: bool = True) -> np.ndarray:
    """FT.py:739 + FT.py:757-772 with train_ratio forced to 1 (FT.py:748-756).

    shuffle: ``random_split(dataset, (N, 0))`` = ``randperm(N)`` drawn from the default CPU generator right after
    ``torch.manual_seed(seed)``.  no-shuffle: file order.
    """
    if not shuffle:
        return np.arange(n_ctx, dtype=np.int64)
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n_ctx, generator=g).numpy().astype(np.int64)


def rank_slice(order: np.ndarray, rank: int, world: int) -> np.ndarray:
    """DistributedSampler(shuffle=False, drop_last=False): pad by wrap-around to a multiple of world, take
    order[rank::world] (torch/utils/data/distributed.py:107-134)."""
    n = len(order)
    total = (n + world - 1) // world * world
    pad = total - n
    if pad:
        reps = (pad + n - 1) // n
        order = np.concatenate([order, np.tile(order, reps)[:pad]])
    return order[rank:total:world]


# ---------------------------------------------------------------------------------------------- lm_head + loss
def lmhead_ce_reference(H: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], targets: torch.Tensor,
                        want_grads: bool = True) -> dict:
    """fp32 CPU restatement of lm_head + HF causal-LM loss on ALREADY SHIFTED targets.

    logits = H @ W.T (+ b); loss = cross_entropy(logits.float(), targets, ignore_index=-100, reduction='mean')
    (transformers loss_utils.fixed_cross_entropy with num_items_in_batch=None, which is what the reference's
    3-argument compute_loss override produces, FT.py:475-479).  Inputs may be bf16: they are widened to fp32
    first, so the oracle sees exactly the values the CUDA path sees.
    Returns loss, lse [M], n_valid and, if want_grads, dH [M,d], dW [V,d], dbias [V].
    """
    Hf = H.detach().to("cpu", torch.float32).clone().requires_grad_(want_grads)
    Wf = W.detach().to("cpu", torch.float32).clone().requires_grad_(want_grads)
    bf = None if bias is None else bias.detach().to("cpu", torch.float32).clone().requires_grad_(want_grads)
    t = targets.detach().to("cpu", torch.int64)
    logits = Hf @ Wf.t()
    if bf is not None:
        logits = logits + bf
    lse = torch.logsumexp(logits.detach(), dim=-1)
    loss = torch.nn.functional.cross_entropy(logits, t, ignore_index=IGNORE, reduction="mean")
    out = {"loss": loss.detach(), "lse": lse, "n_valid": int((t != IGNORE).sum())}
    if want_grads:
        loss.backward()
        out["dH"] = Hf.grad
        out["dW"] = Wf.grad
        out["dbias"] = None if bf is None else bf.grad
    return out


def hf_causal_lm_loss(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """HF's ForCausalLMLoss on UNSHIFTED labels (transformers loss_utils.py): upcast, shift, mean CE."""
    logits = logits.float()
    shift = torch.nn.functional.pad(labels, (0, 1), value=IGNORE)[..., 1:].contiguous()
    return torch.nn.functional.cross_entropy(logits.view(-1, logits.size(-1)), shift.view(-1), ignore_index=IGNORE,
                                             reduction="mean")


# ---------------------------------------------------------------------------------------------- synthetic data
def synth_tokens(n_ctx: int, T: int, vocab: int, seed: int, pad_id: int = 50256, trailing_pads: int = 300) -> np.ndarray:
    """BASELINE.md section 3 / SURVEY section 8(d): uniform uint16 tokens; the final context ends in pads."""
    rng = np.random.default_rng(seed)
    tok = rng.integers(0, vocab, size=n_ctx * T, dtype=np.uint16)
    if trailing_pads:
        tok[n_ctx * T - min(trailing_pads, T):] = pad_id
    return tok


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """Norm-wise relative error ||a-b|| / ||b|| in fp64."""
    a = a.detach().to("cpu", torch.float64)
    b = b.detach().to("cpu", torch.float64)
    den = float(b.norm())
    return float((a - b).norm()) / (den if den > 0 else 1.0)


# ---------------------------------------------------------------------------------------------- optimizer (8f rank 1)
def warmup_lr_reference(it: int, max_lr: float, warmup_steps: int, min_lr: float = 0.0, warmup_type: str = "log") -> float:
    """DeepSpeed 0.9.2 ``runtime/lr_schedules.py`` WarmupLR (third-party, absent from /root/reference; named by the
    reference at finetuner/ds_config.json:19-26): gamma = log(it + 1) / log(warmup_steps) ("log") or
    it / warmup_steps ("linear") during warm-up, 1 afterwards; lr = min + (max - min) * gamma."""
    import math
    warmup_steps = max(2, warmup_steps)
    if it < warmup_steps:
        gamma = math.log(it + 1) / math.log(warmup_steps) if warmup_type == "log" else it / warmup_steps
    else:
        gamma = 1.0
    return min_lr + (max_lr - min_lr) * gamma


def adamw_reference(params, grads_per_step, lrs, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01, max_grad_norm=1.0):
    """torch.optim.AdamW + torch.nn.utils.clip_grad_norm_ on CPU fp32 (what HF/DeepSpeed apply for the reference:
    ds_config.json:10-18 AdamW, :44 gradient_clipping; FT.py:992 weight_decay 0.01).
    params: list of tensors; grads_per_step: list (per step) of lists of tensors; lrs: lr per step.
    Returns the updated parameters (fp32) and the pre-clip global norms."""
    ps = [torch.nn.Parameter(p.detach().to("cpu", torch.float32).clone()) for p in params]
    opt = torch.optim.AdamW(ps, lr=lrs[0], betas=betas, eps=eps, weight_decay=weight_decay)
    norms = []
    for lr, grads in zip(lrs, grads_per_step):
        for g in opt.param_groups:
            g["lr"] = lr
        for p, gr in zip(ps, grads):
            p.grad = gr.detach().to("cpu", torch.float32).clone()
        if max_grad_norm is not None:
            norms.append(float(torch.nn.utils.clip_grad_norm_(ps, max_grad_norm)))
        opt.step()
    return [p.detach() for p in ps], norms


def hf_causal_lm_loss_v428(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """The formulation of the transformers release the reference pins (~=4.28.1, e.g. GPTNeoForCausalLM.forward):
    ``shift_logits = logits[..., :-1, :]``, ``shift_labels = labels[..., 1:]``, ``CrossEntropyLoss()`` on the
    flattened pair after an fp32 upcast.  Restated from the published source of that version, which is not installed
    here; arithmetically the same mean as ``hf_causal_lm_loss`` (5.x pads the labels instead of slicing the logits)."""
    logits = logits.to(torch.float32)
    shift_logits = logits[..., :-1, :].contiguous()
    shift_labels = labels[..., 1:].contiguous()
    return torch.nn.CrossEntropyLoss()(shift_logits.view(-1, shift_logits.size(-1)), shift_labels.view(-1))


# ---------------------------------------------------------------------------------------------- BASELINE configs[0]
GPTNEO_125M = dict(vocab_size=50257, hidden_size=768, num_layers=12, num_heads=12, max_position_embeddings=2048,
                   attention_types=[[["global", "local"], 6]], window_size=256)


def weight_checksum(w: torch.Tensor) -> np.ndarray:
    w = w.detach().to("cpu", torch.float64)
    return np.array([float(w.sum()), float((w * w).sum()), float(w[0].sum()), float(w[-1].sum())])


def gptneo125m_lm_head_weight(seed: int = 42) -> torch.Tensor:
    """lm_head / wte weight [50257, 768] fp32 of the literal GPT-Neo-125M config, random-init under
    ``torch.manual_seed(seed)`` (the reference's default seed, FT.py:148-154) -- how tests/golden/make_golden.py built
    the model behind ``config0_gptneo125m.npz``.  The fixture holds a fingerprint instead of the 154 MB matrix."""
    from transformers import GPTNeoConfig, GPTNeoForCausalLM
    state = torch.random.get_rng_state()
    try:
        torch.manual_seed(seed)
        model = GPTNeoForCausalLM(GPTNeoConfig(**GPTNEO_125M))
    finally:
        torch.random.set_rng_state(state)
    return model.lm_head.weight.detach().clone()
