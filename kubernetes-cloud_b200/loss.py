"""Mirror of ``ModifiedTrainer.compute_loss`` (FT.py:469-493) with the lm_head + loss replaced by the fused kernels.

Reference behaviour kept: ``labels`` are filled with -100 where ``attention_mask`` is False, in place; the loss is the
mean over this micro-batch's valid positions (no ``num_items_in_batch`` normalisation); the result
requires grad; a ``LOSS:`` line goes to stderr every ``2 * gradient_accumulation_steps`` calls on the main process.
"""
from __future__ import annotations

import sys
from typing import Optional

import torch

from .ops import IGNORE, fused_lmhead_ce


def shift_labels(labels: torch.Tensor) -> torch.Tensor:
    """targets[t] = labels[t+1], last position ignored (transformers loss_utils.ForCausalLMLoss)."""
    out = torch.full_like(labels, IGNORE)
    out[..., :-1] = labels[..., 1:]
    return out


def _trunk_and_head(model):
    trunk = getattr(model, "base_model", None)
    head = model.get_output_embeddings() if hasattr(model, "get_output_embeddings") else getattr(model, "lm_head")
    if trunk is None or head is None:
        raise TypeError("model must expose base_model and an output embedding (lm_head)")
    return trunk, head


def compute_loss(model, inputs: dict, return_outputs: bool = False, num_items_in_batch=None,
                 labels_are_shifted: bool = False, loss_scale: float = 1.0):
    """Same signature and meaning as the reference's override (plus the keyword newer Trainers pass).
    ``loss_scale``: returns ``loss_scale * loss``; a loop that accumulates gradients passes 1 / accumulation steps
    here and calls ``.backward()`` on the result directly (see ``ops.fused_lmhead_ce``)."""
    labels = inputs.get("labels")
    if labels is not None and not labels_are_shifted:
        labels.masked_fill_(~inputs["attention_mask"], IGNORE)  # FT.py:476-477, in place like the reference
    trunk, head = _trunk_and_head(model)
    hidden = trunk(input_ids=inputs["input_ids"], attention_mask=inputs["attention_mask"]).last_hidden_state
    targets = labels if labels_are_shifted else shift_labels(labels)
    d = hidden.shape[-1]
    h2 = hidden.reshape(-1, d)
    if h2.dtype != torch.bfloat16:
        h2 = h2.to(torch.bfloat16)
    w = head.weight if head.weight.dtype == torch.bfloat16 else head.weight.to(torch.bfloat16)
    b = getattr(head, "bias", None)
    if b is not None and b.dtype != torch.bfloat16:
        b = b.to(torch.bfloat16)
    loss = fused_lmhead_ce(h2.contiguous(), w.contiguous(), b, targets.reshape(-1).contiguous(), loss_scale)
    if not loss.requires_grad:
        loss.requires_grad_(True)  # FT.py:483
    return (loss, {"loss": loss, "hidden_states": hidden}) if return_outputs else loss


class ModifiedTrainerMixin:
    """Mix into ``transformers.Trainer`` in place of the reference's ``ModifiedTrainer``."""

    def compute_loss(self, model, inputs, return_outputs=False, num_items_in_batch=None):
        results = compute_loss(model, inputs, return_outputs)
        loss = results[0] if return_outputs else results
        self.report_idx = getattr(self, "report_idx", 0) + 1
        gas = getattr(getattr(self, "args", None), "gradient_accumulation_steps", 1)
        if self.report_idx % (2 * gas) == 0 and getattr(self, "is_world_process_zero", lambda: True)():
            # FT.py:486-491 formats the loss tensor, which stalls the host until the GPU catches up.  Here the value
            # goes to pinned memory with an asynchronous copy and the line printed is the previous report's value,
            # whose copy has long finished: same output cadence, no pipeline bubble.
            prev = getattr(self, "_loss_report", None)
            if prev is not None:
                buf, ev = prev
                ev.synchronize()
                print(f"\nLOSS: {float(buf):.3f}", file=sys.stderr, flush=True)
            if loss.is_cuda:
                buf = torch.empty((), dtype=torch.float32).pin_memory()
                buf.copy_(loss.detach().float(), non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                self._loss_report = (buf, ev)
        return results
