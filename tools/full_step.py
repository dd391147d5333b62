#!/usr/bin/env python
"""Context measurement (development / reporting tool): one real finetune step of a whole HF model on one B200, with
HF's own lm_head + loss versus this package's fused path, same trunk, same batch.  Reports step time, tokens/s and
peak memory.  The trunk is third-party code and out of this tier's scope; this only shows what the hot path is worth
inside a full step.   usage: python tools/full_step.py [gpt-neo-125m|gpt2|gpt-neo-1.3b] [batch] [ctx]"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200.finetune import load_model  # noqa: E402
from kubernetes_cloud_b200.loss import compute_loss  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "gpt-neo-1.3b"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
T = int(sys.argv[3]) if len(sys.argv) > 3 else 2048
torch.manual_seed(42)
model = load_model(arch).cuda().to(torch.bfloat16)
model.train()
model.gradient_checkpointing_enable()
V = model.config.vocab_size
g = torch.Generator(device="cuda").manual_seed(1)
ids = torch.randint(0, V, (B, T), device="cuda", generator=g)
mask = torch.ones(B, T, dtype=torch.bool, device="cuda")
opt = torch.optim.AdamW(model.parameters(), lr=1e-5, fused=True)


def step(fused: bool):
    opt.zero_grad(set_to_none=True)
    if fused:
        loss = compute_loss(model, {"input_ids": ids, "attention_mask": mask, "labels": ids.clone()})
    else:
        loss = model(input_ids=ids, attention_mask=mask, labels=ids).loss  # HF lm_head + ForCausalLMLoss
    loss.backward()
    opt.step()
    return loss


out = {"arch": arch, "batch": B, "ctx": T, "tokens": B * T}
variants = [("hf_lm_head_and_loss", False), ("fused_lm_head_ce", True)]
if model.config.model_type == "gpt_neo":
    # + trunk.patch_gpt_neo: whole blocks (LayerNorm, projections, attention, MLP, residual adds) on our kernels
    variants.append(("fused_lm_head_ce_and_ftk_trunk", True))
    # the same without gradient checkpointing: 180 GB of HBM hold the activations the reference's 48 GB cards had to
    # recompute (FT.py:884); identical arithmetic, one forward pass less
    variants.append(("fused_lm_head_ce_and_ftk_trunk_no_recompute", True))
for name, fused in variants:
    if name.endswith("ftk_trunk"):
        from kubernetes_cloud_b200 import trunk
        out["modules_rebound"] = trunk.patch_gpt_neo(model)
    if name.endswith("no_recompute"):
        model.gradient_checkpointing_disable()
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    for _ in range(2):
        loss = step(fused)
    torch.cuda.synchronize()
    t0 = time.time()
    n = 4
    for _ in range(n):
        loss = step(fused)
    torch.cuda.synchronize()
    dt = (time.time() - t0) / n
    out[name] = {"ms_per_step": round(dt * 1e3, 2), "tokens_per_s": round(B * T / dt), "loss": round(float(loss.detach()), 4),
                 "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 1e9, 2)}
    if fused:  # where this library's kernels spend the step (launch timeline of one more step)
        from kubernetes_cloud_b200 import _lib
        _lib.timeline_enable(True)
        step(fused)
        torch.cuda.synchronize()
        agg = {}
        for kname, ms in _lib.timeline_read():
            agg[kname] = agg.get(kname, 0.0) + ms
        _lib.timeline_enable(False)
        out[name]["ftk_kernels_ms"] = {k: round(v, 2) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])}
        out[name]["ftk_kernels_total_ms"] = round(sum(v for k, v in agg.items() if k != "convert"), 2)
print(json.dumps(out))
