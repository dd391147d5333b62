#!/usr/bin/env python
"""Where does a whole finetune step go once the head and the trunk run on this library's kernels?  (development tool)

One GPT-Neo step (patched trunk + fused loss + torch fused AdamW, gradient checkpointing on, as tools/full_step.py)
under torch.profiler; prints the device-time table of every kernel, ours and torch's, so that what is still torch's
(LayerNorm, residual adds, copies, the optimizer) can be ranked.  A profiler run is never a bench number.
    usage: python tools/step_profile.py [gpt-neo-1.3b] [batch] [ctx] [out.txt]"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import trunk  # noqa: E402
from kubernetes_cloud_b200.finetune import load_model  # noqa: E402
from kubernetes_cloud_b200.loss import compute_loss  # noqa: E402

arch = sys.argv[1] if len(sys.argv) > 1 else "gpt-neo-1.3b"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
T = int(sys.argv[3]) if len(sys.argv) > 3 else 2048
out_path = sys.argv[4] if len(sys.argv) > 4 else None
torch.manual_seed(42)
model = load_model(arch).cuda().to(torch.bfloat16)
model.train()
model.gradient_checkpointing_enable()
trunk.patch_gpt_neo(model)
V = model.config.vocab_size
g = torch.Generator(device="cuda").manual_seed(1)
ids = torch.randint(0, V, (B, T), device="cuda", generator=g)
mask = torch.ones(B, T, dtype=torch.bool, device="cuda")
opt = torch.optim.AdamW(model.parameters(), lr=1e-5, fused=True)


def step():
    opt.zero_grad(set_to_none=True)
    loss = compute_loss(model, {"input_ids": ids, "attention_mask": mask, "labels": ids.clone()})
    loss.backward()
    opt.step()
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
table = prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=45, max_name_column_width=90)
if out_path:
    with open(out_path, "w") as f:
        f.write(table)
print(table)
