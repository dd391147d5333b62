#!/usr/bin/env python
"""Experiment: one forward + backward at the bench shape (for ncu dram-byte counting under different knobs)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import ops  # noqa
M, d, V = 16384, 2048, 50257
g = torch.Generator(device="cuda").manual_seed(1)
H = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
W = (torch.randn(V, d, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
t = torch.randint(0, V, (M,), device="cuda", generator=g)
for _ in range(2):
    ls, nv, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
    torch.ops.ftk.lmhead_ce_bwd(H, W, None, t, lse, None, nv, 1.0, False)
torch.cuda.synchronize()
print("ok", float(ls) / int(nv))
