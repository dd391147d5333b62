#!/usr/bin/env python
"""Profiling target: a few calls of the fused forward+backward (ftk_lmhead_ce_fused) at the bench shape, nothing else.
python tools/fused_once.py [iters] [mode]   mode: fused (default) | recompute"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import ops  # noqa: E402,F401

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
mode = sys.argv[2] if len(sys.argv) > 2 else "fused"
M, d, V = 16384, 2048, 50257
g = torch.Generator(device="cuda").manual_seed(1)
H = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
W = (torch.randn(V, d, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
t = torch.randint(0, V, (M,), device="cuda", generator=g)
for _ in range(iters):
    if mode == "fused":
        out = torch.ops.ftk.lmhead_ce_fused(H, W, None, t, 1.0, False, 0)
    else:
        ls, nv, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
        out = (ls, nv) + tuple(torch.ops.ftk.lmhead_ce_bwd(H, W, None, t, lse, None, nv, 1.0, False))
torch.cuda.synchronize()
print("ok", float(out[0]) / int(out[1]))
