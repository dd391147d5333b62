#!/usr/bin/env python
"""Profiling target: attention forward + backward at the GPT-Neo-1.3B shape (batch 8, 16 heads, T 2048, head_dim 128).
python tools/attn_once.py [window] [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import _lib, trunk  # noqa: E402

window = int(sys.argv[1]) if len(sys.argv) > 1 else 0
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
B, T, H, dh = 8, 2048, 16, 128
g = torch.Generator(device="cuda").manual_seed(0)
q, k, v, d_o = ((torch.randn(B, T, H * dh, device="cuda", generator=g) * s).to(torch.bfloat16).requires_grad_(s != 3)
                for s in (0.3, 0.3, 1.0, 3))
for _ in range(2):  # warm-up: module load, function attributes
    trunk.attention(q, k, v, None, H, window, 1.0).backward(d_o)
torch.cuda.synchronize()
_lib.timeline_enable(True)
for _ in range(iters):
    o = trunk.attention(q, k, v, None, H, window, 1.0)
    o.backward(d_o)
torch.cuda.synchronize()
tl = _lib.timeline_read()
print({n: round(sum(ms for nn, ms in tl if nn == n) / iters, 3) for n in ("attn_fwd", "attn_bwd")}, float(o.float().abs().mean()))
