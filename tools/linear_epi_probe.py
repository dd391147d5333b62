#!/usr/bin/env python
"""How much do the linear layers' epilogues cost?  One GEMM shape per GPT-Neo-1.3B layer type, with and without the
fused epilogue work (bias, gelu_new + pre-activation store, residual add, gelu_new' in dX), device times from the launch
timeline.  Development tool."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import _lib, trunk  # noqa: E402

M, d = 16384, 2048
g = torch.Generator(device="cuda").manual_seed(0)


def rnd(*shape, s=1.0):
    return (torch.randn(*shape, device="cuda", generator=g) * s).to(torch.bfloat16)


x, x4 = rnd(M, d), rnd(M, 4 * d)
w_sq, w_up, w_dn = rnd(d, d, s=d ** -0.5), rnd(4 * d, d, s=d ** -0.5), rnd(d, 4 * d, s=(4 * d) ** -0.5)
b_d, b_4d, res, pre4 = rnd(d, s=0.1), rnd(4 * d, s=0.1), rnd(M, d), rnd(M, 4 * d)
dy, dy4 = rnd(M, d), rnd(M, 4 * d)
F = torch.ops.ftk
cases = {
    "fwd 2048->2048 plain": lambda: F.linear_fwd(x, w_sq, None, None, 0, False),
    "fwd 2048->2048 bias": lambda: F.linear_fwd(x, w_sq, b_d, None, 0, False),
    "fwd 2048->2048 bias+residual": lambda: F.linear_fwd(x, w_sq, b_d, res, 0, False),
    "fwd 2048->8192 plain": lambda: F.linear_fwd(x, w_up, None, None, 0, False),
    "fwd 2048->8192 bias+gelu": lambda: F.linear_fwd(x, w_up, b_4d, None, 1, False),
    "fwd 2048->8192 bias+gelu+pre": lambda: F.linear_fwd(x, w_up, b_4d, None, 1, True),
    "fwd 8192->2048 plain": lambda: F.linear_fwd(x4, w_dn, None, None, 0, False),
    "fwd 8192->2048 bias+residual": lambda: F.linear_fwd(x4, w_dn, b_d, res, 0, False),
    "dx 2048<-2048 plain": lambda: F.linear_bwd(x, w_sq, dy, None, None, True, False, False),
    "dx 2048<-2048 +dx_add": lambda: F.linear_bwd(x, w_sq, dy, None, res, True, False, False),
    "dx 8192<-2048 plain": lambda: F.linear_bwd(x4, w_dn, dy, None, None, True, False, False),
    "dx 8192<-2048 gelu'": lambda: F.linear_bwd(x4, w_dn, dy, pre4, None, True, False, False),
    "dx 2048<-8192 plain": lambda: F.linear_bwd(x, w_up, dy4, None, None, True, False, False),
    "dw 2048x2048": lambda: F.linear_bwd(x, w_sq, dy, None, None, False, True, False),
    "dw 8192x2048": lambda: F.linear_bwd(x, w_up, dy4, None, None, False, True, False),
    "dw 2048x8192": lambda: F.linear_bwd(x4, w_dn, dy, None, None, False, True, False),
}
flops = {k: 2 * M * d * d * (1 if "2048->2048" in k or "2048<-2048" in k or "2048x2048" in k else 4) for k in cases}
out = {}
for rounds in range(2):  # the second round runs at sustained (power-capped) clocks
    for name, fn in cases.items():
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        _lib.timeline_enable(True)
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        ms = sum(t for _, t in _lib.timeline_read()) / 10
        _lib.timeline_enable(False)
        out[name] = {"ms": round(ms, 4), "TFs": round(flops[name] / ms / 1e9, 0)}
for k, v in out.items():
    print(f"{k:34s} {v['ms']:8.4f} ms  {v['TFs']:6.0f} TF/s")
