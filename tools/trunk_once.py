#!/usr/bin/env python
"""Profiling target for the trunk kernels at the GPT-Neo-1.3B shape (16,384 tokens, d 2048, 16 heads of 128):
attention forward + backward (causal, then window 256), LayerNorm forward + backward, one linear layer with the
residual add in its epilogue.  Prints device times from the launch timeline and the HBM rates of the LayerNorm kernels.
    python tools/trunk_once.py [iters]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import _lib, trunk  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B, T, H, dh = 8, 2048, 16, 128
M, D = B * T, H * dh
g = torch.Generator(device="cuda").manual_seed(0)


def rnd(*shape, s=1.0):
    return (torch.randn(*shape, device="cuda", generator=g) * s).to(torch.bfloat16)


q, k, v = (rnd(B, T, D, s=s).requires_grad_(True) for s in (0.3, 0.3, 1.0))
d_o = rnd(B, T, D)
x = rnd(M, D).requires_grad_(True)
w_ln, b_ln = rnd(D, s=0.1) + 1.0, rnd(D, s=0.1)
w_ln.requires_grad_(True)
b_ln.requires_grad_(True)
w_lin, b_lin, res = rnd(D, D, s=D ** -0.5).requires_grad_(True), rnd(D, s=0.1).requires_grad_(True), rnd(M, D)
dy = rnd(M, D)


def once():
    for window in (0, 256):
        trunk.attention(q, k, v, None, H, window, 1.0).backward(d_o)
    res_out, y = trunk.layer_norm_fork(x, w_ln, b_ln, 1e-5)
    torch.autograd.backward([res_out, y], [dy, dy])
    trunk.linear(x, w_lin, b_lin, res).backward(dy)


for _ in range(2):
    once()
torch.cuda.synchronize()
_lib.timeline_enable(True)
for _ in range(iters):
    once()
torch.cuda.synchronize()
tl = _lib.timeline_read()
_lib.timeline_enable(False)
agg = {}
for name, ms in tl:
    agg.setdefault(name, []).append(ms)
out = {n: round(sum(v) / iters, 4) for n, v in agg.items()}
# attention launches alternate causal / window 256
for n in ("attn_fwd", "attn_bwd"):
    vals = agg[n]
    out[n + "_causal_ms"] = round(sum(vals[0::2]) / iters, 4)
    out[n + "_window256_ms"] = round(sum(vals[1::2]) / iters, 4)
    del out[n]
elems = M * D
out["layernorm_fwd_GBs"] = round(elems * 4 / (out["layernorm_fwd"] * 1e-3) / 1e9, 1)       # 2 B read + 2 B written
out["layernorm_bwd_GBs"] = round(elems * 8 / (out["layernorm_bwd"] * 1e-3) / 1e9, 1)       # x, dy, dx_add read, dx written
causal_flops = 4 * B * H * dh * (T * (T + 1) // 2)                                          # Q.K^T and P.V, visible pairs only
out["attn_fwd_causal_alg_TFs"] = round(causal_flops / (out["attn_fwd_causal_ms"] * 1e-3) / 1e12, 1)
out["attn_bwd_causal_alg_TFs"] = round(2.5 * causal_flops / (out["attn_bwd_causal_ms"] * 1e-3) / 1e12, 1)
print(json.dumps(out))
