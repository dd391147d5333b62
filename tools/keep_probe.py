#!/usr/bin/env python
"""Timing of the opt-in keep-numerators mode against the default recompute mode (development / reporting tool)."""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import ops  # noqa

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

for (M, d, V) in ((16384, 768, 50257), (16384, 2048, 50257), (16384, 4096, 50400)):
    g = torch.Generator(device="cuda").manual_seed(1)
    H = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(V, d, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    t = torch.randint(0, V, (M,), device="cuda", generator=g)
    def step_recompute():
        ls, nv, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
        return torch.ops.ftk.lmhead_ce_bwd(H, W, None, t, lse, None, nv, 1.0, False)
    def step_keep():
        ls, nv, lse, probs, cref = torch.ops.ftk.lmhead_ce_fwd_keep(H, W, None, t)
        return torch.ops.ftk.lmhead_ce_bwd_kept(H, W, False, t, lse, probs, cref, None, nv, 1.0, False)
    def fwd_keep():
        return torch.ops.ftk.lmhead_ce_fwd_keep(H, W, None, t)
    a = step_recompute(); b = step_keep()
    err_dH = float((a[0].float() - b[0].float()).norm() / a[0].float().norm())
    err_dW = float((a[1].float() - b[1].float()).norm() / a[1].float().norm())
    tr, tk, tfk = timeit(step_recompute), timeit(step_keep), timeit(fwd_keep)
    fl = 6.0 * M * V * d
    print(json.dumps({"M": M, "d": d, "V": V, "recompute_ms": round(tr, 3), "keep_ms": round(tk, 3), "fwd_keep_ms": round(tfk, 3),
                      "speedup": round(tr / tk, 3), "recompute_alg_TFs": round(fl / tr / 1e9, 1), "keep_alg_TFs": round(fl / tk / 1e9, 1),
                      "dH_diff": err_dH, "dW_diff": err_dW}), flush=True)
    del H, W
