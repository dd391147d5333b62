#!/usr/bin/env python
"""BASELINE.json configs[4]: fused lm_head+cross-entropy sweep, 4k-256k tokens x vocab 50k at d_model 768-4096 on one
B200, plus the GPT-2 124M (configs[1]) and GPT-J-6B (configs[3]) head shapes.  Prints one JSON line per shape.
Development / reporting tool (results are copied into BASELINE.md); not part of the product."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import ops  # noqa: E402,F401


def run(M, d, V, bias=False, iters=3, tag=""):
    from kubernetes_cloud_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(M + d)
    H = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(V, d, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    b = (torch.randn(V, device="cuda", generator=g) * 0.1).to(torch.bfloat16) if bias else None
    t = torch.randint(0, V, (M,), device="cuda", generator=g)
    t[::2048] = -100
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, b, t)
    torch.ops.ftk.lmhead_ce_bwd(H, W, b, t, lse, None, n_valid, 1.0, False)
    torch.ops.ftk.lmhead_ce_fused(H, W, b, t, 1.0, False, 0)
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(iters):
        loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, b, t)
    ev[1].record()
    for _ in range(iters):
        torch.ops.ftk.lmhead_ce_bwd(H, W, b, t, lse, None, n_valid, 1.0, False)
    ev[2].record()
    for _ in range(iters):
        out = torch.ops.ftk.lmhead_ce_fused(H, W, b, t, 1.0, False, 0)
    ev[3].record()
    torch.cuda.synchronize()
    tf, tb = ev[0].elapsed_time(ev[1]) / iters, ev[1].elapsed_time(ev[2]) / iters
    tfu = ev[2].elapsed_time(ev[3]) / iters
    fl = 2.0 * M * V * d
    print(json.dumps({"tag": tag, "M": M, "d": d, "V": V, "bias": bias, "fwd_ms": round(tf, 3), "bwd_ms": round(tb, 3),
                      "fwd_TFs": round(fl / tf / 1e9, 1), "recompute_alg_TFs": round(3 * fl / (tf + tb) / 1e9, 1),
                      "fused_ms": round(tfu, 3), "fused_alg_TFs": round(3 * fl / tfu / 1e9, 1),
                      "fused_slab_rows": _lib.lib().ftk_lmhead_ce_fused_slab_rows(M, d, V, 0),
                      "tokens_per_s": round(M / (tfu / 1e3)), "loss": round(float(out[0]) / int(out[1]), 4)}),
          flush=True)
    del H, W, t, lse, out
    torch.cuda.empty_cache()


if __name__ == "__main__":
    run(16 * 1024, 768, 50257, tag="gpt2-124m ctx1024 bs16")
    run(8 * 2048, 2048, 50257, tag="gpt-neo-1.3b ctx2048 bs8")
    run(8 * 2048, 4096, 50400, bias=True, tag="gpt-j-6b ctx2048 bs8")
    if "--small" in sys.argv:
        run(2048, 768, 50257, tag="gpt-neo-125m ctx2048 bs1")
        for d in (768, 1024, 2048, 4096):
            run(4096, d, 50257, tag="small")
    if "--full" in sys.argv:
        for M in (4096, 16384, 65536, 262144):
            for d in (768, 1024, 2048, 4096):
                run(M, d, 50257, iters=2 if M >= 65536 else 3, tag="sweep")
