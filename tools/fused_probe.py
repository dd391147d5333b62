#!/usr/bin/env python
"""Development tool: forward + backward time of the three gradient routes (recompute / keep / fused) through the C ABI
at one shape, and the fused route across slab sizes.  python tools/fused_probe.py [M d V]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import _lib, ops  # noqa: E402

M, d, V = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (16384, 2048, 50257)
g = torch.Generator(device="cuda").manual_seed(1)
H = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
W = (torch.randn(V, d, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
t = torch.randint(0, V, (M,), device="cuda", generator=g)
L = _lib.lib()


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def recompute():
    ls, nv, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
    torch.ops.ftk.lmhead_ce_bwd(H, W, None, t, lse, None, nv, 1.0, False)


def keep():
    ls, nv, lse, probs, cref = torch.ops.ftk.lmhead_ce_fwd_keep(H, W, None, t)
    torch.ops.ftk.lmhead_ce_bwd_kept(H, W, False, t, lse, probs, cref, None, nv, 1.0, False)


def fused():
    torch.ops.ftk.lmhead_ce_fused(H, W, None, t, 1.0, False, 0)


flops = 6.0 * M * V * d
out = {"shape": [M, d, V], "auto_slab": L.ftk_lmhead_ce_fused_slab_rows(M, d, V, 0)}
for name, fn in (("recompute", recompute), ("keep", keep)):
    if name == "keep" and M * V * 2 > 20e9:
        continue
    ms = timed(fn)
    out[name] = {"ms": round(ms, 3), "algorithmic_tflops": round(flops / ms / 1e9, 1)}
for slab in (0, 2048, 4096, 5632, 8192, 9216, 16384):
    if slab > M and slab != 0:
        continue
    ops.set_fused_slab_rows(slab)
    ms = timed(fused)
    out[f"fused_slab{slab}"] = {"ms": round(ms, 3), "algorithmic_tflops": round(flops / ms / 1e9, 1),
                                "rows": L.ftk_lmhead_ce_fused_slab_rows(M, d, V, slab)}
ops.set_fused_slab_rows(0)
# per-kernel breakdown of the automatic configuration from the library's launch timeline
_lib.timeline_enable(True)
for _ in range(5):
    fused()
tl = _lib.timeline_read(with_start=True)
_lib.timeline_enable(False)
agg = {}
for name, ms, _ in tl:
    agg[name] = agg.get(name, 0.0) + ms / 5
out["fused_auto_kernels_ms"] = {k: round(v, 3) for k, v in agg.items()}
per = len(tl) // 5
out["fused_auto_trace_last_call"] = [(n, round(st - tl[-per][2], 3), round(ms, 3)) for n, ms, st in tl[-per:]]
print(json.dumps(out))
