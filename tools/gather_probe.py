#!/usr/bin/env python
"""Experiment (development tool): bulk gather bandwidth for kernel variants and launch patterns."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import _lib, ops  # noqa
T, n_ctx, B = 2048, 4096, 32768
dev = torch.device("cuda")
tok = torch.randint(0, 50257, (n_ctx * T,), device=dev, dtype=torch.int32).to(torch.int16)
idx = torch.randint(0, n_ctx, (B,), device=dev)
ids = torch.empty(B, T, dtype=torch.int64, device=dev); lab = torch.empty_like(ids)
msk = torch.empty(B, T, dtype=torch.bool, device=dev); nv = torch.empty(B, dtype=torch.int32, device=dev)
L = _lib.lib(); s = torch.cuda.current_stream().cuda_stream
def launch():
    _lib.check(L.ftk_gather_windows(tok.data_ptr(), n_ctx, T, idx.data_ptr(), B, 50256, 1, 1, ids.data_ptr(), lab.data_ptr(), msk.data_ptr(), nv.data_ptr(), s), "g")
for var in ("", "cs", "direct"):
    if var: os.environ["FTK_GATHER_VARIANT"] = var
    else: os.environ.pop("FTK_GATHER_VARIANT", None)
    for _ in range(3): launch()
    torch.cuda.synchronize()
    for reps in (1, 10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps): launch()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        print(f"variant={var or 'smem':6s} reps={reps:2d}  {ms*1e3:7.1f} us  {19.0*B*T/ms/1e6:7.1f} GB/s")
