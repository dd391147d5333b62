#!/usr/bin/env python
"""Diagnostic (development tool): does the NCCL all-reduce of dW run beside the next forward?  Records CUDA events on
both streams against a common origin and prints a small timeline.  torchrun --nproc-per-node 2 tools/overlap_probe.py"""
import ctypes, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import _lib, ops  # noqa

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
reserve = int(os.environ.get("RESERVE", "0"))
if reserve:
    os.environ.setdefault("NCCL_MAX_CTAS", str(reserve))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
L = _lib.lib()
buf = ctypes.create_string_buffer(128)
box = [None]
if rank == 0:
    _lib.check(L.ftk_dp_unique_id(buf), "uid"); box = [bytes(buf.raw)]
dist.broadcast_object_list(box, src=0)
h = ctypes.c_void_p()
_lib.check(L.ftk_dp_init(ctypes.create_string_buffer(box[0], 128), rank, world, ctypes.byref(h)), "init")
M, d, V = 16384, 2048, 50257
g = torch.Generator(device=dev).manual_seed(rank)
H = torch.randn(M, d, device=dev, generator=g).to(torch.bfloat16)
W = (torch.randn(V, d, device=dev, generator=g) * 0.02).to(torch.bfloat16)
t = torch.randint(0, V, (M,), device=dev, generator=g)
dW = torch.randn(V, d, device=dev, generator=g).to(torch.bfloat16)
main = torch.cuda.current_stream(dev)
side = torch.cuda.Stream(dev, priority=-1)
E = lambda: torch.cuda.Event(enable_timing=True)

def allreduce_alone():
    a, b = E(), E()
    a.record(main)
    _lib.check(L.ftk_dp_allreduce(h, dW.data_ptr(), dW.numel(), _lib.FTK_BF16, 0.5, main.cuda_stream), "ar")
    _lib.check(L.ftk_dp_join(h, main.cuda_stream), "join")
    b.record(main); torch.cuda.synchronize()
    return a.elapsed_time(b)

for _ in range(3):
    allreduce_alone(); torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
torch.cuda.synchronize(); dist.barrier()
alone = min(allreduce_alone() for _ in range(3))
f0, f1 = E(), E()
f0.record(main); torch.ops.ftk.lmhead_ce_fwd(H, W, None, t); f1.record(main); torch.cuda.synchronize()
fwd_alone = f0.elapsed_time(f1)
# overlapped: issue the all-reduce (comm stream) then the forward (main stream), join afterwards
res = []
for rep in range(3):
    dist.barrier(); torch.cuda.synchronize()
    o, f0, f1, j1 = E(), E(), E(), E()
    o.record(main)
    _lib.check(L.ftk_dp_allreduce(h, dW.data_ptr(), dW.numel(), _lib.FTK_BF16, 0.5, main.cuda_stream), "ar")
    if reserve: L.ftk_set_sm_reserve(reserve)
    f0.record(main); torch.ops.ftk.lmhead_ce_fwd(H, W, None, t); f1.record(main)
    if reserve: L.ftk_set_sm_reserve(0)
    _lib.check(L.ftk_dp_join(h, main.cuda_stream), "join"); j1.record(main)
    torch.cuda.synchronize()
    res.append((o.elapsed_time(f0), o.elapsed_time(f1), o.elapsed_time(j1)))
if rank == 0:
    print(f"reserve={reserve} NCCL_MAX_CTAS={os.environ.get('NCCL_MAX_CTAS')} allreduce alone {alone:.3f} ms, fwd alone {fwd_alone:.3f} ms")
    for r in res:
        print(f"   fwd start {r[0]:.3f}  fwd end {r[1]:.3f}  joined {r[2]:.3f}   (serial would be {alone + fwd_alone:.3f})")
# steady-state loop shaped like bench.py's step: fwd -> join(previous all-reduce) -> bwd -> all-reduce(dW)
for mode in ("prescale", "noscale"):
    scale = 0.5 if mode == "prescale" else 1.0
    ls, nv, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
    times = []
    for it in range(12):
        a, b, c, e = E(), E(), E(), E()
        a.record(main)
        if reserve: L.ftk_set_sm_reserve(reserve)
        ls, nv, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
        if reserve: L.ftk_set_sm_reserve(0)
        b.record(main)
        _lib.check(L.ftk_dp_join(h, main.cuda_stream), "join")
        c.record(main)
        dH, dWn, _ = torch.ops.ftk.lmhead_ce_bwd(H, W, None, t, lse, None, nv, 1.0, False)
        e.record(main)
        _lib.check(L.ftk_dp_allreduce(h, dWn.data_ptr(), dWn.numel(), _lib.FTK_BF16, scale, main.cuda_stream), "ar")
        times.append((a, b, c, e, dWn))
    _lib.check(L.ftk_dp_join(h, main.cuda_stream), "join")
    torch.cuda.synchronize()
    if rank == 0:
        tt = times[4:]
        f = sum(x[0].elapsed_time(x[1]) for x in tt) / len(tt)
        j = sum(x[1].elapsed_time(x[2]) for x in tt) / len(tt)
        bw = sum(x[2].elapsed_time(x[3]) for x in tt) / len(tt)
        step = times[4][0].elapsed_time(times[-1][3]) / (len(times) - 5)
        print(f"   loop[{mode}] fwd {f:.3f}  join-wait {j:.3f}  bwd {bw:.3f}  step {step:.3f}")
L.ftk_dp_destroy(h); dist.destroy_process_group()
