#!/usr/bin/env python
"""G4 roofline: bus bandwidth of ftk_dp_allreduce (ncclAllReduce avg, bf16 and fp32) over bucket sizes, max over ranks.
torchrun --nproc-per-node N tools/allreduce_sweep.py   (reporting tool; results are copied into BASELINE.md)"""
import ctypes, json, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import _lib  # noqa

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
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
s = torch.cuda.current_stream(dev)
sizes_mb = [8, 32, 64, 128, 206, 256, 512, 1024, 2632]   # 206 MB = lm_head dW bf16, 2632 MB = all 1.3 B params bf16
for dt, code, esz in ((torch.bfloat16, _lib.FTK_BF16, 2), (torch.float32, _lib.FTK_F32, 4)):
    for mb in sizes_mb:
        n = mb * (1 << 20) // esz
        x = torch.ones(n, dtype=dt, device=dev)
        def once():
            _lib.check(L.ftk_dp_allreduce(h, x.data_ptr(), n, code, 1.0 / world, s.cuda_stream), "ar")
            _lib.check(L.ftk_dp_join(h, s.cuda_stream), "join")
        for _ in range(3): once()
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 10 if mb <= 512 else 4
        a.record(s)
        for _ in range(iters): once()
        b.record(s); torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(b) / iters], device=dev, dtype=torch.float64)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms)
        ok = bool(torch.all(x[:1024].float() == 1.0))
        if rank == 0:
            nbytes = n * esz
            print(json.dumps({"world": world, "dtype": str(dt).split(".")[-1], "MB": mb, "ms": round(ms, 4),
                              "algbw_GBs": round(nbytes / ms / 1e6, 1),
                              "busbw_GBs": round(2 * (world - 1) / world * nbytes / ms / 1e6, 1), "avg_ok": ok}), flush=True)
        del x
L.ftk_dp_destroy(h); dist.destroy_process_group()
