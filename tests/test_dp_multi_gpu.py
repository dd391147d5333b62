"""G4 on two real GPUs (skipped where fewer are visible): the three reduction forms of ftk_dp_allreduce."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out):
    import ctypes
    import torch.distributed as dist
    from kubernetes_cloud_b200 import _lib
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # only carries the 128-byte unique id
    L = _lib.lib()
    buf = ctypes.create_string_buffer(128)
    box = [None]
    if rank == 0:
        _lib.check(L.ftk_dp_unique_id(buf), "ftk_dp_unique_id")
        box = [bytes(buf.raw)]
    dist.broadcast_object_list(box, src=0)
    h = ctypes.c_void_p()
    _lib.check(L.ftk_dp_init(ctypes.create_string_buffer(box[0], 128), rank, world, ctypes.byref(h)), "ftk_dp_init")
    s = torch.cuda.current_stream().cuda_stream
    errs = []
    for dt, code in ((torch.float32, _lib.FTK_F32), (torch.bfloat16, _lib.FTK_BF16)):
        for scale in (1.0, 1.0 / world, 0.25):  # sum, ncclAvg, pre-multiplied sum
            n = 1_000_003
            g = [torch.Generator().manual_seed(7 + r) for r in range(world)]
            xs = [torch.randn(n, generator=gg).to(dt) for gg in g]
            x = xs[rank].cuda()
            _lib.check(L.ftk_dp_allreduce(h, x.data_ptr(), n, code, scale, s), "ftk_dp_allreduce")
            _lib.check(L.ftk_dp_join(h, s), "ftk_dp_join")
            torch.cuda.synchronize()
            want = sum(v.double() * scale for v in xs)
            err = float((x.cpu().double() - want).abs().max() / want.abs().max())
            errs.append((str(dt), scale, err))
    _lib.check(L.ftk_dp_destroy(h), "ftk_dp_destroy")
    dist.destroy_process_group()
    out[rank] = errs


def test_allreduce_forms_on_two_gpus(cuda_dev):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
        for r in range(2):
            for dt, scale, err in out[r]:
                assert err < (1e-6 if "float32" in dt else 1.5e-2), (dt, scale, err)
