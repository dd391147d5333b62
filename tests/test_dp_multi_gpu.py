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


def _ft_worker(rank, world, port, argv, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    from kubernetes_cloud_b200 import finetune
    captured = {}
    orig_close = None
    # grab the parameters before main() tears everything down: wrap GradSync.close
    from kubernetes_cloud_b200 import dp as dpmod
    orig_close = dpmod.GradSync.close

    def close(self):
        captured["sum"] = float(sum(p.detach().double().sum() for p in self.params))
        captured["abs"] = float(sum(p.detach().double().abs().sum() for p in self.params))
        orig_close(self)

    dpmod.GradSync.close = close
    res = finetune.main(argv)
    out[rank] = {"losses": [h["loss"] for h in res["history"]], **captured}


def test_finetune_cli_data_parallel_two_gpus(cuda_dev, tmp_path):
    """Plain DP end to end: both ranks see different batches, gradients are averaged over NCCL by GradSync, the fused
    optimizer keeps the replicas bit-identical."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    import torch.multiprocessing as mp
    from transformers import GPTNeoConfig
    from kubernetes_cloud_b200 import write_token_file
    T, V = 128, 512
    cfg_dir = tmp_path / "tiny"
    cfg_dir.mkdir()
    GPTNeoConfig(vocab_size=V, hidden_size=64, num_layers=2, num_heads=4, max_position_embeddings=T,
                 attention_types=[[["global", "local"], 1]], window_size=32).save_pretrained(str(cfg_dir))
    pattern = np.random.default_rng(0).integers(0, V - 1, size=37)
    path = str(tmp_path / "d.tokens")
    write_token_file(path, np.tile(pattern, 64 * T // 37 + 1)[: 64 * T], T, pad_id=V - 1)
    argv = ["--run-name", "dp", "--model", str(cfg_dir), "--dataset", path, "--context-size", str(T), "--bs", "2",
            "--gradients", "2", "--lr", "3e-3", "--max-steps", "6", "--pad-token-id", str(V - 1), "--eos-token-id",
            str(V - 1)]
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_ft_worker, args=(2, port, argv, out), nprocs=2, join=True)
        a, b = out[0], out[1]
        assert len(a["losses"]) == 6 and all(np.isfinite(a["losses"])) and all(np.isfinite(b["losses"]))
        assert a["losses"] != b["losses"]                       # different shards of the data
        assert a["sum"] == b["sum"] and a["abs"] == b["abs"]    # replicas stay identical
        assert a["losses"][-1] < a["losses"][0]
