"""Host-side logic that needs no GPU: ordering, and the bucketed gradient sync over gloo with world_size 2."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O


def test_train_order_and_rank_slice_match_oracle_and_fixtures(golden_dir):
    from kubernetes_cloud_b200.ordering import epoch_batches, rank_slice, train_order
    g = np.load(os.path.join(golden_dir, "order_n4096_s42.npz"))
    np.testing.assert_array_equal(train_order(4096, 42, True).numpy(), g["indices"])
    np.testing.assert_array_equal(train_order(7, 1, False).numpy(), np.arange(7))
    for n, world in ((10, 4), (33, 8), (3, 8), (4096, 8)):
        order = train_order(n, 42)
        for r in range(world):
            np.testing.assert_array_equal(rank_slice(order, r, world).numpy(), O.rank_slice(order.numpy(), r, world))
    with pytest.raises(ValueError):
        rank_slice(train_order(4), 4, 4)
    # every rank sees the same number of batches; together they cover the permutation
    seen = []
    for r in range(4):
        b = list(epoch_batches(10, 2, seed=42, rank=r, world=4))
        assert [x.numel() for x in b] == [2, 1]
        seen += [int(v) for x in b for v in x]
    assert set(seen) == set(range(10)) and len(seen) == 12  # 2 wrap-around pads


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kubernetes_cloud_b200.dp import GradSync
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 8))
    sync = GradSync(model.parameters(), rank, world, bucket_bytes=1024, backend="gloo")  # several small buckets
    assert len(sync.buckets) >= 2
    g = torch.Generator().manual_seed(100 + rank)
    xs = [torch.randn(4, 16, generator=g) for _ in range(3)]
    # 3 micro-batches: accumulate locally on the first two, reduce on the last (gradient accumulation)
    sync.zero_grad()
    with sync.no_sync():
        for x in xs[:2]:
            model(x).pow(2).mean().backward()
    assert sync.launched == 0
    model(xs[2]).pow(2).mean().backward()
    sync.finish()
    assert sync.launched == len(sync.buckets)
    grads = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    # expected: mean over ranks of each rank's accumulated gradient, recomputed without GradSync
    torch.manual_seed(0)
    ref = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 8))
    total = None
    for r in range(world):
        gr = torch.Generator().manual_seed(100 + r)
        ref.zero_grad()
        for _ in range(3):
            ref(torch.randn(4, 16, generator=gr)).pow(2).mean().backward()
        v = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
        total = v.clone() if total is None else total + v
    want = total / world
    out[rank] = float((grads - want).abs().max())
    # second optimizer step reuses the buckets
    sync.zero_grad()
    model(xs[0]).pow(2).mean().backward()
    sync.finish()
    assert sync.launched == 2 * len(sync.buckets)
    dist.destroy_process_group()


def test_grad_sync_gloo_world2():
    world = 2
    port = _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        assert len(out) == world
        for r in range(world):
            assert out[r] < 1e-6


def test_grad_sync_rejects_cpu_params_for_ftk_backend():
    from kubernetes_cloud_b200 import FtkError
    from kubernetes_cloud_b200.dp import GradSync
    with pytest.raises(FtkError):
        GradSync(torch.nn.Linear(4, 4).parameters(), 0, 1, backend="ftk")
