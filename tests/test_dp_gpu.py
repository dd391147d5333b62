"""G4 gradient sync through the C ABI's NCCL path on the GPUs that are present (1 on the test box)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("main_grads", [True, False])
def test_grad_sync_ftk_world1_prescale_and_overlap_stream(cuda_dev, main_grads):
    """bf16 parameters; fp32 main gradients (the default, as the reference reduces in fp32) and bf16 buckets."""
    from kubernetes_cloud_b200.dp import GradSync
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.GELU(), torch.nn.Linear(128, 64)).cuda().to(torch.bfloat16)
    ref = [p.detach().clone().float().requires_grad_(True) for p in model.parameters()]
    sync = GradSync(model.parameters(), 0, 1, bucket_bytes=16 << 10, backend="ftk", average=True,
                    main_grads=main_grads)
    assert sync.main_grads == main_grads
    x = torch.randn(32, 64, device="cuda", dtype=torch.bfloat16)
    sync.zero_grad()
    with sync.no_sync():
        model(x).float().pow(2).mean().backward()
    assert sync.launched == 0
    model(x).float().pow(2).mean().backward()
    sync.finish()
    torch.cuda.synchronize()
    assert sync.launched == len(sync.buckets) >= 2
    # world 1: the reduced gradient is the locally accumulated one (two identical micro-batches)
    model2 = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.GELU(), torch.nn.Linear(128, 64)).cuda().to(torch.bfloat16)
    with torch.no_grad():
        for p, q in zip(model2.parameters(), ref):
            p.copy_(q)
    for _ in range(2):
        model2(x).float().pow(2).mean().backward()
    for p, q in zip(model.parameters(), model2.parameters()):
        got = p.main_grad if main_grads else p.grad
        assert got.dtype == (torch.float32 if main_grads else torch.bfloat16)
        assert torch.allclose(got.float(), q.grad.float(), rtol=2e-2, atol=1e-4)
        assert (p.grad is None) == main_grads
    sync.close()


def test_dp_allreduce_c_abi_scales_in_place(cuda_dev):
    import ctypes
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    buf = ctypes.create_string_buffer(128)
    _lib.check(L.ftk_dp_unique_id(buf), "ftk_dp_unique_id")
    h = ctypes.c_void_p()
    _lib.check(L.ftk_dp_init(buf, 0, 1, ctypes.byref(h)), "ftk_dp_init")
    s = torch.cuda.current_stream().cuda_stream
    for dt, code, n in ((torch.float32, _lib.FTK_F32, 1003), (torch.bfloat16, _lib.FTK_BF16, 4099)):
        x = torch.arange(n, device="cuda", dtype=torch.float32).to(dt)
        want = (x.float() * 0.25).to(dt)
        _lib.check(L.ftk_dp_allreduce(h, x.data_ptr(), n, code, 0.25, s), "ftk_dp_allreduce")
        _lib.check(L.ftk_dp_join(h, s), "ftk_dp_join")
        torch.cuda.synchronize()
        assert torch.equal(x, want)
    _lib.check(L.ftk_dp_destroy(h), "ftk_dp_destroy")


def test_lazy_zero_of_the_fused_loss_gradient(cuda_dev):
    """GradSync.zero_grad(overwritten_by_fused_loss=W) zeroes everything in W's bucket except W's own gradient and
    marks that gradient; the next fused loss call overwrites it (no stale content survives), later calls add."""
    from oracle import oracle as O
    from kubernetes_cloud_b200 import ops
    from kubernetes_cloud_b200.dp import GradSync
    M, d, V = 300, 64, 500
    g = torch.Generator().manual_seed(4)
    other = torch.nn.Parameter(torch.randn(37, device="cuda").to(torch.bfloat16))
    W = torch.nn.Parameter((torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16).cuda())
    tail = torch.nn.Parameter(torch.randn(11, device="cuda").to(torch.bfloat16))
    sync = GradSync([other, W, tail], 0, 1, bucket_bytes=1 << 30, backend="ftk")  # one bucket, fp32 main gradients
    assert len(sync.buckets) == 1 and sync.main_grads
    sync.buckets[0].flat.fill_(7.0)  # stale content everywhere
    sync.zero_grad(overwritten_by_fused_loss=W)
    assert float(other.main_grad.abs().max()) == 0.0 and float(tail.main_grad.abs().max()) == 0.0
    assert float(W.main_grad.min()) == 7.0  # not touched by the zeroing pass
    total = torch.zeros(V, d)
    for k in range(2):
        H = torch.randn(M, d, generator=g).to(torch.bfloat16)
        t = torch.randint(0, V, (M,), generator=g)
        total += O.lmhead_ce_reference(H, W.detach().cpu(), None, t)["dW"]
        with sync.no_sync():
            ops.fused_lmhead_ce(H.cuda().requires_grad_(True), W, None, t.cuda()).backward()
    assert O.rel_err(W.main_grad, total) < 5e-3  # first call overwrote the 7s, second call added
    sync.close()
