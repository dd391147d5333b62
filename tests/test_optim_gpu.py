"""Fused AdamW / grad-norm / clip kernels through the C ABI vs torch.optim.AdamW on CPU (oracle.adamw_reference)."""
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _model(dtype):
    torch.manual_seed(3)
    m = torch.nn.Sequential(torch.nn.Linear(48, 96), torch.nn.Tanh(), torch.nn.Linear(96, 17))  # odd sizes: tails
    return m.cuda().to(dtype)


@pytest.mark.parametrize("dtype,main_grads", [(torch.float32, False), (torch.bfloat16, False), (torch.bfloat16, True)])
def test_fused_adamw_matches_torch_adamw(cuda_dev, dtype, main_grads):
    from kubernetes_cloud_b200.dp import GradSync
    from kubernetes_cloud_b200.optim import FusedAdamW, WarmupLR
    model = _model(dtype)
    params0 = [p.detach().clone() for p in model.parameters()]
    sync = GradSync(model.parameters(), 0, 1, bucket_bytes=8 << 10, backend="ftk", main_grads=main_grads)
    assert sync.buckets[0].flat.dtype == (torch.float32 if main_grads else dtype)
    sched = WarmupLR(max_lr=1e-2, warmup_steps=4)
    opt = FusedAdamW(sync, lr=1e-2, weight_decay=0.01, max_grad_norm=0.5, schedule=sched)
    g = torch.Generator().manual_seed(5)
    grads_per_step, lrs, norms = [], [], []
    for step in range(6):
        sync.zero_grad()
        grads = [(torch.randn(p.shape, generator=g) * (3.0 if step % 2 else 0.05)).to(dtype) for p in params0]
        for p, gr in zip(model.parameters(), grads):
            (p.main_grad if main_grads else p.grad).copy_(gr.cuda())  # gradients land in the bucket views
        lrs.append(opt.current_lr())
        opt.step()
        norms.append(float(opt.grad_norm()))
        grads_per_step.append(grads)
    torch.cuda.synchronize()
    want, want_norms = O.adamw_reference(params0, grads_per_step, lrs, weight_decay=0.01, max_grad_norm=0.5)
    for a, b in zip(norms, want_norms):
        assert abs(a - b) <= 1e-4 * b
    ref_src = opt.master if dtype == torch.bfloat16 else [None] * len(sync.buckets)
    for p, w in zip(model.parameters(), want):
        if dtype == torch.float32:
            assert O.rel_err(p.detach().float(), w) < 2e-6
        else:
            assert O.rel_err(p.detach().float(), w) < 4e-3  # bf16 rounding of the stored parameter
    if dtype == torch.bfloat16:  # the fp32 master copy tracks the fp32 reference closely
        flat_want = []
        for b in sync.buckets:
            for p in b.params:
                idx = [i for i, q in enumerate(model.parameters()) if q is p][0]
                flat_want.append((p, want[idx]))
        off_by_bucket = {}
        for bi, b in enumerate(sync.buckets):
            off = 0
            for p in b.params:
                idx = [i for i, q in enumerate(model.parameters()) if q is p][0]
                got = opt.master[bi][off:off + p.numel()].view_as(p)
                assert O.rel_err(got, want[idx]) < 2e-5
                off += (p.numel() + 7) // 8 * 8
    sync.close()


def test_sqnorm_and_clip_c_abi(cuda_dev):
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    s = torch.cuda.current_stream().cuda_stream
    ws = torch.empty(L.ftk_grad_sqnorm_workspace(), dtype=torch.uint8, device="cuda")
    out = torch.zeros(1, device="cuda")
    coef = torch.zeros(1, device="cuda")
    total = 0.0
    for i, (dt, code, n) in enumerate(((torch.float32, _lib.FTK_F32, 1_000_003), (torch.bfloat16, _lib.FTK_BF16, 777_777))):
        x = torch.randn(n, device="cuda").to(dt)
        total += float((x.double() ** 2).sum())
        _lib.check(L.ftk_grad_sqnorm(x.data_ptr(), n, code, out.data_ptr(), int(i > 0), ws.data_ptr(), ws.numel(), s),
                   "ftk_grad_sqnorm")
    torch.cuda.synchronize()
    assert abs(float(out) - total) <= 1e-5 * total
    again = torch.zeros(1, device="cuda")  # deterministic: same bits on a second run
    x = torch.randn(333_333, device="cuda")
    for o in (out, again):
        _lib.check(L.ftk_grad_sqnorm(x.data_ptr(), x.numel(), _lib.FTK_F32, o.data_ptr(), 0, ws.data_ptr(), ws.numel(), s),
                   "ftk_grad_sqnorm")
    torch.cuda.synchronize()
    assert float(out) == float(again)
    _lib.check(L.ftk_clip_coef(out.data_ptr(), 1.0, coef.data_ptr(), s), "ftk_clip_coef")
    torch.cuda.synchronize()
    assert abs(float(coef) - min(1.0, 1.0 / (float(out) ** 0.5 + 1e-6))) < 1e-6
    rc = L.ftk_adamw_step(x.data_ptr(), x.data_ptr(), None, x.data_ptr(), x.data_ptr(), 8, _lib.FTK_BF16, _lib.FTK_BF16,
                          1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1.0, None, s)
    assert rc == -1 and b"master" in L.ftk_last_error()
