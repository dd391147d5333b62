"""ftk_lmhead_ce_fwd_static: the forward for callers that cannot allocate device memory.

Hosts such as a plugin inside a runtime that owns all of HBM, or a loop that must not trigger cudaMalloc's
device-wide synchronisation, have no workspace to pass.  This entry point keeps its scratch in a fixed array inside
the library image.  The test drives it through the C ABI twice in a row (the second call sees the first call's
leftovers in the shared scratch) and compares each result with the ordinary, workspace-taking entry point and with
the fp32 CPU oracle, then once more at the largest size the built-in scratch holds (M = FTK_STATIC_MAX_ROWS, where
the carve-up fills the array to the byte).

The file name sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu
LOSS_RTOL, LSE_ATOL = 1e-3, 2e-3  # the tolerances of tests/test_lmhead_gpu.py


def _inputs(M, d, V, seed, bias):
    g = torch.Generator().manual_seed(seed)
    H = torch.randn(M, d, generator=g).to(torch.bfloat16).cuda()
    W = (torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16).cuda()
    b = (torch.randn(V, generator=g) * 0.5).to(torch.bfloat16).cuda() if bias else None
    t = torch.randint(0, V, (M,), generator=g)
    t[torch.rand(M, generator=g) < 0.2] = -100
    return H, W, b, t.cuda()


def _call_static(L, H, W, b, t, lse, loss_sum, n_valid):
    M, d = H.shape
    return L.ftk_lmhead_ce_fwd_static(H.data_ptr(), W.data_ptr(), 0 if b is None else b.data_ptr(), t.data_ptr(), M, d,
                                      W.shape[0], lse.data_ptr(), loss_sum.data_ptr(), n_valid.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream)


def test_static_scratch_forward_matches_workspace_forward_twice(cuda_dev):
    from kubernetes_cloud_b200 import _lib, ops  # noqa: F401
    L = _lib.lib()
    # two different problems back to back: the second must not be disturbed by what the first left in the scratch
    for M, d, V, seed, bias in ((1000, 256, 5003, 11, True), (300, 128, 777, 12, False)):
        H, W, b, t = _inputs(M, d, V, seed, bias)
        want_sum, want_n, want_lse = torch.ops.ftk.lmhead_ce_fwd(H, W, b, t)  # existing entry point
        lse = torch.empty(M, dtype=torch.float32, device="cuda")  # outputs are the caller's, made before the call
        loss_sum = torch.empty(1, dtype=torch.float32, device="cuda")
        n_valid = torch.empty(1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0] if (M, d) == (300, 128) else None
        _lib.check(_call_static(L, H, W, b, t, lse, loss_sum, n_valid), "ftk_lmhead_ce_fwd_static")
        torch.cuda.synchronize()
        if free_before is not None:  # second call: set-up is done, the device's free memory must not move
            assert torch.cuda.mem_get_info()[0] == free_before
        assert int(n_valid) == int(want_n)
        assert abs(float(loss_sum) - float(want_sum)) <= LOSS_RTOL * abs(float(want_sum))
        assert float((lse - want_lse).abs().max()) <= LSE_ATOL
        # and against the checker that shares no code with either entry point
        ref = O.lmhead_ce_reference(H, W, b, t, want_grads=False)
        assert int(n_valid) == ref["n_valid"]
        assert abs(float(loss_sum) / int(n_valid) - float(ref["loss"])) <= LOSS_RTOL * abs(float(ref["loss"]))
        assert float((lse.cpu() - ref["lse"]).abs().max()) <= LSE_ATOL


def test_static_scratch_forward_at_its_row_limit(cuda_dev):
    """M = 16,384 rows is the most the built-in scratch holds: every byte of the array is in use.  Run it twice with
    different inputs (the second call starts from the first call's leftovers) and check both against the oracle on
    sampled rows and against the workspace entry point in full."""
    from kubernetes_cloud_b200 import _lib, ops  # noqa: F401
    L = _lib.lib()
    M, d, V = 16384, 256, 50257
    for seed in (21, 22):
        H, W, b, t = _inputs(M, d, V, seed, bias=(seed == 22))
        want_sum, want_n, want_lse = torch.ops.ftk.lmhead_ce_fwd(H, W, b, t)
        lse = torch.empty(M, dtype=torch.float32, device="cuda")
        loss_sum = torch.empty(1, dtype=torch.float32, device="cuda")
        n_valid = torch.empty(1, dtype=torch.int32, device="cuda")
        _lib.check(_call_static(L, H, W, b, t, lse, loss_sum, n_valid), "ftk_lmhead_ce_fwd_static")
        torch.cuda.synchronize()
        assert int(n_valid) == int(want_n) == int((t >= 0).sum())
        assert abs(float(loss_sum) - float(want_sum)) <= LOSS_RTOL * abs(float(want_sum))
        assert float((lse - want_lse).abs().max()) <= LSE_ATOL
        rows = torch.arange(0, M, 61, device="cuda")
        ref = O.lmhead_ce_reference(H[rows], W, b, t[rows], want_grads=False)
        assert float((lse[rows].cpu() - ref["lse"]).abs().max()) <= LSE_ATOL


def test_static_scratch_forward_rejects_sizes_over_its_limit(cuda_dev):
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    M = 16384 + 1  # FTK_STATIC_MAX_ROWS + 1
    H, W, b, t = _inputs(M, 64, 128, 13, False)
    lse = torch.empty(M, dtype=torch.float32, device="cuda")
    loss_sum = torch.empty(1, dtype=torch.float32, device="cuda")
    n_valid = torch.empty(1, dtype=torch.int32, device="cuda")
    rc = _call_static(L, H, W, b, t, lse, loss_sum, n_valid)
    assert rc == -3  # FTK_EWORKSPACE
    msg = L.ftk_last_error().decode()
    assert "16384" in msg and "ftk_lmhead_ce_fwd" in msg
