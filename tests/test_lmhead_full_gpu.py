"""Fused lm_head + cross-entropy at the BASELINE.json shapes: dW / db / dH against references that do not come from
the kernel under test, the accumulate paths of the C ABI, and BASELINE configs[0] against the HF model's own numbers.

Tolerances (BASELINE.json north_star): loss 1e-3 relative, gradients 1e-2 relative (norm-wise; bf16 inputs, fp32
accumulate)."""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu
LOSS_RTOL, GRAD_RTOL = 1e-3, 1e-2


def gpu_inputs(M, d, V, seed, bias, wstd=0.02, ignore_every=2048):
    g = torch.Generator(device="cuda").manual_seed(seed)
    H = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(V, d, device="cuda", generator=g) * wstd).to(torch.bfloat16)
    b = (torch.randn(V, device="cuda", generator=g) * 0.5).to(torch.bfloat16) if bias else None
    t = torch.randint(0, V, (M,), device="cuda", generator=g)
    t[ignore_every - 1::ignore_every] = -100  # the last position of every context has no next token
    t[5] = 0
    t[6] = V - 1
    return H, W, b, t


def dense_lse(H, W, b, rows_per_pass=2048):
    """Row log-sum-exp of H.W^T (+b) by plain fp32 torch ops on the GPU (no TF32), a slab of rows at a time."""
    assert not torch.backends.cuda.matmul.allow_tf32
    Wf = W.float()
    out = torch.empty(H.shape[0], dtype=torch.float32, device=H.device)
    for r0 in range(0, H.shape[0], rows_per_pass):
        z = H[r0:r0 + rows_per_pass].float() @ Wf.t()
        if b is not None:
            z += b.float()
        out[r0:r0 + rows_per_pass] = torch.logsumexp(z, -1)
    return out


def sampled_vocab_rows(V, t, n_random=48, seed=0):
    """Row 0, row V-1, rows across the tail tile, targets that occur (the one-hot term) and random rows."""
    rng = np.random.default_rng(seed)
    tail0 = V - (V % 256 or 256)
    tt = t[t >= 0]
    picks = [0, V - 1, tail0, tail0 + 1, (tail0 + V) // 2, V - 2, 255, 256, 257]
    picks += tt[torch.from_numpy(rng.integers(0, tt.numel(), 8)).to(tt.device)].tolist()
    picks += rng.integers(0, V, n_random).tolist()
    return torch.tensor(sorted(set(int(p) for p in picks)), device=t.device)


def dense_dw_rows(H, W, b, t, lse, vs):
    """dW[v, :] = sum_m (exp(z[m, v] - lse[m]) - [t[m] == v]) * H[m, :] / n_valid over valid rows m, and
    db[v] = the same sum without H, for the sampled vocabulary rows vs; fp32 torch ops."""
    valid = t >= 0
    n_valid = int(valid.sum())
    z = H.float() @ W[vs].float().t()  # [M, len(vs)]
    if b is not None:
        z += b[vs].float()
    p = torch.exp(z - lse[:, None])
    p -= (t[:, None] == vs[None, :]).float()
    p[~valid] = 0
    p /= n_valid
    return p.t() @ H.float(), p.sum(0)


@pytest.mark.parametrize("M,d,V,bias", [(16384, 2048, 50257, False),   # GPT-Neo-1.3B head, the bench workload
                                        (16384, 768, 50257, False),    # GPT-2 124M: 16 x 1024
                                        (16384, 4096, 50400, True)])   # GPT-J-6B: untied, bias, 196 tiles + 224
def test_gradients_at_baseline_shapes(cuda_dev, M, d, V, bias):
    from kubernetes_cloud_b200 import ops  # noqa: F401
    H, W, b, t = gpu_inputs(M, d, V, seed=M + d + V, bias=bias)
    lse_ref = dense_lse(H, W, b)
    loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, b, t)
    assert int(n_valid) == int((t >= 0).sum())
    assert float((lse - lse_ref).abs().max()) < 2e-3
    valid = t >= 0
    zt = (H[valid].float() * W[t[valid]].float()).sum(-1)
    if b is not None:
        zt += b[t[valid]].float()
    ref_loss = (lse_ref[valid] - zt).double().sum()
    assert abs(float(loss_sum) - float(ref_loss)) <= 1e-4 * abs(float(ref_loss))
    dH, dW, db = torch.ops.ftk.lmhead_ce_bwd(H, W, b, t, lse, None, n_valid, 1.0, False)
    vs = sampled_vocab_rows(V, t)
    dW_ref, db_ref = dense_dw_rows(H, W, b, t, lse_ref, vs)
    assert O.rel_err(dW[vs].float(), dW_ref) < GRAD_RTOL
    # row by row as well: a single wrong vocabulary row must not hide inside the norm of the others
    per_row = (dW[vs].float() - dW_ref).norm(dim=1) / dW_ref.norm(dim=1).clamp_min(1e-12)
    assert float(per_row.max()) < 2 * GRAD_RTOL, (vs[per_row.argmax()].item(), float(per_row.max()))
    if bias:
        assert O.rel_err(db[vs], db_ref) < GRAD_RTOL
    # dH on sampled rows (every 257th, plus the ignored ones)
    rows = torch.cat([torch.arange(0, M, 257, device="cuda"), torch.arange(2047, M, 2048, device="cuda")])
    z = H[rows].float() @ W.float().t()
    if b is not None:
        z += b.float()
    p = torch.exp(z - lse_ref[rows, None])
    tr = t[rows]
    ok = tr >= 0
    p[ok, tr[ok]] -= 1
    p[~ok] = 0
    dH_ref = (p / float(n_valid)) @ W.float()
    assert O.rel_err(dH[rows].float(), dH_ref) < GRAD_RTOL
    assert float(dH[~valid].abs().max()) == 0.0


def test_full_cpu_oracle_at_the_headline_shape(cuda_dev):
    """One complete run of the fp32 CPU oracle at (16384, 2048, 50257): loss, lse, dH, dW in full, not sampled.
    About 10 TFLOP of fp32 on the host (tens of seconds) and ~15 GB of host memory."""
    from kubernetes_cloud_b200 import ops  # noqa: F401
    M, d, V = 16384, 2048, 50257
    H, W, _, t = gpu_inputs(M, d, V, seed=1, bias=False)
    loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
    dH, dW, _ = torch.ops.ftk.lmhead_ce_bwd(H, W, None, t, lse, None, n_valid, 1.0, True)
    torch.cuda.synchronize()
    want = O.lmhead_ce_reference(H, W, None, t)
    assert int(n_valid) == want["n_valid"]
    loss = float(loss_sum) / int(n_valid)
    assert abs(loss - float(want["loss"])) <= LOSS_RTOL * abs(float(want["loss"]))
    assert float((lse.cpu() - want["lse"]).abs().max()) < 2e-3
    assert O.rel_err(dH.float(), want["dH"]) < GRAD_RTOL
    assert O.rel_err(dW, want["dW"]) < GRAD_RTOL


def _bwd_ctypes(H, W, b, t, lse, n_valid, dW, dw_dtype, accumulate, slab_rows, dbias=None):
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    M, d = H.shape
    V = W.shape[0]
    dH = torch.empty_like(H)
    ws_bytes = L.ftk_lmhead_ce_bwd_workspace(M, d, V, slab_rows)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    rc = L.ftk_lmhead_ce_bwd(H.data_ptr(), W.data_ptr(), 0 if b is None else b.data_ptr(), t.data_ptr(), lse.data_ptr(),
                             M, d, V, 1.0, 0, n_valid.data_ptr(), dH.data_ptr(), dW.data_ptr(), dw_dtype, accumulate,
                             0 if dbias is None else dbias.data_ptr(), ws.data_ptr(), ws_bytes,
                             torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ftk_lmhead_ce_bwd")
    torch.cuda.synchronize()
    return dH


@pytest.mark.parametrize("dw_dtype", ["bf16", "f32"])
@pytest.mark.parametrize("slab_rows", [0, 256])
def test_accumulate_into_a_prefilled_dw(cuda_dev, dw_dtype, slab_rows):
    """accumulate=1 through the C ABI (tied embeddings, gradient accumulation): dW_out = dW_before + dZ^T.H, for both
    gradient dtypes, with one slab and with five (M = 1200 rows over 256-row slabs: the later slabs always add)."""
    from kubernetes_cloud_b200 import _lib, ops  # noqa: F401
    M, d, V = 1200, 256, 5003
    g = torch.Generator().manual_seed(17)
    H = torch.randn(M, d, generator=g).to(torch.bfloat16)
    W = (torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16)
    b = (torch.randn(V, generator=g) * 0.5).to(torch.bfloat16)
    t = torch.randint(0, V, (M,), generator=g)
    t[torch.rand(M, generator=g) < 0.1] = -100
    want = O.lmhead_ce_reference(H, W, b, t)
    Hc, Wc, bc, tc = H.cuda(), W.cuda(), b.cuda(), t.cuda()
    _, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(Hc, Wc, bc, tc)
    tdt = torch.bfloat16 if dw_dtype == "bf16" else torch.float32
    code = _lib.FTK_BF16 if dw_dtype == "bf16" else _lib.FTK_F32
    before = (torch.randn(V, d, generator=g) * float(want["dW"].std())).to(tdt)  # same scale as the gradient itself
    db_before = torch.randn(V, generator=g) * float(want["dbias"].std())
    dW = before.cuda().clone()
    db = db_before.cuda().clone()
    dH = _bwd_ctypes(Hc, Wc, bc, tc, lse, n_valid, dW, code, 1, slab_rows, dbias=db)
    assert O.rel_err(dW.float(), before.float() + want["dW"]) < GRAD_RTOL
    assert O.rel_err(dW.float() - before.float().cuda(), want["dW"]) < (3e-2 if dw_dtype == "bf16" else GRAD_RTOL)
    assert O.rel_err(db, db_before + want["dbias"]) < GRAD_RTOL
    assert O.rel_err(dH.float(), want["dH"]) < GRAD_RTOL
    # accumulate=0 overwrites whatever was there
    dW2 = torch.full_like(dW, float("nan"))
    db2 = torch.full_like(db, float("nan"))
    _bwd_ctypes(Hc, Wc, bc, tc, lse, n_valid, dW2, code, 0, slab_rows, dbias=db2)
    assert O.rel_err(dW2.float(), want["dW"]) < GRAD_RTOL and O.rel_err(db2, want["dbias"]) < GRAD_RTOL


def test_autograd_accumulates_into_existing_grad_in_the_kernel(cuda_dev):
    """Two micro-batches through fused_lmhead_ce: the second backward adds into ``weight.grad`` inside the dW kernel
    (accumulate=1) instead of materialising a second [V, d] tensor; the sum matches the oracle's sum."""
    from kubernetes_cloud_b200 import ops
    M, d, V = 512, 128, 3001
    g = torch.Generator().manual_seed(23)
    W = (torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16)
    Wc = W.cuda().requires_grad_(True)
    total = torch.zeros(V, d)
    for k in range(2):
        H = torch.randn(M, d, generator=g).to(torch.bfloat16)
        t = torch.randint(0, V, (M,), generator=g)
        total += O.lmhead_ce_reference(H, W, None, t)["dW"]
        ops.fused_lmhead_ce(H.cuda(), Wc, None, t.cuda()).backward()
        if k == 0:
            first = Wc.grad
    assert Wc.grad is first  # same tensor, updated in place
    assert O.rel_err(Wc.grad.float(), total) < GRAD_RTOL
    # Megatron-style fp32 main gradient: dW lands in weight.main_grad in fp32, weight.grad stays untouched
    W2 = W.cuda().requires_grad_(True)
    W2.main_grad = torch.zeros(V, d, device="cuda")
    g = torch.Generator().manual_seed(23)
    _ = torch.randn(V, d, generator=g)
    for k in range(2):
        H = torch.randn(M, d, generator=g).to(torch.bfloat16)
        t = torch.randint(0, V, (M,), generator=g)
        ops.fused_lmhead_ce(H.cuda(), W2, None, t.cuda()).backward()
    assert W2.grad is None
    assert O.rel_err(W2.main_grad, total) < 5e-3


def test_config0_gptneo125m_against_the_hf_cpu_model(cuda_dev, golden_dir):
    """BASELINE.json configs[0]: GPT-Neo-125M (literal config), ctx 2048, batch 1.  The fixture holds the fp32 CPU HF
    model's hidden states at the lm_head input, its loss and its gradients at that boundary; the CUDA path gets the
    bf16 rounding of the same hidden states and weight.  North-star tolerances: loss 1e-3, gradients 1e-2."""
    from kubernetes_cloud_b200 import ops
    g = np.load(os.path.join(golden_dir, "config0_gptneo125m.npz"))
    W = O.gptneo125m_lm_head_weight(int(g["seed"]))
    np.testing.assert_allclose(O.weight_checksum(W), g["weight_checksum"], rtol=1e-9)
    labels = torch.from_numpy(g["labels"])
    targets = torch.nn.functional.pad(labels, (0, 1), value=-100)[..., 1:].reshape(-1)
    Hc = torch.from_numpy(g["hidden"]).cuda().to(torch.bfloat16).requires_grad_(True)
    Wc = W.cuda().to(torch.bfloat16).requires_grad_(True)
    loss = ops.fused_lmhead_ce(Hc, Wc, None, targets.cuda())
    loss.backward()
    assert abs(float(loss.detach()) - float(g["loss"])) <= LOSS_RTOL * float(g["loss"])
    h_rows = torch.from_numpy(g["h_rows"]).cuda()
    w_rows = torch.from_numpy(g["w_rows"]).cuda()
    assert O.rel_err(Hc.grad[h_rows].float(), torch.from_numpy(g["d_hidden_rows"])) < GRAD_RTOL
    assert O.rel_err(Wc.grad[w_rows].float(), torch.from_numpy(g["d_weight_rows"])) < GRAD_RTOL
    assert abs(float(Wc.grad.double().norm()) - float(g["d_weight_norm"])) <= GRAD_RTOL * float(g["d_weight_norm"])


def test_target_beyond_vocabulary_poisons_the_loss(cuda_dev):
    """A target id >= V (the uint16 token file can hold ids up to 65535) is a caller error; torch's cross_entropy
    asserts on the device.  Here the loss comes back NaN instead of silently reading stale scratch."""
    from kubernetes_cloud_b200 import ops  # noqa: F401
    M, d, V = 300, 64, 500
    H, W, _, t = gpu_inputs(M, d, V, seed=3, bias=False, ignore_every=64)
    t[17] = V + 3
    loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
    assert torch.isnan(loss_sum).all()


# ------------------------------------------------------------------------------------------------ fused fwd + bwd
def _fused_ctypes(H, W, b, t, dW, dw_code, accumulate, slab_rows, grad_scale=1.0, dbias=None, dw_event=0):
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    M, d = H.shape
    V = W.shape[0]
    lse = torch.empty(M, dtype=torch.float32, device="cuda")
    loss_sum = torch.empty(1, dtype=torch.float32, device="cuda")
    n_valid = torch.empty(1, dtype=torch.int32, device="cuda")
    dH = torch.empty_like(H)
    ws_bytes = L.ftk_lmhead_ce_fused_workspace(M, d, V, slab_rows)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    rc = L.ftk_lmhead_ce_fused(H.data_ptr(), W.data_ptr(), 0 if b is None else b.data_ptr(), t.data_ptr(), M, d, V,
                               grad_scale, 0, lse.data_ptr(), loss_sum.data_ptr(), n_valid.data_ptr(), dH.data_ptr(),
                               dW.data_ptr(), dw_code, accumulate, 0 if dbias is None else dbias.data_ptr(), dw_event,
                               ws.data_ptr(), ws_bytes, slab_rows, torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ftk_lmhead_ce_fused")
    torch.cuda.synchronize()
    return float(loss_sum) / max(int(n_valid), 1), int(n_valid), lse, dH


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("M,d,V,bias,slab", [(128, 64, 256, False, 0), (300, 64, 777, True, 0), (77, 192, 50, False, 0),
                                             (1000, 256, 5003, True, 256), (1200, 128, 3001, False, 512),
                                             (2048, 768, 50257, False, 0), (2048, 768, 50257, False, 768)])
def test_fused_forward_backward_parity(cuda_dev, cta_group, cg, M, d, V, bias, slab):
    """ftk_lmhead_ce_fused (one call, gradients included) through the C ABI against the fp32 CPU oracle: one slab and
    several, ragged slabs, bias, both CTA groupings, with and without the dW-done event (dW before dH in the last slab)."""
    from kubernetes_cloud_b200 import _lib
    cta_group(cg)
    g = torch.Generator().manual_seed(M + V)
    H = torch.randn(M, d, generator=g).to(torch.bfloat16)
    W = (torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16)
    b = (torch.randn(V, generator=g) * 0.5).to(torch.bfloat16) if bias else None
    t = torch.randint(0, V, (M,), generator=g)
    t[torch.rand(M, generator=g) < 0.1] = -100
    want = O.lmhead_ce_reference(H, W, b, t)
    Hc, Wc, tc = H.cuda(), W.cuda(), t.cuda()
    bc = None if b is None else b.cuda()
    dW = torch.full((V, d), float("nan"), device="cuda")
    db = torch.full((V,), float("nan"), device="cuda") if bias else None
    ev = torch.cuda.Event()
    ev.record()  # the handle exists once the event has been recorded
    loss, n_valid, lse, dH = _fused_ctypes(Hc, Wc, bc, tc, dW, _lib.FTK_F32, 0, slab, grad_scale=0.5, dbias=db,
                                           dw_event=int(ev.cuda_event) if cg == 2 else 0)
    if cg == 2:
        assert ev.query()  # recorded between the last slab's dW and dH GEMMs, long finished by now
    assert n_valid == want["n_valid"]
    assert abs(loss - float(want["loss"])) <= LOSS_RTOL * abs(float(want["loss"]))
    assert float((lse.cpu() - want["lse"]).abs().max()) < 2e-3
    assert O.rel_err(dH.float(), 0.5 * want["dH"]) < GRAD_RTOL
    assert O.rel_err(dW, 0.5 * want["dW"]) < GRAD_RTOL
    if bias:
        assert O.rel_err(db, 0.5 * want["dbias"]) < GRAD_RTOL
    assert float(dH[tc == -100].abs().max()) == 0.0
    # bf16 dW accumulated on top of existing content
    before = (torch.randn(V, d, generator=g) * float(want["dW"].std())).to(torch.bfloat16)
    dW16 = before.cuda().clone()
    _fused_ctypes(Hc, Wc, bc, tc, dW16, _lib.FTK_BF16, 1, slab)
    assert O.rel_err(dW16.float(), before.float() + want["dW"]) < GRAD_RTOL


def test_fused_mode_at_the_headline_shape(cuda_dev):
    """The default training path at (16384, 2048, 50257): two 8192-row slabs.  lse / loss / sampled dW rows / sampled
    dH rows against dense fp32 torch recomputations; every quantity against the recompute path of the same library."""
    from kubernetes_cloud_b200 import _lib, ops  # noqa: F401
    M, d, V = 16384, 2048, 50257
    H, W, _, t = gpu_inputs(M, d, V, seed=11, bias=False)
    assert 0 < _lib.lib().ftk_lmhead_ce_fused_slab_rows(M, d, V, 0) < M
    loss_sum, n_valid, lse, dH, dW, _ = torch.ops.ftk.lmhead_ce_fused(H, W, None, t, 1.0, False, 0)
    lse_ref = dense_lse(H, W, None)
    assert int(n_valid) == int((t >= 0).sum())
    assert float((lse - lse_ref).abs().max()) < 2e-3
    vs = sampled_vocab_rows(V, t)
    dW_ref, _ = dense_dw_rows(H, W, None, t, lse_ref, vs)
    assert O.rel_err(dW[vs].float(), dW_ref) < GRAD_RTOL
    ls2, nv2, lse2 = torch.ops.ftk.lmhead_ce_fwd(H, W, None, t)
    dH2, dW2, _ = torch.ops.ftk.lmhead_ce_bwd(H, W, None, t, lse2, None, nv2, 1.0, False)
    assert abs(float(loss_sum) - float(ls2)) <= 1e-5 * abs(float(ls2))
    assert float((lse - lse2).abs().max()) < 1e-4
    assert O.rel_err(dH.float(), dH2.float()) < 5e-3 and O.rel_err(dW.float(), dW2.float()) < 5e-3
    assert float(dH[t < 0].abs().max()) == 0.0


def test_fused_autograd_loss_scale_and_upstream_gradient(cuda_dev):
    """fused_lmhead_ce in its default mode: loss_scale is folded into the gradients made during the forward; an
    upstream gradient other than 1 is applied by the rescaling pass; a weight gradient accumulated in place with an
    upstream gradient other than 1 poisons dH instead of passing silently."""
    from kubernetes_cloud_b200 import ops
    M, d, V = 700, 128, 3001
    g = torch.Generator().manual_seed(5)
    H = torch.randn(M, d, generator=g).to(torch.bfloat16)
    W = (torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16)
    b = (torch.randn(V, generator=g) * 0.5).to(torch.bfloat16)
    t = torch.randint(0, V, (M,), generator=g)
    t[::7] = -100
    want = O.lmhead_ce_reference(H, W, b, t)
    ops.set_fused_slab_rows(256)
    try:
        Hc, Wc, bc = (x.cuda().requires_grad_(True) for x in (H, W, b))
        loss = ops.fused_lmhead_ce(Hc, Wc, bc, t.cuda(), loss_scale=0.25)
        assert abs(float(loss.detach()) - 0.25 * float(want["loss"])) <= LOSS_RTOL * 0.25 * float(want["loss"])
        (loss * 2.0).backward()  # upstream gradient 2 on top of loss_scale 0.25
        assert O.rel_err(Hc.grad.float(), 0.5 * want["dH"]) < GRAD_RTOL
        assert O.rel_err(Wc.grad.float(), 0.5 * want["dW"]) < GRAD_RTOL
        assert O.rel_err(bc.grad.float(), 0.5 * want["dbias"]) < 2e-2  # bf16 bias gradient
        with pytest.raises(Exception):
            loss.backward()
        # in-place accumulation + upstream gradient 1: exact; + upstream gradient != 1: poisoned
        Hc2 = H.cuda().requires_grad_(True)
        ops.fused_lmhead_ce(Hc2, Wc, None, t.cuda()).backward()
        assert torch.isfinite(Hc2.grad).all()
        Hc3 = H.cuda().requires_grad_(True)
        (ops.fused_lmhead_ce(Hc3, Wc, None, t.cuda()) * 3.0).backward()
        assert torch.isnan(Hc3.grad).all()
    finally:
        ops.set_fused_slab_rows(0)


def test_fused_results_are_deterministic_with_split_k(cuda_dev):
    """Shapes whose dH GEMM has few (row block x N tile) units are split over the vocabulary; the partial sums are
    added in a fixed order, so two runs give the same bits, and the result matches the oracle."""
    from kubernetes_cloud_b200 import _lib
    for M, d, V in ((2048, 256, 5003), (10496, 768, 8192)):
        g = torch.Generator().manual_seed(M)
        H = torch.randn(M, d, generator=g).to(torch.bfloat16).cuda()
        W = (torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16).cuda()
        t = torch.randint(0, V, (M,), generator=g).cuda()
        runs = []
        for _ in range(2):
            dW = torch.empty(V, d, device="cuda")
            loss, n_valid, lse, dH = _fused_ctypes(H, W, None, t, dW, _lib.FTK_F32, 0, 0)
            runs.append((loss, lse.clone(), dH.clone(), dW))
        assert runs[0][0] == runs[1][0] and torch.equal(runs[0][1], runs[1][1])
        assert torch.equal(runs[0][2], runs[1][2]) and torch.equal(runs[0][3], runs[1][3])
        want = O.lmhead_ce_reference(H, W, None, t)
        assert O.rel_err(runs[0][2].float(), want["dH"]) < GRAD_RTOL and O.rel_err(runs[0][3], want["dW"]) < GRAD_RTOL
