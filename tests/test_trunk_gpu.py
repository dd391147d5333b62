"""Trunk linear layers (first pieces): ftk_linear_fwd / ftk_linear_bwd through the custom ops against
fp32 CPU torch (the same oracle arithmetic HF's modules run: F.linear, gelu_new), and the patched HF GPT-Neo blocks
against the unpatched model.  Tolerances as for the hot path: 1e-2 norm-wise on gradients, bf16 inputs."""
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-2


def gelu_new(x):
    return 0.5 * x * (1.0 + torch.tanh(0.7978845608028654 * (x + 0.044715 * x ** 3)))


def _mk(M, K, N, seed, bias=True):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(torch.bfloat16)
    b = (torch.randn(N, generator=g) * 0.3).to(torch.bfloat16) if bias else None
    dy = torch.randn(M, N, generator=g).to(torch.bfloat16)
    return x, w, b, dy


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("M,K,N,bias", [(1, 64, 64, True), (300, 128, 192, True), (257, 192, 320, False),
                                        (1000, 256, 1024, True), (4096, 768, 3072, True)])
def test_linear_forward_backward(cuda_dev, cta_group, cg, M, K, N, bias):
    from kubernetes_cloud_b200 import trunk
    cta_group(cg)
    x, w, b, dy = _mk(M, K, N, M + N, bias)
    xr, wr = x.float().requires_grad_(True), w.float().requires_grad_(True)
    br = b.float().requires_grad_(True) if bias else None
    yr = torch.nn.functional.linear(xr, wr, br)
    yr.backward(dy.float())
    xc, wc = x.cuda().requires_grad_(True), w.cuda().requires_grad_(True)
    bc = b.cuda().requires_grad_(True) if bias else None
    y = trunk.linear(xc, wc, bc)
    y.backward(dy.cuda())
    torch.cuda.synchronize()
    assert O.rel_err(y.float(), yr) < RTOL
    assert O.rel_err(xc.grad.float(), xr.grad) < RTOL and O.rel_err(wc.grad.float(), wr.grad) < RTOL
    if bias:
        assert O.rel_err(bc.grad.float(), br.grad) < RTOL


@pytest.mark.parametrize("M,d", [(300, 64), (2048, 256), (4096, 768)])
def test_mlp_gelu_new(cuda_dev, M, d):
    """c_proj(gelu_new(c_fc(x))): bias + GELU in the first GEMM's epilogue, GELU' in the second one's dX epilogue."""
    from kubernetes_cloud_b200 import trunk
    g = torch.Generator().manual_seed(M)
    x = torch.randn(M, d, generator=g).to(torch.bfloat16)
    w1 = (torch.randn(4 * d, d, generator=g) / d ** 0.5).to(torch.bfloat16)
    b1 = (torch.randn(4 * d, generator=g) * 0.3).to(torch.bfloat16)
    w2 = (torch.randn(d, 4 * d, generator=g) / (4 * d) ** 0.5).to(torch.bfloat16)
    b2 = (torch.randn(d, generator=g) * 0.3).to(torch.bfloat16)
    dy = torch.randn(M, d, generator=g).to(torch.bfloat16)
    ref = [t.float().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    yr = torch.nn.functional.linear(gelu_new(torch.nn.functional.linear(ref[0], ref[1], ref[2])), ref[3], ref[4])
    yr.backward(dy.float())
    got = [t.cuda().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    y = trunk.mlp_gelu_new(*got)
    y.backward(dy.cuda())
    torch.cuda.synchronize()
    assert O.rel_err(y.float(), yr) < RTOL
    for name, a, r in zip(("dx", "dw_fc", "db_fc", "dw_proj", "db_proj"), got, ref):
        assert O.rel_err(a.grad.float(), r.grad) < 1.5e-2, name  # the hidden activation is rounded to bf16 once more


@pytest.mark.parametrize("M,K,N", [(300, 128, 192), (4096, 768, 768)])
def test_linear_residual_epilogue(cuda_dev, M, K, N):
    """y = x W^T + b + residual with the add inside the GEMM's epilogue (HF GPTNeoBlock's two residual connections):
    value and all four gradients against fp32 torch; the residual's gradient is dy itself."""
    from kubernetes_cloud_b200 import trunk
    x, w, b, dy = _mk(M, K, N, M + 3)
    res = torch.randn(M, N, generator=torch.Generator().manual_seed(5)).to(torch.bfloat16)
    ref = [t.float().requires_grad_(True) for t in (x, w, b, res)]
    yr = torch.nn.functional.linear(ref[0], ref[1], ref[2]) + ref[3]
    yr.backward(dy.float())
    got = [t.cuda().requires_grad_(True) for t in (x, w, b, res)]
    y = trunk.linear(*got)
    y.backward(dy.cuda())
    torch.cuda.synchronize()
    assert O.rel_err(y.float(), yr) < RTOL
    for name, a, r in zip(("dx", "dw", "db", "dres"), got, ref):
        assert O.rel_err(a.grad.float(), r.grad) < RTOL, name
    assert torch.equal(got[3].grad, dy.cuda())


def test_mlp_residual_and_qkv_chain(cuda_dev):
    """mlp_gelu_new(..., residual) and the three projections of one input whose dX GEMMs chain (dx_add)."""
    from kubernetes_cloud_b200 import trunk
    M, d = 640, 256
    g = torch.Generator().manual_seed(77)
    x = torch.randn(M, d, generator=g).to(torch.bfloat16)
    ws = [(torch.randn(d, d, generator=g) / d ** 0.5).to(torch.bfloat16) for _ in range(3)]
    bs = [None, (torch.randn(d, generator=g) * 0.3).to(torch.bfloat16), None]
    dys = [torch.randn(M, d, generator=g).to(torch.bfloat16) for _ in range(3)]
    xr = x.float().requires_grad_(True)
    wr = [w.float().requires_grad_(True) for w in ws]
    br = [None if b is None else b.float().requires_grad_(True) for b in bs]
    sum(((torch.nn.functional.linear(xr, w, b) * dy.float()).sum() for w, b, dy in zip(wr, br, dys))).backward()
    xc = x.cuda().requires_grad_(True)
    wc = [w.cuda().requires_grad_(True) for w in ws]
    bc = [None if b is None else b.cuda().requires_grad_(True) for b in bs]
    q, k, v = trunk.qkv(xc, wc[0], bc[0], wc[1], bc[1], wc[2], bc[2])
    torch.autograd.backward([q, k, v], [dy.cuda() for dy in dys])
    torch.cuda.synchronize()
    assert O.rel_err(xc.grad.float(), xr.grad) < RTOL
    for a, r in zip(wc, wr):
        assert O.rel_err(a.grad.float(), r.grad) < RTOL
    assert O.rel_err(bc[1].grad.float(), br[1].grad) < RTOL
    # MLP with the residual in c_proj's epilogue
    w1 = (torch.randn(4 * d, d, generator=g) / d ** 0.5).to(torch.bfloat16)
    w2 = (torch.randn(d, 4 * d, generator=g) / (4 * d) ** 0.5).to(torch.bfloat16)
    b1 = (torch.randn(4 * d, generator=g) * 0.3).to(torch.bfloat16)
    b2 = (torch.randn(d, generator=g) * 0.3).to(torch.bfloat16)
    res = torch.randn(M, d, generator=g).to(torch.bfloat16)
    ref = [t.float().requires_grad_(True) for t in (x, w1, b1, w2, b2, res)]
    yr = torch.nn.functional.linear(gelu_new(torch.nn.functional.linear(ref[0], ref[1], ref[2])), ref[3], ref[4]) + ref[5]
    yr.backward(dys[0].float())
    got = [t.cuda().requires_grad_(True) for t in (x, w1, b1, w2, b2, res)]
    y = trunk.mlp_gelu_new(*got)
    y.backward(dys[0].cuda())
    torch.cuda.synchronize()
    assert O.rel_err(y.float(), yr) < RTOL
    for name, a, r in zip(("dx", "dw_fc", "db_fc", "dw_proj", "db_proj", "dres"), got, ref):
        assert O.rel_err(a.grad.float(), r.grad) < 1.5e-2, name


@pytest.mark.parametrize("M,D,bias", [(1, 256, True), (37, 768, True), (1000, 1024, False), (4099, 2048, True),
                                      (515, 4096, True), (130, 8192, True)])
def test_layer_norm_forward_backward(cuda_dev, M, D, bias):
    """ftk_layernorm_fwd / _bwd against torch.nn.functional.layer_norm in fp32 on the CPU: y, dx, dgamma, dbeta, for
    rows that do not fill the last block and every register-chunk variant of the forward."""
    from kubernetes_cloud_b200 import trunk
    g = torch.Generator().manual_seed(M + D)
    x = (torch.randn(M, D, generator=g) * 2.0 + 0.5).to(torch.bfloat16)
    w = (1.0 + 0.2 * torch.randn(D, generator=g)).to(torch.bfloat16)
    b = (0.3 * torch.randn(D, generator=g)).to(torch.bfloat16) if bias else None
    dy = torch.randn(M, D, generator=g).to(torch.bfloat16)
    ref = [t.float().requires_grad_(True) for t in (x, w)] + ([b.float().requires_grad_(True)] if bias else [None])
    yr = torch.nn.functional.layer_norm(ref[0], (D,), ref[1], ref[2], 1e-5)
    yr.backward(dy.float())
    got = [t.cuda().requires_grad_(True) for t in (x, w)] + ([b.cuda().requires_grad_(True)] if bias else [None])
    y = trunk.layer_norm(*got, 1e-5)
    y.backward(dy.cuda())
    torch.cuda.synchronize()
    assert O.rel_err(y.float(), yr) < 5e-3          # one bf16 rounding of an O(1) value
    assert O.rel_err(got[0].grad.float(), ref[0].grad) < RTOL
    assert O.rel_err(got[1].grad.float(), ref[1].grad) < RTOL
    if bias:
        assert O.rel_err(got[2].grad.float(), ref[2].grad) < RTOL


def test_layer_norm_fork_adds_the_residual_gradient(cuda_dev):
    """(x, LN(x)): the backward kernel returns d(residual branch) + d(LayerNorm input) in one pass, deterministically."""
    from kubernetes_cloud_b200 import trunk
    M, D = 777, 768
    g = torch.Generator().manual_seed(3)
    x = torch.randn(M, D, generator=g).to(torch.bfloat16)
    w = (1.0 + 0.2 * torch.randn(D, generator=g)).to(torch.bfloat16)
    b = (0.3 * torch.randn(D, generator=g)).to(torch.bfloat16)
    d_res, d_y = (torch.randn(M, D, generator=g).to(torch.bfloat16) for _ in range(2))
    xr, wr, br = (t.float().requires_grad_(True) for t in (x, w, b))
    yr = torch.nn.functional.layer_norm(xr, (D,), wr, br, 1e-5)
    torch.autograd.backward([xr * 1.0, yr], [d_res.float(), d_y.float()])
    grads = []
    for _ in range(2):
        xc, wc, bc = (t.cuda().requires_grad_(True) for t in (x, w, b))
        res, y = trunk.layer_norm_fork(xc, wc, bc, 1e-5)
        assert res.data_ptr() == xc.data_ptr()  # the residual branch is the input itself, not a copy
        torch.autograd.backward([res, y], [d_res.cuda(), d_y.cuda()])
        grads.append((xc.grad, wc.grad, bc.grad))
    torch.cuda.synchronize()
    assert O.rel_err(grads[0][0].float(), xr.grad) < RTOL
    assert O.rel_err(grads[0][1].float(), wr.grad) < RTOL and O.rel_err(grads[0][2].float(), br.grad) < RTOL
    for a, c in zip(grads[0], grads[1]):
        assert torch.equal(a, c)


def test_linear_accumulates_into_main_grad(cuda_dev):
    from kubernetes_cloud_b200 import trunk
    x, w, b, dy = _mk(512, 128, 256, 9)
    wr = w.float().requires_grad_(True)
    for _ in range(2):
        torch.nn.functional.linear(x.float(), wr, b.float()).backward(dy.float())
    wc = w.cuda().requires_grad_(True)
    wc.main_grad = torch.full((256, 128), 5.0, device="cuda")
    wc.main_grad._ftk_overwrite = True  # first call overwrites (lazy zero), second adds
    for _ in range(2):
        trunk.linear(x.cuda().requires_grad_(True), wc, b.cuda()).backward(dy.cuda())
    assert wc.grad is None and O.rel_err(wc.main_grad, wr.grad) < 5e-3


def attention_reference(q, k, v, key_mask, H, window, scale):
    """HF GPT-Neo's eager ``_attn`` restated in fp32 torch on [B, T, H * dh] operands (modeling_gpt_neo.py: fp32 Q.K^T,
    where(causal / sliding-window bias, s, finfo.min), + additive padding mask, softmax, .V)."""
    B, T, D = q.shape
    dh = D // H
    qh, kh, vh = (t.float().view(B, T, H, dh).permute(0, 2, 1, 3) for t in (q, k, v))
    s = torch.matmul(qh, kh.transpose(-1, -2)) * scale
    bias = torch.tril(torch.ones(T, T, dtype=torch.bool))
    if window > 0:
        bias = torch.bitwise_xor(bias, torch.tril(bias, -window))
    s = torch.where(bias, s, torch.tensor(torch.finfo(torch.float32).min))
    if key_mask is not None:
        s = s + (~key_mask)[:, None, None, :].float() * torch.finfo(torch.float32).min
    o = torch.matmul(torch.softmax(s, -1), vh)
    return o.permute(0, 2, 1, 3).reshape(B, T, D)


@pytest.mark.parametrize("B,T,H,dh,window,scale,masked", [(1, 128, 1, 64, 0, 1.0, False), (2, 256, 2, 64, 0, 1.0, False),
                                                           (2, 384, 2, 128, 0, 0.125, True), (1, 512, 3, 64, 200, 1.0, False),
                                                           (2, 512, 2, 128, 256, 1.0, True), (1, 256, 2, 128, 64, 1.0, False)])
def test_attention_forward_backward(cuda_dev, B, T, H, dh, window, scale, masked):
    """Causal and sliding-window attention against the fp32 restatement of HF's eager path: output and dQ / dK / dV.
    The padded positions (trailing keys of the last sequence) are excluded from the comparison, as the loss ignores
    them."""
    from kubernetes_cloud_b200 import trunk
    g = torch.Generator().manual_seed(T + dh + window)
    q, k, v, d_o = ((torch.randn(B, T, H * dh, generator=g) * s).to(torch.bfloat16) for s in (0.5, 0.5, 1.0, 1.0))
    key_mask = None
    if masked:
        key_mask = torch.ones(B, T, dtype=torch.bool)
        key_mask[-1, T - 37:] = False
    ref = [t.float().requires_grad_(True) for t in (q, k, v)]
    o_ref = attention_reference(*ref, key_mask, H, window, scale)
    keep = torch.ones(B, T, 1) if key_mask is None else key_mask[:, :, None].float()  # padded queries get no gradient
    o_ref.backward(d_o.float() * keep)
    got = [t.cuda().requires_grad_(True) for t in (q, k, v)]
    o = trunk.attention(*got, None if key_mask is None else key_mask.cuda(), H, window, scale)
    o.backward((d_o.float() * keep).to(torch.bfloat16).cuda())
    torch.cuda.synchronize()
    assert torch.isfinite(o).all()
    assert O.rel_err(o.float().cpu() * keep, o_ref * keep) < RTOL
    for name, a, r in zip(("dq", "dk", "dv"), got, ref):
        assert torch.isfinite(a.grad).all(), name
        assert O.rel_err(a.grad.float().cpu() * keep, r.grad * keep) < 1.5e-2, name


def test_attention_rows_without_keys_and_padding_inside_a_block(cuda_dev):
    """Left padding: the first queries of a sequence see no key at all (output 0, lse +inf, no NaN anywhere), and a
    padding boundary that falls inside a 128-key block takes the per-element test while the other blocks of the same
    sequence take the fast path."""
    from kubernetes_cloud_b200 import trunk
    B, T, H, dh = 2, 512, 2, 128
    g = torch.Generator().manual_seed(11)
    q, k, v, d_o = ((torch.randn(B, T, H * dh, generator=g) * s).to(torch.bfloat16) for s in (0.5, 0.5, 1.0, 1.0))
    key_mask = torch.ones(B, T, dtype=torch.bool)
    key_mask[0, :150] = False      # left padding over one whole block and part of the next
    key_mask[1, 300:310] = False   # a hole inside the third block
    keep = key_mask[:, :, None].float()
    ref = [t.float().requires_grad_(True) for t in (q, k, v)]
    o_ref = attention_reference(*ref, key_mask, H, 0, 1.0)
    o_ref.backward(d_o.float() * keep)
    got = [t.cuda().requires_grad_(True) for t in (q, k, v)]
    o = trunk.attention(*got, key_mask.cuda(), H, 0, 1.0)
    o.backward((d_o.float() * keep).to(torch.bfloat16).cuda())
    torch.cuda.synchronize()
    assert torch.isfinite(o).all() and float(o[0, :150].abs().max()) == 0.0
    assert O.rel_err(o.float().cpu() * keep, o_ref * keep) < RTOL
    for name, a, r in zip(("dq", "dk", "dv"), got, ref):
        assert torch.isfinite(a.grad).all(), name
        assert O.rel_err(a.grad.float().cpu() * keep, r.grad * keep) < 1.5e-2, name


def test_attention_long_sequence_is_deterministic(cuda_dev):
    """T = 2048 (the bench context): many tiles through the double-buffered score pipelines and the three-stage operand
    rings of the backward kernels; two runs agree bit for bit, and output and gradients match the fp32 restatement."""
    from kubernetes_cloud_b200 import trunk
    B, T, H, dh = 1, 2048, 2, 128
    g = torch.Generator().manual_seed(21)
    q, k, v, d_o = ((torch.randn(B, T, H * dh, generator=g) * s).to(torch.bfloat16) for s in (0.3, 0.3, 1.0, 1.0))
    for window in (0, 256):
        ref = [t.float().requires_grad_(True) for t in (q, k, v)]
        o_ref = attention_reference(*ref, None, H, window, 1.0)
        o_ref.backward(d_o.float())
        runs = []
        for _ in range(2):
            got = [t.cuda().requires_grad_(True) for t in (q, k, v)]
            o = trunk.attention(*got, None, H, window, 1.0)
            o.backward(d_o.cuda())
            runs.append((o.detach(), *(t.grad for t in got)))
        torch.cuda.synchronize()
        for a, c in zip(*runs):
            assert torch.equal(a, c)
        assert O.rel_err(runs[0][0].float().cpu(), o_ref) < RTOL
        for name, a, r in zip(("dq", "dk", "dv"), runs[0][1:], ref):
            assert O.rel_err(a.float().cpu(), r.grad) < 1.5e-2, (name, window)


@pytest.mark.parametrize("fuse_blocks", [False, True])
def test_patched_gpt_neo_matches_the_hf_blocks(cuda_dev, fuse_blocks):
    """patch_gpt_neo on a small GPT-Neo (global + local attention layers): same loss and gradients as the unpatched
    bf16 model on the GPU, and close to the fp32 CPU model; module by module, and with whole blocks fused (LayerNorm
    forks, chained projections, residual adds in the GEMM epilogues)."""
    from transformers import GPTNeoConfig, GPTNeoForCausalLM
    from kubernetes_cloud_b200 import trunk
    from kubernetes_cloud_b200.loss import compute_loss
    torch.manual_seed(1)
    cfg = GPTNeoConfig(vocab_size=1024, hidden_size=256, num_layers=2, num_heads=4, max_position_embeddings=128,
                       attention_types=[[["global", "local"], 1]], window_size=32)
    ref32 = GPTNeoForCausalLM(cfg)
    ids = torch.randint(0, 1024, (4, 128))
    mask = torch.ones_like(ids, dtype=torch.bool)
    out32 = ref32(input_ids=ids, attention_mask=mask, labels=ids)
    out32.loss.backward()
    models = []
    for patched in (False, True):
        m = GPTNeoForCausalLM(cfg)
        m.load_state_dict(ref32.state_dict())
        m = m.cuda().to(torch.bfloat16)
        m.train()
        if patched:
            # per block: MLP, four projections, the attention itself (+ two LayerNorms and the block); + ln_f
            assert trunk.patch_gpt_neo(m, fuse_blocks=fuse_blocks) == (2 * 9 + 1 if fuse_blocks else 2 * 6)
        inputs = {"input_ids": ids.cuda(), "attention_mask": mask.cuda(), "labels": ids.cuda().clone()}
        loss = compute_loss(m, inputs)
        loss.backward()
        models.append((float(loss.detach()), m))
    (l0, m0), (l1, m1) = models
    assert abs(l1 - l0) <= 3e-3 * abs(l0) and abs(l1 - float(out32.loss)) <= 1e-2 * float(out32.loss)
    for (n0, p0), (n1, p1), (_, p32) in zip(m0.named_parameters(), m1.named_parameters(), ref32.named_parameters()):
        if p0.grad is None:
            continue
        assert O.rel_err(p1.grad.float(), p0.grad.float()) < 5e-2, n0   # bf16 model vs bf16 model
        assert O.rel_err(p1.grad.float(), p32.grad) < 8e-2, n0          # vs the fp32 CPU model (bf16 trunk noise)
