"""G1 window gather through the C ABI vs the reference-pinned oracle: bit-exact (integer / byte work)."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def run_gpu(tokens, T, idx, pad, pad_is_eos, shift):
    from kubernetes_cloud_b200 import ops  # noqa: F401
    t = torch.from_numpy(np.ascontiguousarray(tokens).view(np.int16)).cuda()
    i = torch.from_numpy(np.asarray(idx, dtype=np.int64)).cuda()
    out = torch.ops.ftk.gather_windows(t, i, T, pad, pad_is_eos, shift)
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in out]


@pytest.mark.parametrize("case", ["pad_is_eos", "pad_distinct", "ctx2048", "ragged"])
def test_matches_reference_fixture(cuda_dev, golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"gather_{case}.npz"))
    T, pad, eos = int(g["T"]), int(g["pad_id"]), int(g["eos_id"])
    tokens = g["tokens"][: int(g["n_ctx"]) * T]
    ids, labels, mask, n_valid = run_gpu(tokens, T, g["ctx_idx"], pad, pad == eos, False)
    assert ids.dtype == np.int64 and labels.dtype == np.int64 and mask.dtype == np.bool_
    np.testing.assert_array_equal(ids, g["input_ids"])
    np.testing.assert_array_equal(mask, g["attention_mask"])
    np.testing.assert_array_equal(labels, g["labels_filled"])
    np.testing.assert_array_equal(n_valid, (g["labels_filled"] != -100).sum(1))


@pytest.mark.parametrize("T,n,B", [(16, 3, 5), (128, 7, 7), (1024, 40, 16), (2048, 20, 8), (37, 5, 3), (4096, 5, 4),
                                   (2064, 4, 4), (8, 1, 1)])
@pytest.mark.parametrize("pad_is_eos", [True, False])
@pytest.mark.parametrize("shift", [False, True])
def test_matches_c_oracle(cuda_dev, T, n, B, pad_is_eos, shift):
    tok = O.synth_tokens(n, T, 50257, 1234, trailing_pads=min(300, T // 2))
    tok[::97] = 50256  # pads sprinkled inside contexts too
    rng = np.random.default_rng(T + B)
    idx = rng.integers(0, n, size=B)
    idx[0] = n - 1
    want = O.gather_c(tok, T, idx, 50256, pad_is_eos, shift)
    got = run_gpu(tok, T, idx, 50256, pad_is_eos, shift)
    for w, g in zip(want, got):
        np.testing.assert_array_equal(w, g)


def test_empty_batch_and_dataset_class(cuda_dev, tmp_path):
    from kubernetes_cloud_b200 import TokenizedDataset, collector, DeviceCollator
    T, n = 64, 5
    tok = O.synth_tokens(n, T, 1000, 7, pad_id=3, trailing_pads=9)
    path = tmp_path / "x.tokens"
    with open(path, "wb") as f:
        f.write(tok.tobytes())
        f.write(b"\x01" * 10)  # trailing partial context is dropped (FT.py:647)
    ds = TokenizedDataset(str(path), context_length=T, pad_token_id=3, eos_token_id=3)
    assert len(ds) == n
    ids, mask = ds[n - 1]
    assert ids.dtype == torch.int64 and ids.shape == (T,) and mask.dtype == torch.bool
    want = O.gather_numpy(tok, T, [n - 1], 3, True)
    np.testing.assert_array_equal(ids.cpu().numpy(), want[0][0])
    np.testing.assert_array_equal(mask.cpu().numpy(), want[2][0])
    batch = collector([ds[i] for i in (0, 4, 2)])
    assert batch["labels"].data_ptr() != batch["input_ids"].data_ptr()  # labels do not alias input_ids
    np.testing.assert_array_equal(batch["input_ids"].cpu().numpy(), O.gather_numpy(tok, T, [0, 4, 2], 3, True)[0])
    b2 = DeviceCollator(ds)([0, 4, 2])
    np.testing.assert_array_equal(b2["labels"].cpu().numpy(), O.gather_numpy(tok, T, [0, 4, 2], 3, True)[1])
    assert ds.batch([])["input_ids"].shape == (0, T)
    with pytest.raises(IndexError):
        ds.batch([n])
    with pytest.raises(OSError):
        TokenizedDataset(str(tmp_path / "missing.tokens"), T, pad_token_id=3, eos_token_id=3)


def test_bulk_properties(cuda_dev):
    """Full-size property checks where the CPU oracle would be slow: ids round-trip to the token stream, the mask
    is ids != pad, labels are ids where the mask holds."""
    T, n = 2048, 4096  # 2^23 tokens
    g = torch.Generator(device="cuda").manual_seed(1)
    tok = torch.randint(0, 50257, (n * T,), device="cuda", generator=g, dtype=torch.int32).to(torch.int16)
    idx = torch.randperm(n, device="cuda", generator=g)
    ids, labels, mask, n_valid = torch.ops.ftk.gather_windows(tok, idx, T, 50256, False, False)
    wide = (tok.to(torch.int32) & 0xFFFF).to(torch.int64).view(n, T)
    assert torch.equal(ids, wide[idx])
    assert torch.equal(mask, ids != 50256)
    assert torch.equal(labels, torch.where(mask, ids, torch.full_like(ids, -100)))
    assert torch.equal(n_valid.to(torch.int64), mask.sum(1))


def test_device_batcher_epoch_matches_reference_order(cuda_dev, tmp_path):
    """The pipeline yields exactly the batches the reference would: random_split order, rank slice, resume."""
    from kubernetes_cloud_b200 import DeviceBatcher, TokenizedDataset, write_token_file
    T, n = 64, 21
    toks = np.random.default_rng(3).integers(0, 500, size=n * T - 5)
    path = str(tmp_path / "t.tokens")
    write_token_file(path, toks, T, pad_id=500)
    raw = np.fromfile(path, dtype=np.uint16)
    ds = TokenizedDataset(path, T, pad_token_id=500, eos_token_id=500)
    for rank, world in ((0, 1), (1, 2)):
        order = O.rank_slice(O.train_order(n, 42, True), rank, world)
        bt = DeviceBatcher(ds, batch_size=4, seed=42, shuffle=True, rank=rank, world=world, shift_labels=True, prefetch=3)
        got = list(bt.epoch())
        assert len(got) == bt.batches_per_epoch() == (len(order) + 3) // 4
        for k, b in enumerate(got):
            idx = order[4 * k:4 * k + 4]
            np.testing.assert_array_equal(bt.index_list(k).numpy(), idx)
            want = O.gather_c(raw, T, idx, 500, True, True)
            np.testing.assert_array_equal(b["input_ids"].cpu().numpy(), want[0])
            np.testing.assert_array_equal(b["labels"].cpu().numpy(), want[1])
            np.testing.assert_array_equal(b["attention_mask"].cpu().numpy(), want[2])
        resumed = list(bt.epoch(start_batch=2))
        assert len(resumed) == len(got) - 2
        assert torch.equal(resumed[0]["input_ids"], got[2]["input_ids"])
    ns = DeviceBatcher(ds, batch_size=8, shuffle=False)  # --no-shuffle keeps file order
    np.testing.assert_array_equal(ns.index_list(0).numpy(), np.arange(8))


def test_random_shapes_against_c_oracle(cuda_dev):
    """Randomised sweep over context lengths (all three kernel paths: T % 16 == 0, T % 4 == 0, scalar), batch sizes,
    pad rules and both label forms; bit-exact against the C restatement."""
    rng = np.random.default_rng(2024)
    for case in range(60):
        T = int(rng.choice([4, 8, 12, 16, 20, 33, 48, 64, 100, 127, 512, 520, 2048, 2052, 4100]))
        n = int(rng.integers(1, 12))
        B = int(rng.integers(1, 10))
        pad = int(rng.integers(0, 4))
        tok = rng.integers(0, 5, size=n * T, dtype=np.uint16)
        idx = rng.integers(0, n, size=B)
        pad_is_eos, shift = bool(rng.integers(0, 2)), bool(rng.integers(0, 2))
        want = O.gather_c(tok, T, idx, pad, pad_is_eos, shift)
        got = run_gpu(tok, T, idx, pad, pad_is_eos, shift)
        for w, g in zip(want, got):
            np.testing.assert_array_equal(w, g, err_msg=f"case {case}: T={T} n={n} B={B} pad={pad} eos={pad_is_eos} shift={shift}")
