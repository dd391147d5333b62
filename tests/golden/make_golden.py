#!/usr/bin/env python
"""Generates the golden fixtures in this directory. Run it in the build container only:

    python tests/golden/make_golden.py [/root/reference]

* ``gather_*.npz`` come from EXECUTING THE REFERENCE'S OWN CODE: the ``TokenizedDataset`` class
  (finetuner-workflow/finetuner/finetuner.py:633-695), ``collector`` (:1030-1035) and
  ``ModifiedTrainer.compute_loss`` (:469-493) are lifted from the reference file by AST at run time and exec'd
  with stub globals (``tokenizer``, ``is_main_process``, ``logger``, ``MemoryUsage``, a stub ``Trainer`` base whose
  ``compute_loss`` returns the labels it was handed).  Nothing of the reference is copied into this repository.
* ``order_*.npz`` come from torch's ``random_split`` under ``torch.manual_seed`` exactly as FT.py:739/:761 call it.
* ``lmhead_*.npz`` come from the installed HF transformers (GPT-Neo / GPT-J tiny configs, CPU fp32): hidden states
  at the lm_head input, lm_head weight (untied so its gradient is the pure lm_head gradient), labels, loss and
  gradients.  The reference pins transformers~=4.28.1, which is absent here; the installed version is the
  stand-in (see oracle/oracle.py header).

/root/reference does not exist on the GPU box; tests only read the .npz files committed here.
"""
import ast
import logging
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
FT = os.path.join(REF_ROOT, "finetuner-workflow", "finetuner", "finetuner.py")


def lift(names):
    """Exec the named top-level ClassDef/FunctionDef nodes of the reference file in a stub namespace."""
    import mmap
    from typing import Tuple

    import numpy
    from torch.utils.data import Dataset

    src = open(FT).read()
    tree = ast.parse(src)
    wanted = [n for n in tree.body if isinstance(n, (ast.ClassDef, ast.FunctionDef)) and n.name in names]
    assert {n.name for n in wanted} == set(names), "reference layout changed"

    class StubTrainer:  # stands in for transformers.Trainer: hand back what compute_loss was given
        def compute_loss(self, model, inputs, return_outputs=False):
            out = model(**inputs)
            return (out, None) if return_outputs else out

    class StubMemoryUsage:
        @staticmethod
        def now():
            return "mem"

    ns = {
        "os": os, "mmap": mmap, "numpy": numpy, "torch": torch, "sys": sys, "Tuple": Tuple, "Dataset": Dataset,
        "Trainer": StubTrainer, "MemoryUsage": StubMemoryUsage, "is_main_process": lambda: False,
        "logger": logging.getLogger("golden"), "tokenizer": types.SimpleNamespace(pad_token_id=0, eos_token_id=0),
    }
    mod = ast.Module(body=wanted, type_ignores=[])
    exec(compile(mod, FT, "exec"), ns)
    return ns


def gather_case(ns, name, tokens, T, pad_id, eos_id, ctx_idx, tail_bytes=0):
    ns["tokenizer"] = types.SimpleNamespace(pad_token_id=pad_id, eos_token_id=eos_id)
    with tempfile.NamedTemporaryFile(suffix=".tokens", delete=False) as f:
        f.write(tokens.tobytes())
        f.write(b"\x07" * tail_bytes)  # a trailing partial context the reference must drop
        path = f.name
    try:
        ds = ns["TokenizedDataset"](path, context_length=T)
        batch = ns["collector"]([ds[int(i)] for i in ctx_idx])
        n = len(ds)
        # ModifiedTrainer.compute_loss: labels.masked_fill_(~attention_mask, -100), then super().compute_loss
        trainer = ns["ModifiedTrainer"].__new__(ns["ModifiedTrainer"])
        trainer.args = types.SimpleNamespace(gradient_accumulation_steps=1 << 30)
        inputs = {k: v.clone() for k, v in batch.items()}
        captured = {}

        def model(**kw):
            captured.update(kw)
            return torch.zeros((), requires_grad=True)

        trainer.compute_loss(model, inputs)
        np.savez_compressed(
            os.path.join(HERE, f"gather_{name}.npz"),
            tokens=tokens, tail_bytes=np.int64(tail_bytes), T=np.int64(T), pad_id=np.int64(pad_id),
            eos_id=np.int64(eos_id), ctx_idx=np.asarray(ctx_idx, dtype=np.int64), n_ctx=np.int64(n),
            input_ids=batch["input_ids"].numpy(), attention_mask=batch["attention_mask"].numpy(),
            labels_collated=batch["labels"].numpy(), labels_filled=captured["labels"].numpy())
        print(f"gather_{name}: n_ctx={n} B={len(ctx_idx)} T={T} ignored={(captured['labels'] == -100).sum().item()}")
    finally:
        os.unlink(path)


def make_gather(ns):
    rng = np.random.default_rng(1234)
    # A: pad == eos (the gpt2 tokenizer default): only the final context is masked
    T, n = 128, 7
    tok = rng.integers(0, 50257, size=n * T, dtype=np.uint16)
    tok[3 * T + 5] = 50256  # an eos inside a non-final context stays unmasked
    tok[n * T - 30:] = 50256
    gather_case(ns, "pad_is_eos", tok, T, 50256, 50256, [6, 0, 3, 5, 1, 2, 4])
    # B: pad != eos: every context gets ids != pad
    tok = rng.integers(0, 50258, size=n * T, dtype=np.uint16)
    tok[2 * T + 7] = 50257
    tok[0] = 50257
    tok[n * T - 17:] = 50257
    gather_case(ns, "pad_distinct", tok, T, 50257, 50256, [2, 6, 0, 0, 5])
    # C: BASELINE shape: T=2048, 300 trailing pads -> mask sum 1748 on the last context; token 65535 survives
    T, n = 2048, 3
    tok = rng.integers(0, 50257, size=n * T, dtype=np.uint16)
    tok[17] = 65535
    tok[n * T - 300:] = 50256
    gather_case(ns, "ctx2048", tok, T, 50256, 50256, [2, 0, 1])
    # D: odd context length (scalar path) and a trailing partial context that must be dropped
    T, n = 37, 5
    tok = rng.integers(0, 1000, size=n * T, dtype=np.uint16)
    tok[n * T - 4:] = 3
    gather_case(ns, "ragged", tok, T, 3, 9, [4, 1, 3], tail_bytes=2 * 11)


def make_order():
    from torch.utils.data import random_split
    for n, seed in ((10, 42), (4096, 42), (33, 7)):
        torch.manual_seed(seed)  # FT.py:739
        train, _val = random_split(range(n), (n, 0))  # FT.py:761 with train_ratio forced to 1
        np.savez_compressed(os.path.join(HERE, f"order_n{n}_s{seed}.npz"), n=np.int64(n), seed=np.int64(seed),
                            indices=np.asarray(train.indices, dtype=np.int64))
        print(f"order n={n} seed={seed}: {train.indices[:6]}...")


def make_lmhead():
    from transformers import GPTJConfig, GPTJForCausalLM, GPTNeoConfig, GPTNeoForCausalLM
    rng = np.random.default_rng(99)
    V, d, T, B = 503, 64, 48, 3

    def run(name, model):
        model.train()
        ids = torch.from_numpy(rng.integers(0, V, size=(B, T)).astype(np.int64))
        mask = torch.ones(B, T, dtype=torch.bool)
        mask[-1, -9:] = False
        labels = ids.clone().masked_fill_(~mask, -100)  # FT.py:476-477
        grabbed = {}

        def hook(_m, inp):
            h = inp[0]
            h.retain_grad()
            grabbed["h"] = h

        handle = model.lm_head.register_forward_pre_hook(hook)
        out = model(input_ids=ids, attention_mask=mask, labels=labels)
        out.loss.backward()
        handle.remove()
        h = grabbed["h"]
        bias = model.lm_head.bias
        np.savez_compressed(
            os.path.join(HERE, f"lmhead_{name}.npz"),
            hidden=h.detach().reshape(B * T, d).numpy(), weight=model.lm_head.weight.detach().numpy(),
            bias=(bias.detach().numpy() if bias is not None else np.zeros(0, np.float32)),
            labels=labels.numpy(), loss=np.float32(out.loss.item()),
            d_hidden=h.grad.reshape(B * T, d).numpy(), d_weight=model.lm_head.weight.grad.numpy(),
            d_bias=(bias.grad.numpy() if bias is not None else np.zeros(0, np.float32)),
            logits_row0=out.logits.detach()[0, 0].numpy())
        print(f"lmhead_{name}: loss={out.loss.item():.6f}")

    torch.manual_seed(42)
    run("gptneo_tiny", GPTNeoForCausalLM(GPTNeoConfig(
        vocab_size=V, hidden_size=d, num_layers=2, num_heads=4, max_position_embeddings=T,
        attention_types=[[["global", "local"], 1]], window_size=16, tie_word_embeddings=False)))
    torch.manual_seed(42)
    m = GPTJForCausalLM(GPTJConfig(vocab_size=V, n_embd=d, n_layer=1, n_head=4, rotary_dim=16, n_positions=T))
    with torch.no_grad():
        m.lm_head.bias.normal_(0, 0.5)  # HF zero-inits it; make the bias path observable
    run("gptj_tiny", m)


NEO125M = dict(vocab_size=50257, hidden_size=768, num_layers=12, num_heads=12, max_position_embeddings=2048,
               attention_types=[[["global", "local"], 6]], window_size=256)  # GPT-Neo-125M, literal


def weight_checksum(w: torch.Tensor) -> np.ndarray:
    """Fingerprint of a weight matrix rebuilt from its seed on another machine: sum, sum of squares, two rows."""
    w = w.detach().double()
    return np.array([float(w.sum()), float((w * w).sum()), float(w[0].sum()), float(w[-1].sum())])


def make_config0():
    """BASELINE.json configs[0]: GPT-Neo-125M (literal config, random init under torch.manual_seed(42), the
    reference's default seed FT.py:148-154), ctx 2048, batch 1, fp32 on CPU, tokens from the synthetic stream.
    The model's own loss / gradients at the lm_head boundary, i.e. what ``super().compute_loss`` (FT.py:479) and
    ``loss.backward()`` produce there.  The 154 MB weight is not stored: tests rebuild it from the same seed and check
    the fingerprint.  GPT-Neo ties lm_head to wte, so ``wte.weight.grad`` = lm_head part + embedding part; the
    embedding part is HF's own gradient at the embedding output scattered back to its rows and is subtracted."""
    from transformers import GPTNeoConfig, GPTNeoForCausalLM
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from oracle import oracle as O
    T = 2048
    tok = O.synth_tokens(4, T, 50257, 1234, trailing_pads=300)
    ids_np, _, mask_np, _ = O.gather_numpy(tok, T, [3], 50256, True)  # the padded final context: 300 ignored labels
    ids = torch.from_numpy(ids_np)
    mask = torch.from_numpy(mask_np)
    labels = ids.clone().masked_fill_(~mask, -100)  # FT.py:476-477
    torch.manual_seed(42)
    model = GPTNeoForCausalLM(GPTNeoConfig(**NEO125M))
    model.train()
    grabbed = {}

    def pre_hook(_m, inp):
        inp[0].retain_grad()
        grabbed["h"] = inp[0]

    def emb_hook(_m, _inp, out):
        out.retain_grad()
        grabbed["e"] = out

    h1 = model.lm_head.register_forward_pre_hook(pre_hook)
    h2 = model.transformer.wte.register_forward_hook(emb_hook)
    out = model(input_ids=ids, attention_mask=mask, labels=labels)
    out.loss.backward()
    h1.remove()
    h2.remove()
    W = model.lm_head.weight
    assert W.data_ptr() == model.transformer.wte.weight.data_ptr(), "GPT-Neo ties lm_head to wte"
    emb_part = torch.zeros_like(W).index_add_(0, ids.reshape(-1), grabbed["e"].grad.reshape(T, -1))
    d_weight = W.grad - emb_part  # the lm_head's own share
    h = grabbed["h"].detach().reshape(T, 768)
    dh = grabbed["h"].grad.reshape(T, 768)
    rng = np.random.default_rng(5)
    tgt = labels[0, 1:][labels[0, 1:] >= 0].numpy()
    w_rows = np.unique(np.concatenate([[0, 50256, 50176, 50200], rng.choice(tgt, 20), rng.integers(0, 50257, 40)]))
    h_rows = np.unique(np.concatenate([[0, 1, 1747, 1748, 2046, 2047], rng.integers(0, T, 250)]))
    np.savez_compressed(
        os.path.join(HERE, "config0_gptneo125m.npz"),
        hidden=h.numpy(), labels=labels.numpy(), loss=np.float32(out.loss.item()),
        weight_checksum=weight_checksum(W), seed=np.int64(42),
        h_rows=h_rows.astype(np.int64), d_hidden_rows=dh[h_rows].numpy(),
        w_rows=w_rows.astype(np.int64), d_weight_rows=d_weight[w_rows].numpy(), weight_rows=W.detach()[w_rows].numpy(),
        d_weight_norm=np.float64(d_weight.double().norm()), d_hidden_norm=np.float64(dh.double().norm()))
    print(f"config0_gptneo125m: loss={out.loss.item():.6f} |dW|={float(d_weight.norm()):.5f} rows={len(w_rows)}")


if __name__ == "__main__":
    ns = lift(["TokenizedDataset", "collector", "ModifiedTrainer"])
    make_gather(ns)
    make_order()
    make_lmhead()
    make_config0()
