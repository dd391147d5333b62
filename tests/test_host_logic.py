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


def _worker_main_grads(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kubernetes_cloud_b200.dp import GradSync
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 8)).to(torch.bfloat16)
    sync = GradSync(model.parameters(), rank, world, bucket_bytes=2048, backend="gloo")
    assert sync.main_grads and all(b.flat.dtype == torch.float32 for b in sync.buckets)
    g = torch.Generator().manual_seed(200 + rank)
    xs = [torch.randn(4, 16, generator=g).to(torch.bfloat16) for _ in range(3)]
    sync.zero_grad()
    with sync.no_sync():
        for x in xs[:2]:
            model(x).float().pow(2).mean().backward()
    model(xs[2]).float().pow(2).mean().backward()
    sync.finish()
    assert all(p.grad is None for p in model.parameters())  # micro-batch gradients were folded into main_grad
    got = torch.cat([p.main_grad.reshape(-1) for p in model.parameters()])
    # expected: per rank, the fp32 sum of the three bf16 micro-batch gradients; then the mean over ranks in fp32
    total = None
    for r in range(world):
        torch.manual_seed(0)
        ref = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.Tanh(), torch.nn.Linear(32, 8)).to(torch.bfloat16)
        gr = torch.Generator().manual_seed(200 + r)
        acc = None
        for _ in range(3):
            ref.zero_grad()
            ref(torch.randn(4, 16, generator=gr).to(torch.bfloat16)).float().pow(2).mean().backward()
            v = torch.cat([p.grad.float().reshape(-1) for p in ref.parameters()])
            acc = v if acc is None else acc + v
        total = acc if total is None else total + acc
    out[rank] = float((got - total / world).abs().max())
    dist.destroy_process_group()


def test_grad_sync_fp32_main_grads_gloo_world2():
    """bf16 parameters: micro-batch gradients accumulate into fp32 main gradients and the reduction runs in fp32
    (ds_config.json:45), so nothing is rounded to bf16 between the backward that made it and the optimizer."""
    world = 2
    port = _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker_main_grads, args=(world, port, out), nprocs=world, join=True)
        assert len(out) == world
        for r in range(world):
            assert out[r] < 1e-6


def test_grad_sync_rejects_cpu_params_for_ftk_backend():
    from kubernetes_cloud_b200 import FtkError
    from kubernetes_cloud_b200.dp import GradSync
    with pytest.raises(FtkError):
        GradSync(torch.nn.Linear(4, 4).parameters(), 0, 1, backend="ftk")


def test_warmup_lr_matches_deepspeed_formula():
    from kubernetes_cloud_b200.optim import WarmupLR
    for typ in ("log", "linear"):
        s = WarmupLR(max_lr=5e-5, warmup_steps=10, min_lr=1e-6, warmup_type=typ)
        for it in range(0, 25):
            assert abs(s.lr_at(it) - O.warmup_lr_reference(it, 5e-5, 10, 1e-6, typ)) < 1e-18
        assert s.lr_at(0) == pytest.approx(1e-6) and s.lr_at(10) == pytest.approx(5e-5) and s.lr_at(1000) == pytest.approx(5e-5)
    with pytest.raises(ValueError):
        WarmupLR(1e-3, 10, warmup_type="cosine")


def test_token_file_writer_and_validator(tmp_path):
    from kubernetes_cloud_b200 import validate_token_file, write_token_file
    T = 32
    rng = np.random.default_rng(0)
    toks = rng.integers(0, 1000, size=5 * T - 7)
    p = str(tmp_path / "a.tokens")
    assert write_token_file(p, toks, T, pad_id=1000) == 5
    raw = np.fromfile(p, dtype=np.uint16)
    assert raw.size == 5 * T and (raw[-7:] == 1000).all() and (raw[:toks.size] == toks).all()
    rep = validate_token_file(p, T, vocab_size=1001, pad_id=1000, eos_id=1000)
    assert rep.ok and rep.n_contexts == 5 and rep.trailing_bytes == 0 and rep.max_token == 1000
    assert rep.padded_contexts == [4] and rep.convention_ok is True and rep.vocab_ok is True
    # the oracle's reader agrees with what was written (same format the reference reads)
    ids, labels, mask, _ = O.gather_numpy(raw, T, [4], 1000, True)
    assert mask[0].sum() == T - 7 and (labels[0, -7:] == -100).all()
    # violations: vocabulary, trailing partial context, padding in a non-final context when pad == eos
    assert not validate_token_file(p, T, vocab_size=1000).ok
    with open(p, "ab") as f:
        f.write(b"\x00" * 6)
    rep = validate_token_file(p, T)
    assert rep.trailing_bytes == 6 and rep.n_contexts == 5 and not rep.ok
    raw[2 * T - 3:2 * T] = 1000
    raw.tofile(p)
    rep = validate_token_file(p, T, pad_id=1000, eos_id=1000)
    assert rep.convention_ok is False and rep.padded_contexts == [1, 4]
    assert validate_token_file(p, T, pad_id=1000, eos_id=999).ok  # distinct pad: every context is masked by id
    with pytest.raises(ValueError):
        write_token_file(p, [70000], T, 0)


def test_sampler_epoch_order_matches_distributed_sampler_over_the_split():
    """world > 1: HF Trainer wraps the reference's Subset in DistributedSampler(seed=args.seed); reproduce its walk."""
    from torch.utils.data import Subset, random_split
    from torch.utils.data.distributed import DistributedSampler
    from kubernetes_cloud_b200.ordering import sampler_epoch_order, train_order
    n, seed = 37, 42
    torch.manual_seed(seed)
    train, _ = random_split(range(n), (n, 0))          # FT.py:739, :761
    split = train_order(n, seed)
    assert list(split) == list(train.indices)
    for world in (2, 8):
        for epoch in (0, 3):
            for rank in range(world):
                smp = DistributedSampler(train, num_replicas=world, rank=rank, seed=seed, shuffle=True)
                smp.set_epoch(epoch)
                want = [train[i] for i in smp]           # dataset items = context indices
                got = sampler_epoch_order(split, seed, epoch, rank, world)
                assert [int(v) for v in got] == want


# ------------------------------------------------------------------------------------------------ CLI boundary
# The flag string the workflow hands to finetuner.py (finetune-workflow.yaml:267-300), templates as written there,
# and the workflow's own default parameter values (finetune-workflow.yaml:8-199).
WORKFLOW_FLAGS = """
--run-name={{workflow.parameters.run_name}}
--model=/{{workflow.parameters.pvc}}/models/{{workflow.parameters.model}}
--trust-remote-code={{workflow.parameters.trust_remote_code}}
--dataset {{DATASET}}
--tensorizer-uri='{{workflow.parameters.tensorizer_uri}}'
--lr={{workflow.parameters.learn_rate}}
--epochs={{workflow.parameters.epochs}}
--train-ratio=1.0
--eot='{{workflow.parameters.eot_token}}'
--pad='{{workflow.parameters.pad_token}}'
--warmup-ratio={{workflow.parameters.warmup_ratio}}
--bs={{workflow.parameters.batch_size}}
--bs-divisor={{workflow.parameters.batch_size_divisor}}
--gradients={{workflow.parameters.gradients}}
--zero-stage={{workflow.parameters.zero_stage}}
--seed={{workflow.parameters.random_seed}}
--output-path=/{{workflow.parameters.pvc}}/finetunes/
--no-resume={{workflow.parameters.no_resume}}
--cache=/{{workflow.parameters.pvc}}/cache/
--save-steps={{workflow.parameters.save_steps}}
--context-size={{workflow.parameters.context}}
--project-id={{workflow.parameters.project_id}}
--logs=/{{workflow.parameters.pvc}}/{{workflow.parameters.logs}}
--fp16={{workflow.parameters.force_fp16}}
--no-shuffle={{workflow.parameters.no_shuffle}}
--prompt-file={{PROMPT_FILE}}
--prompt-every={{workflow.parameters.prompt_every}}
--prompt-tokens={{workflow.parameters.prompt_tokens}}
--prompt-samples={{workflow.parameters.prompt_samples}}
--top-k={{workflow.parameters.top_k}}
--top-p={{workflow.parameters.top_p}}
--temperature={{workflow.parameters.temperature}}
--repetition-penalty={{workflow.parameters.repetition_penalty}}
"""
WORKFLOW_DEFAULTS = {"run_name": "my-run", "pvc": "finetune-data", "model": "EleutherAI/pythia-2.8b-deduped",
                     "trust_remote_code": "false", "tensorizer_uri": "", "no_shuffle": "false", "eot_token": "",
                     "pad_token": "", "context": "2048", "prompt_every": "0", "prompt_tokens": "200",
                     "prompt_samples": "5", "top_k": "50", "top_p": "0.95", "temperature": "1.0",
                     "repetition_penalty": "1.1", "warmup_ratio": "0.1", "batch_size": "-1", "force_fp16": "false",
                     "batch_size_divisor": "1.0", "random_seed": "42", "learn_rate": "5e-5", "epochs": "1",
                     "gradients": "5", "zero_stage": "3", "save_steps": "500", "no_resume": "false", "logs": "logs",
                     "project_id": "huggingface"}


def _workflow_argv(dataset_path, **overrides):
    import re
    import shlex
    params = dict(WORKFLOW_DEFAULTS, **overrides)
    text = WORKFLOW_FLAGS.replace("{{DATASET}}", dataset_path).replace("{{PROMPT_FILE}}", "")
    text = re.sub(r"\{\{workflow\.parameters\.(\w+)\}\}", lambda m: params[m.group(1)], text)
    return shlex.split(text) + ["--local_rank=0"]  # the DeepSpeed launcher appends this one (FT.py:261-266)


def test_workflow_flag_string_parses_verbatim(tmp_path):
    from decimal import Decimal
    from kubernetes_cloud_b200.finetune import build_parser
    ds = tmp_path / "dataset-EleutherAI_pythia_2_8b_deduped-2048-b-1-7c82f97.tokens"
    ds.write_bytes(b"\0" * 4096)
    a = build_parser().parse_args(_workflow_argv(str(ds)))
    assert a.run_name == "my-run" and a.model == "/finetune-data/models/EleutherAI/pythia-2.8b-deduped"
    assert a.dataset == str(ds) and a.trust_remote_code is False and a.tensorizer_uri == ""
    assert a.lr == 5e-5 and a.epochs == 1 and a.train_ratio == Decimal("1.0") and a.warmup_ratio == Decimal("0.1")
    assert a.bs == -1 and a.bs_divisor == Decimal("1.0") and a.gradients == 5 and a.zero_stage == 3 and a.seed == 42
    assert a.resume is True and a.shuffle is True and a.fp16 is False and a.fp16_full_eval is False
    assert a.save_steps == 500 and a.context_size == 2048 and a.prompt_file == "" and a.prompt_every == 0
    assert a.top_p == 0.95 and a.repetition_penalty == 1.1 and a.local_rank == 0 and a.log_level == "INFO"
    # the workflow's boolean parameters flipped: inverted-dest flags turn their destination off
    b = build_parser().parse_args(_workflow_argv(str(ds), no_shuffle="true", no_resume="True", force_fp16="1",
                                                 trust_remote_code="yes"))
    assert b.shuffle is False and b.resume is False and b.fp16 is True and b.trust_remote_code is True


def test_fuzzy_bool_grammar_and_validators(tmp_path):
    from kubernetes_cloud_b200.finetune import build_parser
    ds = tmp_path / "x.tokens"
    ds.write_bytes(b"\0" * 64)
    base = ["--run_name", "r", "--model", "gpt2", "--dataset", str(ds)]  # underscore aliases are accepted
    parse = lambda *extra: build_parser().parse_args(base + list(extra))  # noqa: E731
    assert parse().shuffle is True and parse().bs == -1 and parse().resume is True  # the reference's defaults
    assert parse("--no-shuffle").shuffle is False                 # bare flag
    for text in ("0", "no", "F", "False"):                        # "as if absent"
        assert parse(f"--no-shuffle={text}").shuffle is True
        assert parse(f"--fp16={text}").fp16 is False
    for text in ("1", "true", "yes", "anything"):                 # "as if present"
        assert parse(f"--no_shuffle={text}").shuffle is False
        assert parse(f"--fp16={text}").fp16 is True
    for bad in (["--bs=0"], ["--bs=-2"], ["--epochs=0"], ["--lr=-1"], ["--warmup-ratio=1.5"], ["--zero-stage=4"],
                ["--seed=4294967296"], ["--dataset", str(tmp_path / "missing.tokens")], ["--log-level=loud"],
                ["--prompt-every=-2"]):
        with pytest.raises(SystemExit) as e:
            parse(*bad)
        assert e.value.code == 2  # argparse errors exit 2, as in the reference
    assert parse("--log-level=debug").log_level == "DEBUG" and parse("--prompt-every=-1").prompt_every == -1


def test_fp16_request_is_rejected_loudly(tmp_path):
    from kubernetes_cloud_b200 import finetune
    ds = tmp_path / "x.tokens"
    ds.write_bytes(b"\0" * 64)
    with pytest.raises(SystemExit) as e:
        finetune.main(["--run-name", "r", "--model", "gpt2", "--dataset", str(ds), "--fp16"])
    assert e.value.code == 1  # runtime failures exit 1 (FT.py:442-444)


def test_fused_slab_planner():
    """ftk_lmhead_ce_fused_plan (host logic, no device): whole 256-row tiles, one slab within the 1 GiB cap, the
    bench shape split as 6,912 + 9,472 rows (37 row blocks x 8 dH tiles = exactly 4 waves over 74 SM pairs), small
    batches kept in one slab (short slabs make the dW GEMM epilogue-paced)."""
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    assert L.ftk_lmhead_ce_fused_plan(16384, 2048, 50257, 148, 2) == 9472
    for M in (1, 300, 2048, 4096):
        for d in (768, 2048, 4096):
            assert L.ftk_lmhead_ce_fused_plan(M, d, 50257, 148, 2) == (M + 255) // 256 * 256
    for M in (8192, 16384, 65536, 262144):
        for d in (768, 1024, 2048, 4096):
            for V in (50257, 50400, 32000):
                r = L.ftk_lmhead_ce_fused_plan(M, d, V, 148, 2)
                assert r % 256 == 0 and 2816 <= r <= M
                assert r * ((V + 63) // 64 * 64) * 2 <= 1 << 30
    assert L.ftk_lmhead_ce_fused_plan(0, 2048, 50257, 148, 2) == 0 and L.ftk_lmhead_ce_fused_plan(4096, 2048, 50257, 148, 3) == 0


def test_recompute_flag_parses():
    """--ftk-recompute: this command line's own switch for the reference's always-on gradient checkpointing (FT.py:884,
    :995): auto (default) / on / off; anything else is an argparse error (exit code 2 as for the reference's flags)."""
    from kubernetes_cloud_b200.finetune import build_parser
    base = ["--run-name", "r", "--model", "gpt-neo-125m", "--dataset", __file__]
    assert build_parser().parse_args(base).ftk_recompute == "auto"
    for v in ("on", "off", "auto"):
        assert build_parser().parse_args(base + [f"--ftk-recompute={v}"]).ftk_recompute == v
    with pytest.raises(SystemExit) as e:
        build_parser().parse_args(base + ["--ftk-recompute=maybe"])
    assert e.value.code == 2


def test_patch_gpt_neo_host_side_rules():
    """trunk.patch_gpt_neo only rebinds forwards (no kernel runs): which modules it takes over for which configuration,
    and what it refuses -- a model it does not know, dropout (the fused blocks have none), whole-block fusion without
    the attention kernels or with a hidden size the LayerNorm kernels do not take."""
    from transformers import GPT2Config, GPT2LMHeadModel, GPTNeoConfig, GPTNeoForCausalLM
    from kubernetes_cloud_b200 import FtkError, trunk

    def neo(**kw):
        cfg = dict(vocab_size=128, hidden_size=256, num_layers=2, num_heads=4, max_position_embeddings=128,
                   attention_types=[[["global", "local"], 1]], window_size=32)
        cfg.update(kw)
        return GPTNeoForCausalLM(GPTNeoConfig(**cfg))

    m = neo()                                        # head_dim 64, hidden 256: everything, whole blocks fused
    assert trunk.patch_gpt_neo(m) == 2 * 9 + 1 and m.config.use_cache is False
    assert trunk.patch_gpt_neo(neo(), fuse_blocks=False) == 2 * 6
    assert trunk.patch_gpt_neo(neo(hidden_size=128, num_heads=2)) == 2 * 6       # LayerNorm kernels need hidden % 256 == 0
    assert trunk.patch_gpt_neo(neo(hidden_size=256, num_heads=8)) == 2 * 5       # head_dim 32: HF's attention stays
    with pytest.raises(FtkError):
        trunk.patch_gpt_neo(neo(hidden_size=256, num_heads=8), attention_kernels=True)
    with pytest.raises(FtkError):
        trunk.patch_gpt_neo(neo(hidden_size=128, num_heads=2), fuse_blocks=True)
    with pytest.raises(FtkError):
        trunk.patch_gpt_neo(neo(resid_dropout=0.1))
    with pytest.raises(FtkError):
        trunk.patch_gpt_neo(neo(attention_dropout=0.1))
    with pytest.raises(FtkError):
        trunk.patch_gpt_neo(neo(activation_function="gelu"))
    with pytest.raises(FtkError):
        trunk.patch_gpt_neo(GPT2LMHeadModel(GPT2Config(vocab_size=128, n_embd=64, n_layer=1, n_head=2, n_positions=64)))
    # the rebound modules refuse what they do not implement instead of ignoring it (CPU tensors never reach a kernel)
    blk = m.base_model.h[0]
    x = torch.zeros(1, 128, 256)
    with pytest.raises(FtkError):
        blk(x, use_cache=True)
    with pytest.raises(FtkError):
        blk(x, head_mask=torch.ones(4))
    with pytest.raises(FtkError):
        blk(x)                                       # CPU tensors: no fallback
