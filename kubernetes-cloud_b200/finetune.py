"""Command-line finetuner on the B200 hot path, with the reference's flag names.

    python -m kubernetes_cloud_b200.finetune --run-name r --model gpt-neo-125m --dataset x.tokens --bs 1 ...
    torchrun --nproc-per-node 8 -m kubernetes_cloud_b200.finetune ...          (plain data parallel)

It keeps the parts of ``finetuner.py`` that decide the step's results (flags at FT.py:61-274; seeds FT.py:739-741;
dataset and ordering FT.py:746-772; micro-batch / accumulation / warm-up / weight decay from TrainingArguments at
FT.py:987-998; per-step samples/s as ``PerformanceCallback`` computes it, FT.py:512-522) and swaps in this package's
pieces: ``TokenizedDataset`` + ``DeviceBatcher`` for the dataset/collator/H2D path, ``compute_loss`` for lm_head + loss,
``GradSync`` for DeepSpeed's gradient reduction, ``FusedAdamW`` + ``WarmupLR`` for its optimizer.  The transformer
trunk is the HF model (third-party, out of this tier's scope).  Flags of the reference that configure things outside
the hot path (wandb, Tensorizer, prompt sampling, checkpoint discovery) are accepted and ignored with a log line, so
the workflow's flag string (finetune-workflow.yaml:267-300) parses unchanged.

There is no network in the build environment: ``--model`` is an architecture name (random-init from the literal HF
config, as BASELINE.json's configs ask), or a local directory / config.json.
"""
from __future__ import annotations

import argparse
import json
import logging
import os
import sys
import time
from typing import List, Optional

import torch

logger = logging.getLogger("finetune")

# literal configs: no network, so architectures are built from numbers
ARCHS = {
    "gpt-neo-125m": ("gpt_neo", dict(vocab_size=50257, hidden_size=768, num_layers=12, num_heads=12,
                                      max_position_embeddings=2048, attention_types=[[["global", "local"], 6]],
                                      window_size=256)),
    "gpt-neo-1.3b": ("gpt_neo", dict()),   # GPTNeoConfig() defaults: 2048 / 24 layers / 16 heads, V 50257
    "gpt2": ("gpt2", dict()),              # GPT2Config() defaults: 768 / 12 / 12, ctx 1024, V 50257
    "gpt-j-6b": ("gptj", dict()),          # GPTJConfig() defaults: 4096 / 28 / 16, V 50400, untied, lm_head bias
}

IGNORED_FLAGS = ["--cache", "--output-path", "--logs", "--save-steps", "--project-id", "--prompt-file",
                 "--prompt-every", "--prompt-tokens", "--prompt-samples", "--top-k", "--top-p", "--temperature",
                 "--repetition-penalty", "--tensorizer-uri", "--zero-stage", "--train-ratio", "--force-fp32-attn"]
IGNORED_BOOLS = ["--no-resume", "--fp16"]


class DashParser(argparse.ArgumentParser):
    """Accepts ``--flag-name`` and ``--flag_name`` alike, as the reference's parser does (utils.py:163-202)."""

    def add_argument(self, *names, **kw):
        extra = []
        for n in names:
            if n.startswith("--") and "-" in n[2:]:
                extra.append("--" + n[2:].replace("-", "_"))
        return super().add_argument(*names, *extra, **kw)


def build_parser() -> DashParser:
    p = DashParser(description="B200 finetuner (reference-compatible flags)")
    p.add_argument("--run-name", type=str, required=True)
    p.add_argument("--model", type=str, required=True, help="architecture name, local model directory or config.json")
    p.add_argument("--dataset", type=str, required=True, help="pre-tokenized uint16 file (FT.py:81-86)")
    p.add_argument("--lr", type=float, default=5e-5)
    p.add_argument("--epochs", type=int, default=1)
    p.add_argument("--bs", type=int, default=1, help="micro-batch per GPU; the reference's -1 (estimate) is not kept")
    p.add_argument("--gradients", type=int, default=5, help="gradient accumulation steps (FT.py:135-140)")
    p.add_argument("--seed", type=int, default=42)
    p.add_argument("--context-size", type=int, default=2048)
    p.add_argument("--warmup-ratio", type=float, default=0.1)
    p.add_argument("--eot", type=str, default="")
    p.add_argument("--pad", type=str, default="")
    p.add_argument("--pad-token-id", type=int, default=50256, help="used instead of a tokenizer download")
    p.add_argument("--eos-token-id", type=int, default=50256)
    p.add_argument("--no-shuffle", action="store_false", dest="shuffle")
    p.add_argument("--local-rank", type=int, default=-1)
    p.add_argument("--max-steps", type=int, default=-1, help="stop after this many optimizer steps")
    p.add_argument("--max-grad-norm", type=float, default=1.0)
    p.add_argument("--report-json", type=str, default="", help="write per-step metrics here")
    for f in IGNORED_FLAGS:
        p.add_argument(f, type=str, default=None)
    for f in IGNORED_BOOLS:
        p.add_argument(f, nargs="?", const="true", default=None)
    return p


def load_model(spec: str):
    from transformers import AutoConfig, AutoModelForCausalLM
    key = spec.lower().split("/")[-1]
    if key in ARCHS:
        kind, kw = ARCHS[key]
        from transformers import GPT2Config, GPTJConfig, GPTNeoConfig
        cfg = {"gpt_neo": GPTNeoConfig, "gpt2": GPT2Config, "gptj": GPTJConfig}[kind](**kw)
    elif os.path.exists(spec):
        cfg = AutoConfig.from_pretrained(spec)
    else:
        raise SystemExit(f"--model {spec!r}: not a known architecture {sorted(ARCHS)} nor a local path (no network)")
    cfg.use_cache = False  # FT.py:888
    return AutoModelForCausalLM.from_config(cfg)


def main(argv: Optional[List[str]] = None) -> dict:
    args = build_parser().parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(name)s %(levelname)s %(message)s")
    for f in IGNORED_FLAGS + IGNORED_BOOLS:
        v = getattr(args, f[2:].replace("-", "_"))
        if v is not None:
            logger.info("flag %s=%s is outside the hot path and is ignored here", f, v)

    import torch.distributed as dist
    from . import DeviceBatcher, FusedAdamW, GradSync, TokenizedDataset, WarmupLR, validate_token_file
    from .loss import compute_loss

    world = int(os.environ.get("WORLD_SIZE", "1"))  # FT.py:318
    rank = int(os.environ.get("RANK", "0"))
    local = args.local_rank if args.local_rank != -1 else int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)  # FT.py:699
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    main_proc = rank == 0

    torch.manual_seed(args.seed)  # FT.py:739-741
    import random
    import numpy
    random.seed(args.seed)
    numpy.random.seed(args.seed)

    rep = validate_token_file(args.dataset, args.context_size, pad_id=args.pad_token_id, eos_id=args.eos_token_id)
    for prob in rep.problems:
        logger.warning("DATASET: %s", prob)
    ds = TokenizedDataset(args.dataset, context_length=args.context_size, pad_token_id=args.pad_token_id,
                          eos_token_id=args.eos_token_id, device=dev)
    if main_proc:
        logger.info("DATASET: %s, %d tokens, %d contexts", args.dataset, len(ds) * args.context_size, len(ds))

    model = load_model(args.model).to(dev).to(torch.bfloat16)
    model.train()  # FT.py:831
    if hasattr(model, "gradient_checkpointing_enable"):
        model.gradient_checkpointing_enable()  # FT.py:884
    if rep.max_token >= model.config.vocab_size:
        raise SystemExit(f"token id {rep.max_token} does not fit the model's vocabulary {model.config.vocab_size}")

    batcher = DeviceBatcher(ds, args.bs, seed=args.seed, shuffle=args.shuffle, rank=rank, world=world, shift_labels=True)
    micro_per_epoch = batcher.batches_per_epoch()
    steps_per_epoch = max(1, micro_per_epoch // args.gradients)
    total_steps = steps_per_epoch * args.epochs
    if args.max_steps > 0:
        total_steps = min(total_steps, args.max_steps)
    sync = GradSync(model.parameters(), rank, world)
    opt = FusedAdamW(sync, lr=args.lr, weight_decay=0.01, max_grad_norm=args.max_grad_norm,
                     schedule=WarmupLR(args.lr, max(2, int(args.warmup_ratio * total_steps))))

    history = []
    step = 0
    for epoch in range(args.epochs):
        it = batcher.epoch()
        for _ in range(steps_per_epoch):
            if step >= total_steps:
                break
            t0 = time.time()
            sync.zero_grad()
            losses = []
            for g in range(args.gradients):
                batch = next(it)
                n_valid = batch.pop("n_valid")  # noqa: F841  (per-row counts; the fused loss recounts)
                last = g == args.gradients - 1
                ctx = sync.no_sync() if not last else _null()
                with ctx:
                    # HF Trainer scales the loss by 1/GAS before backward; here the factor goes into the fused loss,
                    # whose gradients are made during the forward call
                    loss = compute_loss(model, batch, labels_are_shifted=True, loss_scale=1.0 / args.gradients)
                    loss.backward()
                losses.append(loss.detach() * args.gradients)
            sync.finish()
            lr_used = opt.current_lr()
            opt.step()
            torch.cuda.synchronize(dev)
            dt = time.time() - t0
            mean_loss = float(torch.stack(losses).mean())
            rank_sps = args.bs * args.gradients / dt  # FT.py:516-522
            rec = {"step": step, "loss": mean_loss, "lr": lr_used, "grad_norm": float(opt.grad_norm()),
                   "rank_samples_per_second": rank_sps, "world_samples_per_second": world * rank_sps,
                   "tokens_per_second": world * rank_sps * args.context_size, "step_time": dt}
            history.append(rec)
            if main_proc:
                print(f"\nLOSS: {mean_loss:.3f} step {step} {rec['tokens_per_second']:.0f} tokens/s", file=sys.stderr,
                      flush=True)  # FT.py:486-491
            step += 1
    sync.close()
    out = {"run_name": args.run_name, "steps": step, "history": history}
    if args.report_json and main_proc:
        with open(args.report_json, "w") as f:
            json.dump(out, f)
    if world > 1:
        dist.destroy_process_group()
    return out


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


if __name__ == "__main__":
    main()
