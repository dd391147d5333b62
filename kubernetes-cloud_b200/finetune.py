"""Command-line finetuner on the B200 hot path, with the reference's flag names.

    python -m kubernetes_cloud_b200.finetune --run-name r --model gpt-neo-125m --dataset x.tokens --bs 1 ...
    torchrun --nproc-per-node 8 -m kubernetes_cloud_b200.finetune ...          (plain data parallel)

It keeps the parts of ``finetuner.py`` that decide the step's results (flags at FT.py:61-274; seeds FT.py:739-741;
dataset and ordering FT.py:746-772; micro-batch / accumulation / warm-up / weight decay from TrainingArguments at
FT.py:987-998; per-step samples/s as ``PerformanceCallback`` computes it, FT.py:512-522) and swaps in this package's
pieces: ``TokenizedDataset`` + ``DeviceBatcher`` for the dataset/collator/H2D path, ``compute_loss`` for lm_head + loss,
``GradSync`` for DeepSpeed's gradient reduction, ``FusedAdamW`` + ``WarmupLR`` for its optimizer.  The transformer
trunk is the HF model; for GPT-Neo its blocks run on this package's kernels (``trunk.patch_gpt_neo``).  Flags of the reference that configure things outside
the hot path (wandb, Tensorizer, prompt sampling, checkpoint discovery) are parsed with the reference's own types,
defaults and validation and then only logged, so the workflow's flag string (finetune-workflow.yaml:267-300) parses
verbatim (tests/test_host_logic.py feeds it through ``build_parser``).  ``--fp16=true`` is rejected with exit code 1:
the kernels compute in bf16 and serving fp16 requests in bf16 silently would misreport the run.

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

class FuzzyBool(argparse.Action):
    """Boolean flag with the reference's grammar (utils.py:229-292): bare ``--flag`` or ``--flag=<text>``, where the
    texts 0 / no / f / false (any case) mean "as if the flag were absent" and anything else "as if it were present".
    Present = ``not default``, so a flag declared with ``default=True`` and another ``dest`` (``--no-shuffle`` ->
    ``shuffle``) is inverted: ``--no-shuffle`` and ``--no-shuffle=true`` give shuffle False, ``--no-shuffle=false``
    gives True."""

    def __init__(self, option_strings, dest, default=False, help=None, **_):
        if not isinstance(default, bool):
            raise ValueError("FuzzyBool needs a boolean default")
        absent, present = default, not default
        super().__init__(option_strings, dest, nargs="?", const=present, default=absent, metavar="BOOL", help=help,
                         type=lambda text: absent if text.strip().lower() in ("0", "no", "f", "false") else present)

    def __call__(self, parser, namespace, values, option_string=None):
        setattr(namespace, self.dest, values)


def _number(kind, low=None, low_open=False, high=None, also=None):
    """argparse ``type=`` that parses ``kind`` and enforces a range, with one optional extra allowed value
    (the reference's validators, utils.py:297-356: positive, non-negative, at most 1, -1 as a special value)."""
    def parse(text: str):
        v = kind(text)
        if also is not None and v == also:
            return v
        if low is not None and (v <= low if low_open else v < low):
            raise argparse.ArgumentTypeError(f"must be {'greater than' if low_open else 'at least'} {low}, not {v}")
        if high is not None and v > high:
            raise argparse.ArgumentTypeError(f"must be at most {high}, not {v}")
        return v
    return parse


def _file(optional: bool):
    def parse(text: str) -> str:
        if not text:
            if optional:
                return ""
            raise argparse.ArgumentTypeError("must be specified")
        if not os.path.isfile(text):
            raise argparse.ArgumentTypeError(f"file not found: {text}")
        return text
    return parse


class DashParser(argparse.ArgumentParser):
    """Accepts ``--flag-name`` and ``--flag_name`` alike, as the reference's parser does (utils.py:163-202)."""

    def add_argument(self, *names, **kw):
        extra = []
        for n in names:
            if n.startswith("--") and "-" in n[2:]:
                extra.append("--" + n[2:].replace("-", "_"))
        return super().add_argument(*names, *extra, **kw)


# flags of the reference that configure things outside the hot path: parsed with the reference's types and defaults,
# then only logged
OUTSIDE_HOT_PATH = ["trust_remote_code", "tensorizer_uri", "train_ratio", "zero_stage", "output_path", "resume", "cache",
                    "save_steps", "project_id", "logs", "ds_config", "fp16_full_eval", "prompt_file", "prompt_every",
                    "prompt_tokens", "prompt_samples", "top_k", "top_p", "temperature", "repetition_penalty"]


def build_parser() -> DashParser:
    """Every flag of the reference's parser (FT.py:61-274) with its type, default and validation, so the flag string the
    workflow builds (finetune-workflow.yaml:267-300) parses verbatim; plus a few of this CLI's own at the end."""
    from decimal import Decimal
    p = DashParser(description="B200 finetuner (reference-compatible flags)")
    p.add_argument("--run-name", type=str, required=True)
    p.add_argument("--model", type=str, required=True, help="architecture name, local model directory or config.json")
    p.add_argument("--trust-remote-code", action=FuzzyBool, default=False)
    p.add_argument("--dataset", type=_file(optional=False), required=True, help="pre-tokenized uint16 file")
    p.add_argument("--tensorizer-uri", type=str, default="")
    p.add_argument("--lr", type=_number(float, low=0), default=5e-5)
    p.add_argument("--epochs", type=_number(int, low=0, low_open=True), default=1)
    p.add_argument("--train-ratio", type=_number(Decimal, low=0, high=1), default=Decimal("0.9"))
    p.add_argument("--warmup-ratio", type=_number(Decimal, low=0, high=1), default=Decimal("0.1"))
    p.add_argument("--eot", type=str, default="")
    p.add_argument("--pad", type=str, default="")
    p.add_argument("--bs", type=_number(int, low=0, low_open=True, also=-1), default=-1,
                   help="micro-batch per GPU (-1 = size it from free device memory)")
    p.add_argument("--bs-divisor", type=_number(Decimal, low=0, low_open=True), default=Decimal(1))
    p.add_argument("--gradients", type=_number(int, low=0, low_open=True), default=5,
                   help="gradient accumulation steps (FT.py:135-140)")
    p.add_argument("--zero-stage", type=int, default=3, choices=range(0, 4))
    p.add_argument("--seed", type=_number(int, low=0, high=(1 << 32) - 1), default=42)
    p.add_argument("--output-path", type=str, default="./")
    p.add_argument("--no-resume", action=FuzzyBool, dest="resume", default=True)
    p.add_argument("--cache", type=str, default="/tmp")
    p.add_argument("--save-steps", type=_number(int, low=0), default=500)
    p.add_argument("--context-size", type=_number(int, low=0, low_open=True), default=2048)
    p.add_argument("--project-id", type=str, default="huggingface")
    p.add_argument("--logs", type=str, default="./logs")
    p.add_argument("--ds-config", type=_file(optional=True), default="")
    p.add_argument("--fp16", action=FuzzyBool, default=False)
    p.add_argument("--fp16-full-eval", action=FuzzyBool, default=False)
    p.add_argument("--no-shuffle", action=FuzzyBool, dest="shuffle", default=True)
    p.add_argument("--prompt-file", type=_file(optional=True))
    p.add_argument("--prompt-every", type=_number(int, low=0, also=-1), default=0)
    p.add_argument("--prompt-tokens", type=_number(int, low=0), default=200)
    p.add_argument("--prompt-samples", type=_number(int, low=0), default=5)
    p.add_argument("--top-k", type=_number(int, low=0), default=50)
    p.add_argument("--top-p", type=_number(float, low=0, high=1), default=0.95)
    p.add_argument("--temperature", type=_number(float, low=0, low_open=True), default=1.0)
    p.add_argument("--repetition-penalty", type=_number(float, low=0, low_open=True), default=1.1)
    p.add_argument("--local-rank", type=_number(int, low=0, also=-1), default=-1)
    p.add_argument("--log-level", type=str.upper, default="INFO",
                   choices=("DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"))
    # this CLI's own (no network here, so ids instead of a tokenizer download; bounded runs; a metrics file)
    p.add_argument("--pad-token-id", type=int, default=50256, help="used instead of a tokenizer download")
    p.add_argument("--eos-token-id", type=int, default=50256)
    p.add_argument("--max-steps", type=int, default=-1, help="stop after this many optimizer steps")
    p.add_argument("--max-grad-norm", type=float, default=1.0)
    p.add_argument("--report-json", type=str, default="", help="write per-step metrics here")
    p.add_argument("--ftk-recompute", choices=("auto", "on", "off"), default="auto",
                   help="gradient checkpointing (FT.py:884, :995: always on in the reference, sized for 48 GB cards). "
                        "'off' keeps every block's activations for the backward instead of re-running its forward; "
                        "'auto' does so when they fit in half of the free device memory (180 GB of HBM usually do)")
    p.add_argument("--ftk-sm-reserve", type=_number(int, low=0, high=64), default=0,
                   help="data parallel only: SMs the persistent GEMM kernels leave idle so that the gradient "
                        "all-reduce's NCCL kernels find room beside them (their CTAs cannot share an SM with a GEMM "
                        "CTA that holds the whole register file); 0 = the all-reduce runs in the gaps between kernels")
    p.add_argument("--no-ftk-trunk", action=FuzzyBool, dest="ftk_trunk", default=True,
                   help="keep HF's eager GPT-Neo blocks; by default their MLP, projections and attention run on this "
                        "package's kernels (trunk.patch_gpt_neo) when the context size is a multiple of 128")
    return p


def estimate_batch_size(device, divisor, per_context_bytes: int) -> int:
    """--bs=-1.  The reference sizes the micro-batch as ceil(free device memory / (memory torch holds after loading the
    model x divisor)) (FT.py:447-466).  Same quantities and rounding here; the result is additionally capped by what
    the step itself is known to need per context (gradient-checkpointed activations + the hot path's per-token
    buffers), so the first step cannot run out of memory on a card with 180 GB."""
    import math
    torch.cuda.synchronize(device)
    torch.cuda.empty_cache()
    free, _total = torch.cuda.mem_get_info(device)
    used = max(torch.cuda.memory_allocated(device), 1)
    by_reference_rule = math.ceil(free / (used * float(divisor)))
    by_step_need = max(1, int(free * 0.6 // max(per_context_bytes, 1)))
    return max(1, min(by_reference_rule, by_step_need))


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
    parser = build_parser()
    args = parser.parse_args(argv)
    logging.basicConfig(level=getattr(logging, args.log_level), format="%(asctime)s %(name)s %(levelname)s %(message)s")
    defaults = vars(parser.parse_args(["--run-name", "x", "--model", "x", "--dataset", args.dataset]))
    for name in OUTSIDE_HOT_PATH:
        if getattr(args, name) != defaults[name]:
            logger.info("flag --%s=%s configures something outside the hot path and has no effect here",
                        name.replace("_", "-"), getattr(args, name))
    if args.fp16:
        # FT.py:195-200, :775-779, :791: fp16 weights with DeepSpeed's dynamic loss scaling.  The kernels here take
        # bf16 operands (same exponent range as fp32, so no loss scaling); serving the flag in bf16 silently would
        # change the run's numerics behind the user's back.
        logger.error("--fp16 is not supported by the B200 hot path (bf16 operands, fp32 accumulation); "
                     "drop the flag or pass --fp16=false")
        sys.exit(1)

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
    checkpointing = hasattr(model, "gradient_checkpointing_enable")
    if checkpointing:
        model.gradient_checkpointing_enable()  # FT.py:884
    if (args.ftk_trunk and getattr(model.config, "model_type", "") == "gpt_neo" and model.config.hidden_size % 64 == 0
            and getattr(model.config, "activation_function", "") == "gelu_new"):
        from . import trunk
        # the attention kernels additionally need head_dim 64 / 128 and a context that is a multiple of 128
        attn_ok = args.context_size % 128 == 0 and model.config.hidden_size // model.config.num_heads in (64, 128)
        n_mod = trunk.patch_gpt_neo(model, attention_kernels=attn_ok)
        if main_proc:
            logger.info("TRUNK: %d GPT-Neo modules rebound to this package's linear / MLP / attention / LayerNorm kernels",
                        n_mod)
    if rep.max_token >= model.config.vocab_size:
        raise SystemExit(f"token id {rep.max_token} does not fit the model's vocabulary {model.config.vocab_size}")

    bs = args.bs
    if bs == -1:  # FT.py:892-895
        cfg = model.config
        d_model = getattr(cfg, "hidden_size", getattr(cfg, "n_embd", 2048))
        n_layer = getattr(cfg, "num_layers", getattr(cfg, "n_layer", getattr(cfg, "num_hidden_layers", 24)))
        # per context: one checkpointed residual per layer + ~12 live activations of the layer being recomputed
        # (bf16), the hot path's hidden / dH rows and its share of the numerators slab
        per_ctx = args.context_size * (d_model * 2 * (n_layer + 12 + 2) + 2 * 2 * ((cfg.vocab_size + 63) // 64 * 64))
        bs = estimate_batch_size(dev, args.bs_divisor, per_ctx)
    if main_proc:
        logger.info("BS: %d", bs)
        logger.info("GAS: %d", args.gradients)
    # the split order is drawn once (FT.py:761); every epoch the Trainer's sampler re-shuffles it (seed + epoch)
    batcher = DeviceBatcher(ds, bs, seed=args.seed, shuffle=args.shuffle, rank=rank, world=world, shift_labels=True,
                            trainer_sampler=args.shuffle)
    micro_per_epoch = batcher.batches_per_epoch()
    # HF Trainer steps once every GAS micro-batches and once more on a shorter tail at the end of the epoch
    steps_per_epoch = (micro_per_epoch + args.gradients - 1) // args.gradients
    total_steps = steps_per_epoch * args.epochs
    if args.max_steps > 0:
        total_steps = min(total_steps, args.max_steps)
    if world > 1 and args.ftk_sm_reserve > 0:
        from . import _lib
        _lib.set_option("sm_reserve", args.ftk_sm_reserve)
    sync = GradSync(model.parameters(), rank, world)
    opt = FusedAdamW(sync, lr=args.lr, weight_decay=0.01, max_grad_norm=args.max_grad_norm,
                     schedule=WarmupLR(args.lr, max(2, int(float(args.warmup_ratio) * total_steps))))
    if checkpointing and args.ftk_recompute != "on":
        # kept per block and token without recomputation: ~16 bf16 [d] rows (LayerNorm inputs, q / k / v / o, the MLP's
        # input, pre-activation and activation at 4 d each, the residuals)
        cfg = model.config
        d_model = getattr(cfg, "hidden_size", getattr(cfg, "n_embd", 2048))
        n_layer = getattr(cfg, "num_layers", getattr(cfg, "n_layer", getattr(cfg, "num_hidden_layers", 24)))
        need = bs * args.context_size * n_layer * d_model * 2 * 16
        free, _total = torch.cuda.mem_get_info(dev)
        if args.ftk_recompute == "off" or need <= free // 2:
            model.gradient_checkpointing_disable()
            checkpointing = False
    if main_proc:
        logger.info("RECOMPUTE: gradient checkpointing %s", "on (as the reference)" if checkpointing else
                    "off (activations kept in HBM)")

    history = []
    step = 0
    for epoch in range(args.epochs):
        batcher.set_epoch(epoch)
        it = batcher.epoch()
        for s_in_epoch in range(steps_per_epoch):
            if step >= total_steps:
                break
            t0 = time.time()
            sync.zero_grad()
            losses = []
            # the last step of an epoch may have fewer micro-batches than GAS (micro_per_epoch % GAS of them)
            n_micro = min(args.gradients, micro_per_epoch - s_in_epoch * args.gradients)
            for g in range(n_micro):
                batch = next(it)
                n_valid = batch.pop("n_valid")  # noqa: F841  (per-row counts; the fused loss recounts)
                last = g == n_micro - 1
                ctx = sync.no_sync() if not last else _null()
                with ctx:
                    # HF Trainer scales the loss by 1/GAS before backward; here the factor goes into the fused loss,
                    # whose gradients are made during the forward call
                    loss = compute_loss(model, batch, labels_are_shifted=True, loss_scale=1.0 / n_micro)
                    loss.backward()
                losses.append(loss.detach() * n_micro)
            sync.finish()
            lr_used = opt.current_lr()
            opt.step()
            torch.cuda.synchronize(dev)
            dt = time.time() - t0
            mean_loss = float(torch.stack(losses).mean())
            rank_sps = bs * n_micro / dt  # FT.py:516-522
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
