#!/usr/bin/env python
"""BASELINE.json's metric itself: tokens/s/GPU of a whole GPT-Neo-1.3B finetune step (context 2048, 8 contexts per GPU
and step, bf16) through the reference-compatible command line (DeviceBatcher -> patched trunk -> fused loss ->
GradSync over every gradient bucket -> FusedAdamW), on 1 .. 8 GPUs of one box.  Synthetic uint16 token file, random-init
weights.  Under torchrun every rank runs this script; rank 0 prints one JSON line.
    python tools/finetune_scale.py [on|off|auto] [steps]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/finetune_scale.py off"""
import json
import os
import statistics
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from kubernetes_cloud_b200 import write_token_file  # noqa: E402
from kubernetes_cloud_b200.finetune import main  # noqa: E402

recompute = sys.argv[1] if len(sys.argv) > 1 else "auto"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
model = sys.argv[3] if len(sys.argv) > 3 else "gpt-neo-1.3b"
reserve = sys.argv[4] if len(sys.argv) > 4 else "0"  # --ftk-sm-reserve
world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
T, bs, V = 2048, 8, 50257
n_ctx = steps * bs * world + 3
path = os.path.join(tempfile.gettempdir(), f"ftk_scale_{world}_{rank}.tokens")  # same bytes on every rank
write_token_file(path, np.random.default_rng(1234).integers(0, V - 1, size=n_ctx * T - 300), T, pad_id=V - 1)
out = main(["--run-name", "scale", "--model", model, "--dataset", path, "--context-size", str(T), "--bs", str(bs),
            "--gradients", "1", "--max-steps", str(steps), "--lr", "1e-5", "--pad-token-id", str(V - 1),
            "--eos-token-id", str(V - 1), "--ftk-recompute", recompute, "--ftk-sm-reserve", reserve, "--log-level", "WARNING"])
os.remove(path)
if rank == 0:
    hist = out["history"][3:]  # the first steps include allocator growth and lazy initialisation
    dt = statistics.median(h["step_time"] for h in hist)
    print(json.dumps({"model": model, "n_gpus": world, "recompute": recompute, "sm_reserve": int(reserve),
                      "nccl_max_ctas": os.environ.get("NCCL_MAX_CTAS"), "context": T, "contexts_per_gpu_per_step": bs,
                      "ms_per_step_median": round(dt * 1e3, 2), "tokens_per_s": round(world * bs * T / dt),
                      "tokens_per_s_per_gpu": round(bs * T / dt), "loss_first": round(out["history"][0]["loss"], 4),
                      "loss_last": round(out["history"][-1]["loss"], 4), "steps": out["steps"],
                      "step_ms_all": [round(h["step_time"] * 1e3, 1) for h in out["history"]]}))
