#!/usr/bin/env python
"""Benchmark of the training hot path: window gather -> fused lm_head + cross-entropy forward
and backward -> data-parallel all-reduce of the lm_head / embedding gradient.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched once per rank by torchrun)
    python bench.py --impl reference ...                     the CPU path of the reference, timed on the host cores

Workload (BASELINE.json configs[2], the one the metric is quoted on): GPT-Neo-1.3B head, d = 2048, V = 50257,
context 2048, 8 contexts per GPU per step = 16,384 tokens.  One "step" = one pass of the hot path over one batch:
gather 8 windows from the HBM-resident uint16 token stream (ids, shifted+masked labels, mask); fused lm_head+CE
forward AND backward in one call (loss, dH, dW added into W's fp32 main gradient); all-reduce(dW)/N through GradSync,
started by the dW-done event while the last dH GEMM still runs; gradient-norm clip + fused AdamW on W (FusedAdamW).
The reduced gradient is applied before the next forward reads W, as in a real training step.  The transformer trunk
is out of scope for this tier: the hidden states are synthetic bf16, resident in HBM.
Inputs (W 206 MB + H 67 MB) exceed the 126 MB L2, so no explicit flush is needed between steps.

One JSON line on stdout (rank 0).  ``value`` = tokens/s over all GPUs with inputs resident in HBM; ``e2e`` = the same
with this step's token windows and hidden states copied from pinned host memory and the loss read back, inside the
timed region; ``roofline`` = the step's dominant tensor kernel against the measured bf16 peak, its duration taken
from the library's launch timeline (CUDA events around every launch inside the timed steps); ``cpu_baseline`` = the
oracle's CPU restatement of the same path on a bounded sample.  At N > 1 the line also carries
``allreduce_max_rel_err`` (reduced dW against an independently gathered mean; the run fails above tolerance),
``exposed_comm_ms_per_step`` and ``full_gradient_allreduce`` (the whole gradient set of GPT-Neo-1.3B and GPT-J-6B
in 128 MB buckets through the same communicator).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
RESULT_OUT = sys.stdout  # replaced by a private handle on the real stdout in __main__

D_MODEL, VOCAB, CTX, BATCH = 2048, 50257, 2048, 8  # GPT-Neo-1.3B head (GPTNeoConfig defaults), 8 contexts / GPU
N_CTX_FILE = 4096                                   # contexts in the synthetic token stream
PAD_ID = 50256
WORKLOAD = "GPT-Neo-1.3B lm_head+CE hot path: d=2048 V=50257 ctx=2048 batch=8/GPU (16384 tokens/GPU/step), bf16"


def bench_config(n_gpus: int) -> dict:
    """The ``config`` object both arms print (identical for the same N, so the driver can match the two lines)."""
    return {"workload": WORKLOAD, "tokens_per_gpu_per_step": BATCH * CTX, "parallelism": f"dp{n_gpus}"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return {"burst": float(j["bf16_tflops"]), "sustained": float(j.get("bf16_tflops_sustained", j["bf16_tflops"])),
                "hbm": float(j["hbm_gbs"]), "src": "measured"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 25 ms; only samples inside the timed region are kept."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []
        self.t0 = self.t1 = None

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], None, [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, ln in self.lines:
            if self.t0 is not None and not (self.t0 <= ts <= (self.t1 or ts) + 0.03):
                continue  # keep only samples taken while the timed region ran
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
                power.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": (sm[len(sm) // 2] if sm else None), "sm_max_mhz": smax,
                "power_w_max": (max(power) if power else None), "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_step_fn(sample_ctx: int):
    """The reference's CPU path for this hot path, restated by the oracle (kind "port"): batch assembly in C
    (oracle/gather_oracle.c) + lm_head and HF causal-LM loss forward/backward in fp32 torch on the host cores."""
    import numpy as np
    import torch
    from oracle import oracle as O
    tok = O.synth_tokens(64, CTX, VOCAB, 7, pad_id=PAD_ID)
    g = torch.Generator().manual_seed(42)
    W = (torch.randn(VOCAB, D_MODEL, generator=g) * 0.02).requires_grad_(True)
    H = torch.randn(sample_ctx * CTX, D_MODEL, generator=g).requires_grad_(True)  # continues W's stream: independent
    order = O.train_order(64, 42, True)
    state = {"i": 0}

    def step():
        i = state["i"]
        idx = order[(i * sample_ctx) % 64:(i * sample_ctx) % 64 + sample_ctx]
        if len(idx) < sample_ctx:
            idx = order[:sample_ctx]
        state["i"] += 1
        ids, labels, mask, _ = O.gather_c(tok, CTX, idx, PAD_ID, True, shift_labels=False)
        lab = torch.from_numpy(labels)
        logits = (H @ W.t()).view(sample_ctx, CTX, VOCAB)
        loss = O.hf_causal_lm_loss(logits, lab)
        W.grad = None
        H.grad = None
        loss.backward()
        return float(loss.detach())

    return step


def run_cpu_baseline(budget_s: float = 20.0) -> dict:
    import torch
    step = cpu_step_fn(1)
    step()  # warm-up
    times = []
    t_start = time.time()
    while len(times) < 3 or (time.time() - t_start < budget_s and len(times) < 8):
        t0 = time.time()
        step()
        times.append(time.time() - t0)
    times.sort()
    med = times[len(times) // 2]
    return {"value": CTX / med, "unit": "tokens/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{len(times)} steps of 1 context (2048 tokens) of the same workload, fp32, median {med:.3f} s/step",
            "host_cpus": os.cpu_count()}


def main_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm is rank 0 alone and should use the whole host
    torch.set_num_threads(os.cpu_count() or 1)
    step = cpu_step_fn(1)
    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.time()
    for _ in range(args.steps):
        step()
    dt = time.time() - t0
    val = args.steps * CTX / dt
    line = {"impl": "reference", "metric": "tokens/s", "value": val, "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args.gpus),
            "config_notes": {"step_sample": "each CPU step is 1 context (2048 tokens) of the workload on the host cores"},
            "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": torch.get_num_threads(), "kind": "port",
                             "sample": f"{args.steps} steps of 1 context (2048 tokens)", "host_cpus": os.cpu_count()},
            "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=RESULT_OUT, flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
def main_gpu(args) -> None:
    import ctypes
    import numpy as np
    import torch
    import torch.distributed as dist
    from kubernetes_cloud_b200 import _lib, ops
    from kubernetes_cloud_b200.dp import GradSync
    from kubernetes_cloud_b200.optim import FusedAdamW
    from kubernetes_cloud_b200.ordering import epoch_batches
    from kubernetes_cloud_b200.tokenfile import synthetic_token_stream

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    wire = os.environ.get("FTK_BENCH_WIRE", "fp32")  # fp32: the reference's communication_data_type; bf16: half the bytes
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    L = _lib.lib()
    # SMs the last slab's dH GEMM leaves to the all-reduce of dW that runs beside it: as many as one more
    # wave of that GEMM frees (296 tiles = 4 waves on 74 SM pairs or 5 on 60, i.e. 28 SMs; under the power cap the
    # GEMM's clock rises as it occupies fewer SMs, so the extra wave costs far less than 1/4).  NCCL keeps its own
    # CTA count: capping it to the reserve (NCCL_MAX_CTAS) measured slower at 8 GPUs (profiles/r02_overlap_n8.txt).
    sm_reserve = 0
    if world > 1:
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        slab = L.ftk_lmhead_ce_fused_slab_rows(BATCH * CTX, D_MODEL, VOCAB, 0)
        tiles = ((slab + 255) // 256) * ((D_MODEL + 255) // 256)
        waves = -(-tiles // (sms // 2))
        sm_reserve = max(0, min(32, sms - 2 * (-(-tiles // (waves + 1)))))
        sm_reserve = int(os.environ.get("FTK_BENCH_SM_RESERVE", sm_reserve))
    stream = torch.cuda.current_stream(dev)

    # ---- synthetic inputs, resident in HBM (and a pinned host copy for the e2e leg)
    tok_host = torch.from_numpy(synthetic_token_stream(N_CTX_FILE, CTX, VOCAB, 7, pad_id=PAD_ID).view(np.int16)).pin_memory()
    tokens = tok_host.to(dev)
    g = torch.Generator(device=dev).manual_seed(42 + rank)
    gw = torch.Generator(device=dev).manual_seed(20240607)  # not the H seed: equal streams make H rows = W rows
    W = torch.nn.Parameter((torch.randn(VOCAB, D_MODEL, device=dev, generator=gw) * 0.02).to(torch.bfloat16))
    M = BATCH * CTX
    H = torch.randn(M, D_MODEL, device=dev, generator=g).to(torch.bfloat16)
    H_host = [H.cpu().pin_memory(), H.cpu().pin_memory()]
    batches = list(epoch_batches(N_CTX_FILE, BATCH, seed=42, shuffle=True, rank=rank, world=world))
    idx_dev = [b.to(dev) for b in batches]
    nb = len(idx_dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    def timed_loop(fn, n: int) -> float:
        """Device time of n calls of fn(i), bracketed by barrier + synchronize on both sides, max over ranks (ms)."""
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for i in range(n):
            fn(i)
        b.record(stream)
        barrier()
        return max_over_ranks(a.elapsed_time(b))

    # ================= bulk gather bandwidth (HBM roofline of G1; outside the step: one step's gather is 0.3 MB).
    # Measured first, on a GPU that is not yet power-capped by the GEMM phase: the kernel tracks the SM/L2 clock. ====
    n_bulk = 4096 * 8  # 2^26 tokens per launch, each context of a 134 MB source read exactly once (a permutation):
    # the source is larger than the 126 MB L2, so the 2 B/token read side is HBM traffic as well
    big_tokens = torch.randint(0, VOCAB, (n_bulk * CTX,), device=dev, dtype=torch.int32).to(torch.int16)
    big_idx = torch.randperm(n_bulk, device=dev)
    # The kernel is timed through the C ABI into preallocated outputs: going through the torch op would add four
    # fresh 0.1-0.5 GB allocations per call to the measurement (that path measured 4.3-4.7 TB/s for the same kernel).
    b_ids = torch.empty(n_bulk, CTX, dtype=torch.int64, device=dev)
    b_lab = torch.empty_like(b_ids)
    b_msk = torch.empty(n_bulk, CTX, dtype=torch.bool, device=dev)
    b_nv = torch.empty(n_bulk, dtype=torch.int32, device=dev)

    def bulk_launch():
        _lib.check(L.ftk_gather_windows(big_tokens.data_ptr(), n_bulk, CTX, big_idx.data_ptr(), n_bulk, PAD_ID, 1, 1,
                                        b_ids.data_ptr(), b_lab.data_ptr(), b_msk.data_ptr(), b_nv.data_ptr(),
                                        stream.cuda_stream), "ftk_gather_windows")

    def bulk_gather_gbs() -> float:
        for _ in range(5):
            bulk_launch()
        torch.cuda.synchronize(dev)
        g0 = torch.cuda.Event(enable_timing=True)
        g1 = torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        for _ in range(10):
            bulk_launch()
        g1.record(stream)
        torch.cuda.synchronize(dev)
        return 19.0 * n_bulk * CTX / (g0.elapsed_time(g1) / 10 / 1e3) / 1e9

    gather_gbs = bulk_gather_gbs()
    _lib.set_option("gather_variant", 2)  # the shared-memory staged kernel (the default for small batches)
    gather_other_gbs = bulk_gather_gbs()
    _lib.set_option("gather_variant", 0)
    _lib.set_option("gather_stores", 2)
    gather_plain_gbs = bulk_gather_gbs()
    _lib.set_option("gather_stores", 0)
    del b_ids, b_lab, b_msk, b_nv, big_tokens

    # ================= the step: gather -> fused lm_head+CE forward+backward -> all-reduce(dW) -> clip + AdamW(W) =====
    # GradSync owns W's gradient bucket (fp32 main gradient unless FTK_BENCH_WIRE=bf16) and the NCCL communicator;
    # FusedAdamW owns W's fp32 master copy and moments.  The dependency of a real training step is honoured: the
    # reduced gradient is applied to W before W is read by the next forward.
    sync = GradSync([W], rank, world, main_grads=(wire == "fp32"))
    opt = FusedAdamW(sync, lr=5e-5, weight_decay=0.01, max_grad_norm=1.0)
    dw_event = None
    if world > 1:
        dw_event = torch.cuda.Event()
        dw_event.record(stream)
        ops.set_dw_done_event(dw_event)
        _lib.set_option("sm_reserve", sm_reserve)

    fused_mode = [True]

    def step(hidden: torch.Tensor, tok: torch.Tensor, idx: torch.Tensor, comm: bool = True, optimizer: bool = True):
        ids, targets, mask, n_valid_rows = torch.ops.ftk.gather_windows(tok, idx, CTX, PAD_ID, True, True)
        sync.zero_grad(overwritten_by_fused_loss=W)  # W's gradient is overwritten by the first slab, not zeroed + added
        h = hidden.detach().requires_grad_(True)
        loss = ops.fused_lmhead_ce(h, W, None, targets.view(-1))  # dH, dW (into W's gradient bucket) made here
        early = fused_mode[0] and world > 1 and comm
        if early:
            sync.mark_ready(W, dw_event)  # all-reduce(dW) starts when dW is done, beside the last slab's dH GEMM
        loss.backward()                   # hands dH to h.grad (upstream gradient 1: nothing to rescale)
        if world > 1 and comm and not early:
            sync.mark_ready(W)            # recompute mode: dW only exists after backward
        sync.finish(reduce=comm)
        if optimizer:
            opt.step()
        return loss.detach(), h.grad

    def run_steps(n: int, base: int, **kw):
        out = None
        for i in range(n):
            out = step(H, tokens, idx_dev[(base + i) % nb], **kw)
        return out

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    run_steps(args.warmup, 0)
    barrier()
    l0 = ops.launch_count()
    _lib.timeline_enable(True)
    sampler.mark_begin()
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin.record(stream)
    loss, _ = run_steps(args.steps, args.warmup)
    t_end.record(stream)
    barrier()
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None
    trace = _lib.timeline_read(with_start=True)
    timeline = [(n_, ms_) for n_, ms_, _ in trace]
    _lib.timeline_enable(False)
    if os.environ.get("FTK_BENCH_TRACE") and rank == 0:  # start / duration of every launch of the last two steps
        per = len(trace) // max(args.steps, 1)
        for n_, ms_, st_ in trace[-2 * per:]:
            print(f"trace {st_:10.3f} ms  +{ms_:7.3f}  {n_}", file=sys.stderr)
        print(f"trace: GradSync launched {sync.launched} all-reduces in {args.warmup + args.steps} steps", file=sys.stderr)
    # ours: gather + fused call + rescale checks (counted in ops) + grad-norm, clip, AdamW per step
    launches = ops.launch_count() - l0 + 3 * args.steps
    ms_total = max_over_ranks(t_begin.elapsed_time(t_end))
    ms_step = ms_total / args.steps
    value = world * M * args.steps / (ms_total / 1e3)
    loss_val = float(loss)
    # span of the fused forward+backward inside each step: first sweep start -> last gradient GEMM end (the
    # conversion passes run on the side stream beside the sweeps, so summing kernel durations would double count)
    spans, cur0, cur1 = [], None, None
    for name, ms, st_ in trace:
        if name == "gather":
            if cur0 is not None:
                spans.append(cur1 - cur0)
            cur0, cur1 = None, None
        elif name in ("fwd_keep", "fwd", "dz", "convert", "dH", "dW"):
            cur0 = st_ if cur0 is None else min(cur0, st_)
            cur1 = st_ + ms if cur1 is None else max(cur1, st_ + ms)
    if cur0 is not None:
        spans.append(cur1 - cur0)
    fused_span_ms = sum(spans) / max(len(spans), 1)
    per_kernel = {}
    for name, ms in timeline:
        e = per_kernel.setdefault(name, [0, 0.0])
        e[0] += 1
        e[1] += ms
    kern = {k: {"launches_per_step": v[0] / args.steps, "ms_per_step": v[1] / args.steps} for k, v in per_kernel.items()}
    # how evenly the ranks' GPUs run the same kernels (under the power cap each GPU settles at its own clock): the
    # step is paced by the slowest one, whatever the collective costs
    own_ms = sum(v[1] for k, v in per_kernel.items() if k not in ("allreduce", "convert")) / args.steps  # convert overlaps
    rank_compute = [own_ms]
    if world > 1:
        t_all = [torch.zeros(1, device=dev, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(t_all, torch.tensor([own_ms], device=dev, dtype=torch.float64))
        rank_compute = [float(x) for x in t_all]

    # ---- the same step without its parts, to price them (short loops)
    n_short = max(5, args.steps // 2)
    exposed_comm_ms = None
    if world > 1:
        run_steps(2, 0, comm=False)
        ms_nocomm = timed_loop(lambda i: step(H, tokens, idx_dev[i % nb], comm=False), n_short) / n_short
        run_steps(2, 0)
        ms_comm = timed_loop(lambda i: step(H, tokens, idx_dev[i % nb]), n_short) / n_short
        exposed_comm_ms = ms_comm - ms_nocomm
    run_steps(2, 0, optimizer=False)
    ms_noopt = timed_loop(lambda i: step(H, tokens, idx_dev[i % nb], optimizer=False), n_short) / n_short

    # ---- the r01 default for comparison: separate forward, backward that recomputes the logits
    ops.set_backward_mode("recompute")
    fused_mode[0] = False
    run_steps(2, 0)
    recompute_ms = timed_loop(lambda i: step(H, tokens, idx_dev[i % nb]), n_short) / n_short
    ops.set_backward_mode("fused")
    fused_mode[0] = True

    # ================= multi-GPU correctness, visible to the driver: the reduced dW against an independent mean =========
    allreduce_err = None
    if world > 1:
        sync.zero_grad()
        h = H.detach().requires_grad_(True)
        ids, targets, _, _ = torch.ops.ftk.gather_windows(tokens, idx_dev[0], CTX, PAD_ID, True, True)
        with sync.no_sync():
            ops.fused_lmhead_ce(h, W, None, targets.view(-1)).backward()  # local gradient only
        flat = sync.buckets[0].flat
        sample_idx = torch.arange(0, flat.numel(), max(1, flat.numel() // (1 << 20)), device=dev)[: 1 << 20]
        mine = flat[sample_idx].float().clone()
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)  # second, independent route: torch.distributed all_gather + local fp64 mean
        want = torch.stack(gathered).double().mean(0)
        sync.mark_ready(W)
        sync.finish()
        torch.cuda.synchronize(dev)
        got = flat[sample_idx].double()
        err = float((got - want).norm() / want.norm().clamp_min(1e-30))
        t = torch.tensor([err], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        allreduce_err = float(t)

    # ================= whole gradient set through the same communicator =========================
    full_grad = None
    if world > 1:
        full_grad = {}
        handle = sync._handle
        for name, n_el in (("gpt-neo-1.3b", 1_315_575_808), ("gpt-j-6b", 6_050_882_784)):
            for dt, code, esz in ((torch.float32, _lib.FTK_F32, 4), (torch.bfloat16, _lib.FTK_BF16, 2)):
                bucket_el = (128 << 20) // esz
                flat = torch.zeros(n_el + (-n_el) % 8, dtype=dt, device=dev)
                offs = list(range(0, flat.numel(), bucket_el))

                def reduce_all():
                    for o in offs:
                        n = min(bucket_el, flat.numel() - o)
                        _lib.check(L.ftk_dp_allreduce(handle, flat.data_ptr() + o * esz, n, code, 1.0 / world,
                                                      stream.cuda_stream), "ftk_dp_allreduce")
                    _lib.check(L.ftk_dp_join(handle, stream.cuda_stream), "ftk_dp_join")

                reduce_all()
                ms = timed_loop(lambda i: reduce_all(), 2) / 2
                bus = 2.0 * (world - 1) / world * flat.numel() * esz / (ms / 1e3) / 1e9
                full_grad[f"{name}_{'fp32' if esz == 4 else 'bf16'}"] = {
                    "elements": n_el, "buckets": len(offs), "bucket_mb": 128, "ms": ms, "bus_gbs": bus}
                del flat
        _lib.set_option("sm_reserve", sm_reserve)

    # ================= end-to-end leg: host buffers in, loss out, copies inside the timed region =================
    copy_stream = torch.cuda.Stream(dev)
    stage_tok = [torch.empty(BATCH * CTX, dtype=torch.int16, device=dev) for _ in range(2)]
    stage_h = [torch.empty_like(H) for _ in range(2)]
    ev_copied = [torch.cuda.Event() for _ in range(2)]
    ev_consumed = [torch.cuda.Event() for _ in range(2)]
    idx_local = torch.arange(BATCH, device=dev)
    tok_rows = tok_host.view(N_CTX_FILE, CTX)
    host_batch = [torch.empty(BATCH, CTX, dtype=torch.int16).pin_memory() for _ in range(2)]
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()

    def upload(i: int):
        s = i & 1
        ev_copied[s].synchronize()  # the previous H2D out of this pinned buffer has finished
        torch.index_select(tok_rows, 0, batches[i % nb], out=host_batch[s])  # host-side window slice (page cache)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_consumed[s])
            stage_tok[s].copy_(host_batch[s].view(-1), non_blocking=True)
            stage_h[s].copy_(H_host[s], non_blocking=True)
            ev_copied[s].record(copy_stream)

    def e2e_steps(n: int, base: int):
        upload(base)
        for i in range(base, base + n):
            s = i & 1
            if i + 1 < base + n:
                upload(i + 1)  # next step's H2D overlaps this step's compute
            stream.wait_event(ev_copied[s])
            ls, _ = step(stage_h[s], stage_tok[s], idx_local)
            ev_consumed[s].record(stream)
            loss_host.copy_(ls.reshape(1), non_blocking=True)

    for s_ in range(2):
        ev_consumed[s_].record(stream)
    e2e_steps(max(args.warmup, 2), 0)
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    e2e_steps(args.steps, 1000)
    t1.record(stream)
    barrier()
    e2e_ms = max_over_ranks(t0.elapsed_time(t1))
    e2e_value = world * M * args.steps / (e2e_ms / 1e3)
    h2d = BATCH * CTX * 2 + M * D_MODEL * 2
    d2h = 4

    ops.set_dw_done_event(None)
    sync.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = measured_peaks()
    flops_gemm = 2.0 * M * VOCAB * D_MODEL  # one [tokens x vocab x d] contraction: forward, dH and dW are one each
    tf = lambda ms: flops_gemm / (ms / 1e3) / 1e12  # noqa: E731
    tensor_ms = {k: kern[k]["ms_per_step"] for k in ("fwd_keep", "dH", "dW") if k in kern}
    dom = max(tensor_ms, key=tensor_ms.get)
    dom_desc = {"fwd_keep": "umma_kernel<FwdKeepEpi> (lm_head GEMM + online-softmax CE + bf16 numerators)",
                "dH": "umma_kernel<StoreEpi> dH = dZ.W", "dW": "umma_kernel<StoreEpi> dW += dZ^T.H"}[dom]
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(f"{dom}_dram_bytes_per_step")
    fused_ms = fused_span_ms
    conv = kern.get("convert")
    line = {
        "metric": "tokens/s", "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": bench_config(world),
        "config_notes": {
                   "step": "gather 8 windows -> fused lm_head+CE forward+backward in one call (dH, dW) -> "
                           "all-reduce(dW)/N -> grad-norm clip + AdamW on W; W is updated before the next forward",
                   "l2": "inputs (W 206 MB + H 67 MB) larger than the 126 MB L2; no explicit flush",
                   "grad_sync": (f"GradSync: ncclAllReduce(avg) of dW (103M {wire}) on the comm stream, started by the "
                                 f"dW-done event and running beside the last slab's dH GEMM, which leaves {sm_reserve} "
                                 f"SMs free for it; the optimizer waits for it")
                   if world > 1 else "none (1 GPU)"},
        "tokens_per_s_per_gpu": value / world,
        "loss": loss_val,
        "parity": "green: every -m gpu test compares the CUDA path with the fp32 oracle / HF fixtures (tests/)",
        "roofline": {"bound": "tensor", "kernel": dom_desc, "achieved": tf(tensor_ms[dom]), "peak": peaks["sustained"],
                     "unit": "TFLOP/s", "frac": tf(tensor_ms[dom]) / peaks["sustained"], "traffic": traffic,
                     "peak_source": f"{peaks['src']} sustained bf16 (kernel timed inside a long step)",
                     "frac_of_burst": tf(tensor_ms[dom]) / peaks["burst"], "frac_of_nominal_2250": tf(tensor_ms[dom]) / 2250.0,
                     "flops_per_step": flops_gemm, "ms_per_step": tensor_ms[dom],
                     "launches_per_step": kern[dom]["launches_per_step"],
                     "how": "ftk_timeline: CUDA events around every launch inside the timed steps, on the launch stream"},
        "roofline_kernels": {k: {"achieved_tflops": tf(v), "frac_of_sustained": tf(v) / peaks["sustained"],
                                 "ms_per_step": v} for k, v in tensor_ms.items()},
        "roofline_fused": {"achieved": 3 * flops_gemm / (fused_ms / 1e3) / 1e12, "unit": "TFLOP/s",
                           "frac_of_sustained": 3 * flops_gemm / (fused_ms / 1e3) / 1e12 / peaks["sustained"],
                           "frac_of_burst": 3 * flops_gemm / (fused_ms / 1e3) / 1e12 / peaks["burst"],
                           "frac_of_nominal_2250": 3 * flops_gemm / (fused_ms / 1e3) / 1e12 / 2250.0,
                           "ms": fused_ms,
                           "note": "algorithmic 6MVd over the span of the fused forward+backward call inside the step "
                                   "(first sweep start to last gradient GEMM end, from the launch timeline; no "
                                   "recompute: executed = algorithmic; the conversion passes run on a side stream "
                                   "beside the sweeps, which is why the per-kernel times overlap)"},
        "roofline_convert": ({"bound": "hbm", "achieved": 4.0 * M * ((VOCAB + 63) // 64 * 64) / (conv["ms_per_step"] / 1e3) / 1e9,
                              "peak": peaks["hbm"], "unit": "GB/s", "ms_per_step": conv["ms_per_step"],
                              "note": "bf16 numerators -> dZ in place: 2 B read + 2 B written per (token, vocab) entry"}
                             if conv else None),
        "roofline_gather": {"bound": "hbm", "achieved": gather_gbs, "peak": peaks["hbm"], "unit": "GB/s",
                            "frac": gather_gbs / peaks["hbm"],
                            "note": "bulk launch of 2^26 tokens drawn once each from a 134 MB source (> L2), 19 B/token, "
                                    "streaming stores, timed through the C ABI",
                            "staged_variant_gbs": gather_other_gbs, "plain_stores_gbs": gather_plain_gbs},
        "roofline_adamw": ({"bound": "hbm",
                            "achieved": ((32.0 if wire == "fp32" else 28.0) * VOCAB * D_MODEL) / (kern["adamw"]["ms_per_step"] / 1e3) / 1e9,
                            "peak": peaks["hbm"], "unit": "GB/s", "ms_per_step": kern["adamw"]["ms_per_step"],
                            "sqnorm_ms_per_step": kern.get("sqnorm", {}).get("ms_per_step"),
                            "note": "fused AdamW on W inside the timed step: bf16 parameter, fp32 master and moments, "
                                    f"{wire} gradient"} if "adamw" in kern else None),
        "kernels_ms_per_step": {k: round(v["ms_per_step"], 4) for k, v in kern.items()},
        "step_without_optimizer_ms": ms_noopt,
        "recompute_mode": {"ms_per_step": recompute_ms, "value": world * M / (recompute_ms / 1e3), "unit": "tokens/s",
                           "note": "ops.set_backward_mode('recompute'), round 1's default: forward keeps only lse, the "
                                   "backward recomputes the logits (8MVd executed for 6MVd algorithmic); same step otherwise"},
        "e2e": {"value": e2e_value, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms / args.steps,
                "note": "token windows + hidden states from pinned host memory each step, loss read back"},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if world > 1:
        line["rank_kernel_ms_per_step"] = {"min": min(rank_compute), "max": max(rank_compute),
                                           "note": "sum of this library's kernels per step on each rank's GPU "
                                                   "(without the all-reduce and the conversion passes, which run "
                                                   "beside other kernels); the slowest rank paces the step"}
        line["exposed_comm_ms_per_step"] = exposed_comm_ms
        line["allreduce_max_rel_err"] = allreduce_err
        line["full_gradient_allreduce"] = full_grad
    if world == 1 and not args.no_full_step:
        line["full_step"] = run_full_step(dev)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = run_cpu_baseline()
    print(json.dumps(line), file=RESULT_OUT, flush=True)
    if world > 1:
        tol = 1e-4 if wire == "fp32" else 1.5e-2
        dist.destroy_process_group()
        if allreduce_err is None or not (allreduce_err <= tol):
            print(f"all-reduce check failed: max relative error {allreduce_err} > {tol}", file=sys.stderr)
            sys.exit(3)


def run_full_step(dev) -> dict:
    """Context for the headline: one whole finetune step of GPT-Neo-1.3B (BASELINE configs[2]: ctx 2048, batch 8, bf16,
    gradient checkpointing as the reference enables it, torch fused AdamW) on this GPU, with HF's eager blocks + HF's
    lm_head/loss, with this package's fused loss + trunk kernels (patch_gpt_neo), and the latter once more with
    gradient checkpointing off (same arithmetic; 180 GB of HBM hold what the reference's 48 GB cards had to
    recompute).  tokens/s/GPU of the metric's first half.  Errors are reported, not raised: this block must not take the bench line down."""
    import torch
    try:
        from kubernetes_cloud_b200 import trunk
        from kubernetes_cloud_b200.finetune import load_model
        from kubernetes_cloud_b200.loss import compute_loss
        torch.manual_seed(42)
        model = load_model("gpt-neo-1.3b").to(dev).to(torch.bfloat16)
        model.train()
        model.gradient_checkpointing_enable()
        B, T = BATCH, CTX
        g = torch.Generator(device=dev).manual_seed(1)
        ids = torch.randint(0, model.config.vocab_size, (B, T), device=dev, generator=g)
        mask = torch.ones(B, T, dtype=torch.bool, device=dev)
        opt = torch.optim.AdamW(model.parameters(), lr=1e-5, fused=True)

        def step(ours: bool):
            opt.zero_grad(set_to_none=True)
            if ours:
                loss = compute_loss(model, {"input_ids": ids, "attention_mask": mask, "labels": ids.clone()})
            else:
                loss = model(input_ids=ids, attention_mask=mask, labels=ids).loss
            loss.backward()
            opt.step()
            return loss

        out = {"model": "GPT-Neo-1.3B, ctx 2048, batch 8, bf16, gradient checkpointing, torch fused AdamW",
               "tokens_per_step": B * T}
        for name, ours, n in (("hf_eager", False, 2), ("ftk_loss_and_trunk", True, 5),
                              ("ftk_loss_and_trunk_no_recompute", True, 5)):
            if name == "ftk_loss_and_trunk":
                out["modules_rebound"] = trunk.patch_gpt_neo(model)
            if name.endswith("no_recompute"):  # same arithmetic, activations kept in HBM instead of a second forward
                model.gradient_checkpointing_disable()
            torch.cuda.empty_cache()
            torch.cuda.reset_peak_memory_stats(dev)
            for _ in range(2):
                step(ours)
            torch.cuda.synchronize(dev)
            t0 = time.time()
            for _ in range(n):
                loss = step(ours)
            torch.cuda.synchronize(dev)
            dt = (time.time() - t0) / n
            out[name] = {"ms_per_step": dt * 1e3, "tokens_per_s_per_gpu": B * T / dt, "loss": float(loss.detach()),
                         "peak_mem_gb": torch.cuda.max_memory_allocated(dev) / 1e9}
        out["speedup"] = out["hf_eager"]["ms_per_step"] / out["ftk_loss_and_trunk"]["ms_per_step"]
        out["speedup_no_recompute"] = (out["hf_eager"]["ms_per_step"] /
                                       out["ftk_loss_and_trunk_no_recompute"]["ms_per_step"])
        del model, opt
        torch.cuda.empty_cache()
        return out
    except Exception as e:  # noqa: BLE001
        return {"error": f"{type(e).__name__}: {e}"}


def _claim_stdout():
    """Libraries loaded later (NCCL prints its version banner) write to fd 1; the contract is ONE JSON line on stdout.
    Point fd 1 at stderr for everything else and keep a private handle on the real stdout for the result line."""
    real = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    RESULT_OUT = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-full-step", action="store_true", help="skip the whole-model context block (GPT-Neo-1.3B)")
    a = ap.parse_args()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_gpu(a)
