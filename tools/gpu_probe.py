#!/usr/bin/env python
"""Stage-by-stage GPU bring-up probe (development tool, not a test): each stage runs in its own subprocess with a
timeout so a device trap or hang in one stage does not hide the others.  Usage: python tools/gpu_probe.py [stage...]"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def stage_gather():
    import numpy as np
    import torch
    from oracle import oracle as O
    from kubernetes_cloud_b200 import ops  # noqa
    res = {}
    for T, n, B in ((128, 7, 7), (2048, 9, 8), (37, 5, 3), (1024, 40, 16)):
        tok = O.synth_tokens(n, T, 50257, 1234, trailing_pads=min(300, T // 2))
        idx = np.random.default_rng(0).integers(0, n, size=B)
        idx[0] = n - 1
        for pad_is_eos in (True, False):
            for shift in (False, True):
                want = O.gather_c(tok, T, idx, 50256, pad_is_eos, shift)
                t = torch.from_numpy(tok.view(np.int16)).cuda()
                got = torch.ops.ftk.gather_windows(t, torch.from_numpy(idx).cuda(), T, 50256, pad_is_eos, shift)
                torch.cuda.synchronize()
                ok = all(np.array_equal(w, g.cpu().numpy()) for w, g in zip(want, got))
                res[f"T{T}_B{B}_eos{int(pad_is_eos)}_shift{int(shift)}"] = ok
    return res


def _mk(M, d, V, seed=0, bias=False, frac_ignore=0.1):
    import torch
    g = torch.Generator().manual_seed(seed)
    H = torch.randn(M, d, generator=g).to(torch.bfloat16)
    W = (torch.randn(V, d, generator=g) * 0.05).to(torch.bfloat16)
    b = (torch.randn(V, generator=g) * 0.5).to(torch.bfloat16) if bias else None
    t = torch.randint(0, V, (M,), generator=g)
    t[torch.rand(M, generator=g) < frac_ignore] = -100
    return H, W, b, t


def stage_fwd(shapes):
    import torch
    from oracle import oracle as O
    from kubernetes_cloud_b200 import ops  # noqa
    res = {}
    for (M, d, V, bias) in shapes:
        H, W, b, t = _mk(M, d, V, bias=bias)
        want = O.lmhead_ce_reference(H, W, b, t, want_grads=False)
        loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(H.cuda(), W.cuda(), None if b is None else b.cuda(), t.cuda())
        torch.cuda.synchronize()
        loss = float(loss_sum) / max(int(n_valid), 1)
        res[f"M{M}_d{d}_V{V}_b{int(bias)}"] = {
            "loss": loss, "want": float(want["loss"]), "n_valid": int(n_valid), "want_n": want["n_valid"],
            "lse_maxabs": float((lse.cpu() - want["lse"]).abs().max()),
            "ok": abs(loss - float(want["loss"])) <= 1e-3 * abs(float(want["loss"])) and int(n_valid) == want["n_valid"]}
    return res


def stage_bwd(shapes):
    import torch
    from oracle import oracle as O
    from kubernetes_cloud_b200 import ops
    res = {}
    for (M, d, V, bias, slab) in shapes:
        ops.set_bwd_slab_rows(slab)
        H, W, b, t = _mk(M, d, V, bias=bias)
        want = O.lmhead_ce_reference(H, W, b, t, want_grads=True)
        Hc, Wc, tc = H.cuda(), W.cuda(), t.cuda()
        bc = None if b is None else b.cuda()
        loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(Hc, Wc, bc, tc)
        for dw32 in (False, True):
            dH, dW, db = torch.ops.ftk.lmhead_ce_bwd(Hc, Wc, bc, tc, lse, None, n_valid, 1.0, dw32)
            torch.cuda.synchronize()
            r = {"dH": O.rel_err(dH.float(), want["dH"]), "dW": O.rel_err(dW.float(), want["dW"])}
            if b is not None:
                r["db"] = O.rel_err(db, want["dbias"])
            r["ok"] = all(v < 1e-2 for v in r.values())
            res[f"M{M}_d{d}_V{V}_b{int(bias)}_slab{slab}_f32{int(dw32)}"] = r
    ops.set_bwd_slab_rows(0)
    return res


def stage_time():
    import torch
    from kubernetes_cloud_b200 import ops  # noqa
    res = {}
    for (M, d, V) in ((16384, 768, 50257), (16384, 2048, 50257)):
        H, W, b, t = _mk(M, d, V)
        Hc, Wc, tc = H.cuda(), W.cuda(), t.cuda()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        for _ in range(3):
            loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(Hc, Wc, None, tc)
            torch.ops.ftk.lmhead_ce_bwd(Hc, Wc, None, tc, lse, None, n_valid, 1.0, False)
        torch.cuda.synchronize()
        n = 5
        ev[0].record()
        for _ in range(n):
            loss_sum, n_valid, lse = torch.ops.ftk.lmhead_ce_fwd(Hc, Wc, None, tc)
        ev[1].record()
        for _ in range(n):
            torch.ops.ftk.lmhead_ce_bwd(Hc, Wc, None, tc, lse, None, n_valid, 1.0, False)
        ev[2].record()
        torch.cuda.synchronize()
        tf, tb = ev[0].elapsed_time(ev[1]) / n, ev[1].elapsed_time(ev[2]) / n
        fl = 2.0 * M * V * d
        res[f"M{M}_d{d}"] = {"fwd_ms": tf, "bwd_ms": tb, "fwd_TF": fl / tf / 1e9, "bwd_TF_alg": 2 * fl / tb / 1e9,
                             "total_TF_alg": 3 * fl / (tf + tb) / 1e9}
    return res


SMALL_F = [(128, 64, 256, False), (256, 128, 1000, False), (300, 64, 777, True), (2048, 768, 50257, False)]
SMALL_B = [(128, 64, 256, False, 0), (256, 128, 1000, False, 0), (300, 64, 777, True, 0), (1024, 256, 5000, False, 256),
           (2048, 768, 50257, False, 0)]
def stage_san():
    """Small shapes for compute-sanitizer runs."""
    r = {}
    r.update(stage_fwd([(256, 128, 1000, False), (300, 64, 777, True)]))
    r.update(stage_bwd([(256, 128, 1000, False, 0), (300, 64, 777, True, 0), (600, 64, 300, False, 256)]))
    g = stage_gather()
    r["gather_all_ok"] = all(g.values())
    return r


STAGES = {
    "san": lambda: stage_san(),
    "gather": lambda: stage_gather(),
    "fwd": lambda: stage_fwd(SMALL_F),
    "bwd": lambda: stage_bwd(SMALL_B),
    "time": lambda: stage_time(),
}

if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--run":
        if os.environ.get("FTK_PROBE_CTA_GROUP"):
            from kubernetes_cloud_b200 import _lib
            _lib.set_option("cta_group", int(os.environ["FTK_PROBE_CTA_GROUP"]))
        out = STAGES[sys.argv[2]]()
        print("RESULT " + json.dumps(out))
        sys.exit(0)
    names = sys.argv[1:] or ["gather", "fwd@1", "bwd@1", "fwd@2", "bwd@2", "time@1", "time@2"]
    for nm in names:
        st, _, cg = nm.partition("@")
        env = dict(os.environ)
        if cg:
            env["FTK_PROBE_CTA_GROUP"] = cg  # the child applies it with _lib.set_option("cta_group", ...)
        t0 = time.time()
        try:
            p = subprocess.run([sys.executable, __file__, "--run", st], env=env, capture_output=True, text=True, timeout=240)
            tail = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
            print(f"=== {nm} rc={p.returncode} {time.time() - t0:.1f}s")
            if tail:
                for k, v in json.loads(tail[-1][7:]).items():
                    print(f"   {k}: {v}")
            if p.returncode != 0 or not tail:
                print("   stdout:", p.stdout[-1500:])
                print("   stderr:", p.stderr[-3000:])
        except subprocess.TimeoutExpired as e:
            print(f"=== {nm} TIMEOUT; stdout={(e.stdout or b'')[-800:]} stderr={(e.stderr or b'')[-800:]}")
        sys.stdout.flush()
