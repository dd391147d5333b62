"""ctypes binding of the C ABI declared in include/ftk.h (lib/libftk.so).

There is no fallback: if the library is missing or its ABI version differs, importing an op raises.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import POINTER, c_char_p, c_float, c_int32, c_int64, c_size_t, c_void_p

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libftk.so")
CSRC = os.path.join(_PKG, "csrc")
ABI_VERSION = 1

FTK_BF16, FTK_F32 = 0, 1

# name -> (restype, argtypes); mirrors include/ftk.h one to one
SIGNATURES = {
    "ftk_version": (c_int32, []),
    "ftk_last_error": (c_char_p, []),
    "ftk_gather_windows": (c_int32, [c_void_p, c_int64, c_int32, c_void_p, c_int32, c_int32, c_int32, c_int32,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ftk_lmhead_ce_fwd_workspace": (c_size_t, [c_int64, c_int32, c_int32]),
    "ftk_lmhead_ce_fwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "ftk_lmhead_ce_probs_bytes": (c_size_t, [c_int64, c_int32]),
    "ftk_lmhead_ce_cref_bytes": (c_size_t, [c_int64, c_int32]),
    "ftk_lmhead_ce_fwd_keep": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p,
                                         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "ftk_lmhead_ce_bwd_kept": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32,
                                         c_int32, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32,
                                         c_void_p, c_void_p]),
    "ftk_lmhead_ce_fwd_static": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p,
                                           c_void_p, c_void_p, c_void_p]),
    "ftk_lmhead_ce_fused_slab_rows": (c_int64, [c_int64, c_int32, c_int32, c_int64]),
    "ftk_lmhead_ce_fused_plan": (c_int64, [c_int64, c_int32, c_int32, c_int32, c_int32]),
    "ftk_lmhead_ce_fused_workspace": (c_size_t, [c_int64, c_int32, c_int32, c_int64]),
    "ftk_lmhead_ce_fused": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_float,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32,
                                      c_void_p, c_void_p, c_void_p, c_size_t, c_int64, c_void_p]),
    "ftk_scale_unless_one": (c_int32, [c_void_p, c_size_t, c_int32, c_void_p, c_int32, c_void_p]),
    "ftk_lmhead_ce_bwd_workspace": (c_size_t, [c_int64, c_int32, c_int32, c_int64]),
    "ftk_lmhead_ce_bwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32,
                                    c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_void_p,
                                    c_void_p, c_size_t, c_void_p]),
    "ftk_set_option": (c_int32, [c_int32, c_int32]),
    "ftk_timeline_enable": (c_int32, [c_int32]),
    "ftk_timeline_read": (c_int32, [c_void_p, c_void_p, c_void_p, c_int32]),
    "ftk_linear_fwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_int32, c_void_p,
                                 c_void_p, c_void_p]),
    "ftk_linear_bwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p,
                                 c_void_p, c_int32, c_int32, c_void_p, c_void_p]),
    "ftk_layernorm_fwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_float, c_void_p, c_void_p, c_void_p,
                                    c_void_p]),
    "ftk_layernorm_bwd_workspace": (c_size_t, [c_int64, c_int32]),
    "ftk_layernorm_bwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "ftk_attention_fwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32,
                                    c_float, c_void_p, c_void_p, c_void_p]),
    "ftk_attention_bwd_workspace": (c_size_t, [c_int32, c_int32, c_int32]),
    "ftk_attention_bwd": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32,
                                    c_int32, c_int32, c_int32, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                    c_void_p]),
    "ftk_set_sm_reserve": (c_int32, [c_int32]),
    "ftk_dp_unique_id": (c_int32, [c_void_p]),
    "ftk_dp_init": (c_int32, [c_void_p, c_int32, c_int32, POINTER(c_void_p)]),
    "ftk_dp_allreduce": (c_int32, [c_void_p, c_void_p, c_size_t, c_int32, c_float, c_void_p]),
    "ftk_dp_allreduce_after": (c_int32, [c_void_p, c_void_p, c_size_t, c_int32, c_float, c_void_p]),
    "ftk_dp_join": (c_int32, [c_void_p, c_void_p]),
    "ftk_dp_destroy": (c_int32, [c_void_p]),
    "ftk_grad_sqnorm_workspace": (c_size_t, []),
    "ftk_grad_sqnorm": (c_int32, [c_void_p, c_size_t, c_int32, c_void_p, c_int32, c_void_p, c_size_t, c_void_p]),
    "ftk_clip_coef": (c_int32, [c_void_p, c_float, c_void_p, c_void_p]),
    "ftk_adamw_step": (c_int32, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int32, c_int32, c_float,
                                 c_float, c_float, c_float, c_float, c_int64, c_float, c_void_p, c_void_p]),
}


OPTIONS = {"cta_group": 0, "gather_variant": 1, "gather_stores": 2, "sm_reserve": 3}  # FTK_OPT_* in include/ftk.h


class FtkError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into lib/libftk.so (nvcc cross-compiles without a GPU)."""
    res = subprocess.run(["make", "-C", CSRC, "-j4"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise FtkError("building libftk.so failed")
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FtkError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU or eager fallback)")
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        v = l.ftk_version()
        if v != ABI_VERSION:
            raise FtkError(f"libftk.so ABI version {v}, binding expects {ABI_VERSION}: rebuild the library")
        _lib = l
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().ftk_last_error()
        raise FtkError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")


def set_option(name: str, value: int) -> None:
    """ftk_set_option by name: "cta_group" (1|2), "gather_variant" (0 auto|1 direct|2 staged), "gather_stores"
    (0 auto|1 streaming|2 plain), "sm_reserve" (0..64)."""
    check(lib().ftk_set_option(OPTIONS[name], int(value)), f"ftk_set_option({name})")


TIMELINE_TAGS = {1: "gather", 2: "fwd", 3: "fwd_keep", 4: "finalize", 5: "dz", 6: "convert", 7: "dH", 8: "dW",
                 9: "dbias", 10: "adamw", 11: "sqnorm", 12: "allreduce",
                 13: "linear_fwd", 14: "linear_dx", 15: "attn_fwd", 16: "attn_bwd", 17: "layernorm_fwd",
                 18: "layernorm_bwd"}  # FTK_TL_* in include/ftk.h


def timeline_enable(on: bool) -> None:
    check(lib().ftk_timeline_enable(int(on)), "ftk_timeline_enable")


def timeline_read(max_n: int = 1 << 16, with_start: bool = False):
    """[(kernel name, milliseconds)] for every tagged launch since ``timeline_enable(True)``; waits for them.
    ``with_start``: triples (name, milliseconds, start in ms after the first launch) -- a small trace across streams."""
    tags = (c_int32 * max_n)()
    ms = (c_float * max_n)()
    st = (c_float * max_n)()
    n = lib().ftk_timeline_read(tags, ms, st if with_start else None, max_n)
    if n < 0:
        check(n, "ftk_timeline_read")
    if with_start:
        return [(TIMELINE_TAGS.get(tags[i], str(tags[i])), ms[i], st[i]) for i in range(n)]
    return [(TIMELINE_TAGS.get(tags[i], str(tags[i])), ms[i]) for i in range(n)]
