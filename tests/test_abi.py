"""The C-ABI library loads on a machine without a GPU and exports every symbol include/ftk.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ftk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ftk_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for s in ("ftk_version", "ftk_last_error", "ftk_gather_windows", "ftk_lmhead_ce_fwd", "ftk_lmhead_ce_bwd",
              "ftk_lmhead_ce_fwd_workspace", "ftk_lmhead_ce_bwd_workspace", "ftk_dp_init", "ftk_dp_allreduce"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    from kubernetes_cloud_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "build the library first (__graft_entry__.build())"
    l = ctypes.CDLL(_lib.LIB_PATH)
    for s in declared_symbols():
        assert hasattr(l, s), f"{s} declared in ftk.h but not exported"
    assert set(_lib.SIGNATURES) == set(declared_symbols()), "ctypes binding and header disagree"
    assert _lib.lib().ftk_version() == _lib.ABI_VERSION


def test_workspace_queries_are_host_arithmetic():
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    assert L.ftk_lmhead_ce_fwd_workspace(0, 768, 50257) == 0
    a = L.ftk_lmhead_ce_fwd_workspace(2048, 768, 50257)
    b = L.ftk_lmhead_ce_fwd_workspace(16384, 2048, 50257)
    assert 0 < a < b
    # one bf16 dZ slab: rows rounded to 256, pitch rounded to 64 columns
    assert L.ftk_lmhead_ce_bwd_workspace(1000, 768, 50257, 0) == 1024 * 50304 * 2
    assert L.ftk_lmhead_ce_bwd_workspace(100000, 768, 50257, 512) == 512 * 50304 * 2


def test_bad_arguments_return_error_codes_without_a_gpu():
    from kubernetes_cloud_b200 import _lib
    L = _lib.lib()
    rc = L.ftk_gather_windows(None, 1, 8, None, 1, 0, 0, 0, None, None, None, None, None)
    assert rc == -1 and b"null pointer" in L.ftk_last_error()
    with pytest.raises(_lib.FtkError):
        _lib.check(rc, "ftk_gather_windows")


def test_ops_refuse_cpu_tensors():
    import torch
    from kubernetes_cloud_b200 import FtkError, ops
    with pytest.raises(FtkError):
        ops.fused_lmhead_ce(torch.zeros(4, 64, dtype=torch.bfloat16), torch.zeros(8, 64, dtype=torch.bfloat16), None,
                            torch.zeros(4, dtype=torch.int64))
