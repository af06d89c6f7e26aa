"""CPU: the C-ABI library loads and exports every symbol include/rtb200.h declares; compute
entry points fail loudly (no fallback) when no CUDA device is usable."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def test_header_symbols_exported(rt):
    hdr = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    declared = sorted(set(re.findall(r"\b(rtb200_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 15
    L = rt.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert sorted(rt.EXPORTS) == declared
    assert L.rtb200_abi_version() == 1


def test_struct_sizes_match_header(rt):
    assert C.sizeof(rt.abi.Material) == 40 and C.sizeof(rt.abi.Light) == 28
    assert C.sizeof(rt.abi.Frame) == 4 * 4 + 8 + 12 * 5 + 8 + 4 + 4
    assert C.sizeof(rt.abi.Stats) == 8 * 4 + 8 * 9 + 4 * 4 + 8 * 3 + 8 * 3 + 8


def test_no_cpu_fallback(rt):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(rt.RtError):
        rt.Renderer(0)


def test_error_strings(rt):
    L = rt.lib()
    assert L.rtb200_error_string(0) == b"ok"
    assert b"CUDA" in L.rtb200_error_string(-2)
