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
    assert C.sizeof(rt.abi.Stats) == 8 * 4 + 8 * 9 + 4 * 4 + 8 * 3 + 8 * 3 + 8 + 8 + 8 * 2


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


def test_validate_scene_without_a_gpu(rt):
    """rtb200_validate_scene = the checks of rtb200_set_scene, pure host code: the order arrays must
    each be strictly increasing and together a permutation of 0..N-1 (insertion indices), materials
    must exist, counts must be sane."""
    from importlib import import_module
    import numpy as np
    synth = import_module("rtb200.synth")
    base = synth.soup(40, 3, 1, seed=2)
    ok, why = rt.validate_scene(base)
    assert ok and why == ""
    for n in range(1, 10):                                   # every bundled scene passes
        ps = rt.PackedScene.load(f"{__import__('os').path.dirname(__file__)}/golden/packed/scene{n}.npz")
        assert rt.validate_scene(ps)[0]
    T, Q, P, _ = base.counts
    rng = np.random.default_rng(0)
    slots = rng.permutation(T + Q + P)                       # interleaved insertion order is fine
    arrays = {k: base[k] for k in base.a}
    arrays["tri_order"] = np.sort(slots[:T]).astype(np.int32)
    arrays["quad_order"] = np.sort(slots[T:T + Q]).astype(np.int32)
    arrays["plane_order"] = np.sort(slots[T + Q:]).astype(np.int32)
    assert rt.validate_scene(rt.PackedScene(arrays, base.frame))[0]
    cases = {
        "duplicate order": ("tri_order", lambda a: np.concatenate([a[:1], a[:-1]])),
        "decreasing order": ("tri_order", lambda a: a[::-1].copy()),
        "order out of range": ("plane_order", lambda a: a + 7),
        "negative order": ("quad_order", lambda a: np.where(np.arange(len(a)) == 0, -1, a)),
        "two types claim one slot": ("quad_order", lambda a: np.concatenate([[0], a[1:]])),
        "material out of range": ("tri_material", lambda a: np.where(np.arange(len(a)) == 5, 10 ** 6, a)),
        "negative material": ("plane_material", lambda a: a * 0 - 1),
    }
    for name, (key, bad) in cases.items():
        arrays = {k: base[k].copy() for k in base.a}
        arrays[key] = np.asarray(bad(arrays[key])).astype(np.int32)
        ok, why = rt.validate_scene(rt.PackedScene(arrays, base.frame))
        assert not ok and "invalid" in why, name
    empty = rt.PackedScene({"materials": np.zeros((0, 10)), "lights": np.zeros((0, 7))}, base.frame)
    assert rt.validate_scene(empty)[0]                       # an empty scene renders the background
