"""CPU: the pure-Python decisions of bench.py -- which workload a line is about, crops, the clock-sample windows."""
import importlib.util
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["bench_under_test"] = mod
    spec.loader.exec_module(mod)
    return mod


def test_workload_rule(monkeypatch):
    b = _bench()
    assert b.pick_workload("c4", 1)[0:2] == ("c4", [])
    assert b.pick_workload("auto", 8)[0:2] == ("c5", [])                 # N > 1: config 5 strong, nothing nested
    monkeypatch.setattr(b, "physical_gpus", lambda: 1)
    key, nested, rule = b.pick_workload("auto", 1)
    assert key == "c3" and nested == ["c1", "c2", "c4", "c5"] and "config 3" in rule
    monkeypatch.setattr(b, "physical_gpus", lambda: 8)
    key, nested, rule = b.pick_workload("auto", 1)                       # N = 1 point of a scaling curve
    assert key == "c5" and nested == ["c3"] and "8 GPUs" in rule
    monkeypatch.setattr(b, "physical_gpus", lambda: 0)                   # nvidia-smi unavailable: the one-GPU rule
    assert b.pick_workload("auto", 1)[0] == "c3"


def test_crops_are_centred_and_clamped():
    b = _bench()
    assert b.crop_rect({}, 3840, 2160, None) == (0, 0, 3840, 2160)
    assert b.crop_rect({}, 3840, 2160, (256, 144)) == (1792, 1008, 256, 144)
    assert b.crop_rect({}, 100, 50, (256, 144)) == (0, 0, 100, 50)


def test_clock_sampler_keeps_only_samples_inside_the_timed_windows():
    b = _bench()
    s = b.ClockSampler(0)
    t0 = time.time()
    row = lambda mhz, reason="Not Active": ["0", str(mhz), "1965", "300.0", "0x0", reason, "Not Active", "Not Active", "Not Active"]
    s.rows = [(t0 + 0.0, row(500)), (t0 + 1.0, row(1900)), (t0 + 1.1, row(1950)), (t0 + 1.2, row(1965)), (t0 + 1.3, row(1965, "Active")),
              (t0 + 3.0, row(400))]
    out = s.stop([(t0 + 0.95, t0 + 1.35)])
    assert out["samples"] == 4 and out["sm_mhz"] == 1957.5 and out["reasons"] == ["hw_slowdown"] and "inside" in out["sampled"]
    out = s.stop([(t0 + 10.0, t0 + 10.1)])                               # region shorter than the sampling period: all samples, and it says so
    assert out["samples"] == 6 and "whole run" in out["sampled"]
