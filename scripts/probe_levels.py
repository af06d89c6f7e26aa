#!/usr/bin/env python
"""Config 3 rendered with max_bounces = 0..3: stage times and executed (ray, triangle) tests per depth
(differences between consecutive depths = the cost of one wavefront level)."""
import os
import sys
from importlib import import_module

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

rt = load_package()
synth = import_module("rtb200.synth")
ps = synth.soup(10000, 64, 4, seed=1234)
r = rt.Renderer(0)
r.set_scene(ps)
W, H = 3840, 2160
for depth in (0, 1, 2, 3):
    fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=rt.F_FORCE_STREAM)
    r.render(fr, want_f32=False, want_rgba8=False)
    _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
    print(f"depth {depth}: frame {st.ms_render:8.3f} ms  nearest {st.ms_trace_nearest:8.3f}  shadow {st.ms_trace_shadow:8.3f}  shade {st.ms_shade:6.3f} | "
          f"rays p/s/rl/rr {st.rays_primary}/{st.rays_shadow}/{st.rays_reflect}/{st.rays_refract} | exec {st.tests_triangle_executed:.4g} "
          f"= {st.tests_triangle_executed / max(st.tests_triangle, 1):.4f} of the scan, fixups {st.shadow_fixups}", flush=True)
