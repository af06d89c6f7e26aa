"""Timing probe of the streamed path on a synthetic soup (BASELINE config 3 shape)."""
import argparse, os, sys, json
from importlib import import_module
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
ap = argparse.ArgumentParser()
ap.add_argument("--tris", type=int, default=10000); ap.add_argument("--quads", type=int, default=64)
ap.add_argument("--lights", type=int, default=4); ap.add_argument("-x", type=int, default=3840); ap.add_argument("-y", type=int, default=2160)
ap.add_argument("--depth", type=int, default=3); ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--fused", action="store_true"); ap.add_argument("--exact", action="store_true")
a = ap.parse_args()
rt = load_package(); synth = import_module("rtb200.synth")
ps = synth.soup(a.tris, a.quads, a.lights)
r = rt.Renderer(0); r.set_scene(ps)
peak, mhz = r.fp32_peak()
flags = rt.F_FORCE_STREAM | (rt.F_FUSED_FP if a.fused else 0) | (rt.F_EXACT_FP if a.exact else 0)
fr = rt.make_frame(ps.frame, a.x, a.y, max_bounces=a.depth, flags=flags)
for i in range(a.frames):
    _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
    flop = st.tests_triangle * 46.0 + st.tests_quadric * 65.0 + st.tests_plane * 12.0
    prim_flop = st.rays_primary * (a.tris * 46.0 + a.quads * 65.0 + 12.0)
    print(json.dumps({"frame": i, "ms": st.ms_render, "Mrays": st.rays / 1e6, "Mrays_per_s": st.rays / st.ms_render / 1e3,
                      "rays": [st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract], "levels": st.depth_levels,
                      "launches": st.kernel_launches, "TFLOPs_alg": flop / st.ms_render / 1e9, "frac_of_peak": flop / st.ms_render / 1e9 / peak,
                      "k_nearest_l0_ms": st.ms_primary_kernel, "k_nearest_l0_TFLOPs": prim_flop / st.ms_primary_kernel / 1e9,
                      "k_nearest_l0_frac": prim_flop / st.ms_primary_kernel / 1e9 / peak, "peak": peak}))
