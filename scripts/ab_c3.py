#!/usr/bin/env python
"""One library variant on config 3 (soup 10 k + 64 quadrics, 4 lights, 4K, depth 3) and on a 20 k-triangle height field
(8 lights, 4K, depth 2) through the culled scan: ms per frame (median of N), stage times, counters and the SHA-1 of the
float frame -- variants must print the same hashes.  The library is chosen by RTB200_LIB (scripts/ab_c3.sh loops).
usage: RTB200_LIB=ab/<name>/librtb200.so python scripts/ab_c3.py [frames]"""
import hashlib
import os
import statistics
import sys
from importlib import import_module

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

rt = load_package()
synth = import_module("rtb200.synth")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 7
tag = os.environ.get("RTB200_LIB", "default").split("/")[-2] if os.environ.get("RTB200_LIB") else "default"
r = rt.Renderer(0)
W, H = 3840, 2160
for name, ps, depth in (("c3", synth.soup(10000, 64, 4, seed=1234), 3), ("hf20k", synth.heightfield(20000, 300.0, 8, seed=9), 2)):
    r.set_scene(ps)
    fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=rt.F_FORCE_STREAM)
    f32, _, st = r.render(fr, want_rgba8=False)
    sha = hashlib.sha1(f32.tobytes()).hexdigest()[:12]
    ms, near, shad, shade = [], [], [], []
    for _ in range(N):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
        ms.append(st.ms_render); near.append(st.ms_trace_nearest); shad.append(st.ms_trace_shadow); shade.append(st.ms_shade)
    med = statistics.median
    print(f"{tag:10s} {name:6s} frame {med(ms):7.3f} ms (min {min(ms):7.3f})  nearest {med(near):6.3f}  shadow {med(shad):6.3f}  shade {med(shade):5.3f} | "
          f"exec {st.tests_triangle_executed / max(st.tests_triangle, 1):.5f} fixups {st.shadow_fixups} sha1 {sha}", flush=True)
r.close()
