#!/usr/bin/env python
"""One library variant on config 5 (100 k-triangle height field, 4 lights, 8192 x 8192 at 16 spp, depth 2, LBVH) and
config 4 (1 M triangles, 8 lights, 4K, depth 4): ms per frame, stage times, SHA-1 of the RGBA8 frame.
usage: RTB200_LIB=ab/<name>/librtb200.so python scripts/ab_c5.py [frames]"""
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
N = int(sys.argv[1]) if len(sys.argv) > 1 else 3
tag = os.environ.get("RTB200_LIB", "default").split("/")[-2] if os.environ.get("RTB200_LIB") else "default"
r = rt.Renderer(0)
for name, ps, W, H, depth, spp in (("c5", synth.heightfield(100000, 350.0, 4, seed=5), 8192, 8192, 2, 4),
                                   ("c4", synth.heightfield(1000000, 1000.0, 8, seed=4321), 3840, 2160, 4, 1)):
    r.set_scene(ps)
    r.build_bvh()
    fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp, flags=rt.F_USE_BVH)
    _, rgba, st = r.render(fr, want_f32=False)
    sha = hashlib.sha1(rgba.tobytes()).hexdigest()[:12]
    ms, near, shad, shade = [], [], [], []
    for _ in range(N):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
        ms.append(st.ms_render); near.append(st.ms_trace_nearest); shad.append(st.ms_trace_shadow); shade.append(st.ms_shade)
    med = statistics.median
    print(f"{tag:10s} {name:4s} frame {med(ms):8.3f} ms (min {min(ms):8.3f})  nearest {med(near):7.3f}  shadow {med(shad):7.3f}  shade {med(shade):6.3f} | "
          f"chunks {st.chunks} sha1 {sha}", flush=True)
r.close()
