#!/usr/bin/env python
"""A/B of the warp-bundle culling of the tiled brute-force kernels on BASELINE config 3 (and a height
field through the same kernels): frame time with / without (RTB200_NO_BUNDLE_CULL is read per launch),
and the frame compared bit for bit with RTB200_F_EXACT_FP (no culling of any kind).
usage: python scripts/ab_bundle.py [W H]"""
import os
import sys
from importlib import import_module

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402


def main():
    rt = load_package()
    synth = import_module("rtb200.synth")
    W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
    r = rt.Renderer(0)
    scenes = {"c3 soup 10k+64q, 4 lights, d3": (synth.soup(10000, 64, 4, seed=1234), 3),
              "height field 20k, 4 lights, d2": (synth.heightfield(20000, 400.0, 4, seed=3), 2)}
    for name, (ps, depth) in scenes.items():
        r.set_scene(ps)
        fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=rt.F_FORCE_STREAM)
        res = {}
        for tag, env in (("cull", "0"), ("nocull", "1")):
            os.environ["RTB200_NO_BUNDLE_CULL"] = env
            r.render(fr, want_f32=False, want_rgba8=False)
            img, _, st = r.render(fr)
            res[tag] = (img, st)
            print(f"{name} [{tag}]: {st.ms_render:.2f} ms (nearest {st.ms_trace_nearest:.2f}, shadow {st.ms_trace_shadow:.2f}, shade {st.ms_shade:.2f}), "
                  f"rays {st.rays_primary + st.rays_shadow + st.rays_reflect + st.rays_refract}, tests executed {st.tests_triangle_executed / max(st.tests_triangle, 1):.4f} of {st.tests_triangle:.3e}", flush=True)
        os.environ["RTB200_NO_BUNDLE_CULL"] = "0"
        a, b = res["cull"][0], res["nocull"][0]
        print("   cull vs nocull: differing floats =", int((a.view(np.uint32) != b.view(np.uint32)).sum()), flush=True)
        fx = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=rt.F_FORCE_STREAM | rt.F_EXACT_FP)
        ex, _, stx = r.render(fx)
        print(f"   cull vs EXACT_FP ({stx.ms_render:.1f} ms): differing floats =", int((a.view(np.uint32) != ex.view(np.uint32)).sum()),
              "| shadow rays", res["cull"][1].rays_shadow, stx.rays_shadow, flush=True)
    r.close()


if __name__ == "__main__":
    main()
