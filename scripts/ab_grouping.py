#!/usr/bin/env python
"""A/B of the queue order of the tiled path (RTB200_QUERY_GROUPING, read per frame): level-0 rays by 16x16 patches
and each CTA's shadow queries light by light -- and of RTB200_NEAREST_CONES=4 (warps of secondary rays cull with up to
4 cones).  Frames must be bit-identical (only the ORDER of the queues / the set of culled triangles changes);
prints frame / stage times and the share of (ray, triangle) tests executed.
usage: python scripts/ab_grouping.py [quick]"""
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
    r = rt.Renderer(0)
    cases = [("c3 soup 10k+64q, 4 lights, d3, 3840x2160", synth.soup(10000, 64, 4, seed=1234), 3840, 2160, 3, 1),
             ("soup 600+24q, 3 lights, d3, 251x131 spp 2x2", synth.soup(600, 24, 3, seed=2024, radius=45.0), 251, 131, 3, 2),
             ("height field 20k, 8 lights, d2, 1000x562", synth.heightfield(20000, 400.0, 8, seed=3), 1000, 562, 2, 1)]
    if len(sys.argv) > 1 and sys.argv[1] == "quick":
        cases = cases[:2]
    for name, ps, W, H, depth, spp in cases:
        r.set_scene(ps)
        fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp, flags=rt.F_FORCE_STREAM)
        res = {}
        for tag, env, cones in (("ray order", "0", "1"), ("grouped", "1", "1"), ("grouped, 4 cones", "1", "4")):
            os.environ["RTB200_QUERY_GROUPING"] = env
            os.environ["RTB200_NEAREST_CONES"] = cones      # secondary rays: up to 4 cones per warp (k_trace_nearest MULTI)
            r.render(fr, want_f32=False, want_rgba8=False)
            img, img8, st = r.render(fr)
            res[tag] = (img, img8, st)
            print(f"{name} [{tag}]: {st.ms_render:.2f} ms (nearest {st.ms_trace_nearest:.2f}, shadow {st.ms_trace_shadow:.2f}, shade {st.ms_shade:.2f}), "
                  f"rays p/s/rl/rr {st.rays_primary}/{st.rays_shadow}/{st.rays_reflect}/{st.rays_refract}, "
                  f"tests executed {st.tests_triangle_executed / max(st.tests_triangle, 1):.4f} of {st.tests_triangle:.3e}", flush=True)
        os.environ.pop("RTB200_QUERY_GROUPING"); os.environ.pop("RTB200_NEAREST_CONES")
        a = res["ray order"]
        for tag in ("grouped", "grouped, 4 cones"):
            b = res[tag]
            print(f"   ray order vs {tag}: differing floats =", int((a[0].view(np.uint32) != b[0].view(np.uint32)).sum()),
                  "| differing bytes =", int((a[1] != b[1]).sum()),
                  "| ray counts equal =", (a[2].rays_shadow, a[2].rays_reflect, a[2].rays_refract) == (b[2].rays_shadow, b[2].rays_reflect, b[2].rays_refract), flush=True)
    r.close()


if __name__ == "__main__":
    main()
