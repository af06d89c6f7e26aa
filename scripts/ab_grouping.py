#!/usr/bin/env python
"""A/B of the switches of the streamed path, inside one process (all read per frame): the default culled scan, the
same without the 16x16 patches / light-by-light queues (RTB200_QUERY_GROUPING=0), without the Morton sort of the
levels >= 1 (RTB200_SORT_SECONDARY=0), and the plain tiled scan (RTB200_NO_BUNDLE_CULL=1).  Frames must be bit-identical
(only the ORDER of the work / the set of culled triangles changes); prints frame / stage times and the share of
(ray, triangle) tests entered.
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
        variants = (("default", {}), ("scanline queues", {"RTB200_QUERY_GROUPING": "0"}), ("levels >= 1 unsorted", {"RTB200_SORT_SECONDARY": "0"}),
                    ("plain scan", {"RTB200_NO_BUNDLE_CULL": "1"}))
        for tag, env in variants:
            os.environ.update(env)
            r.render(fr, want_f32=False, want_rgba8=False)
            img, img8, st = r.render(fr)
            for k in env:
                os.environ.pop(k)
            res[tag] = (img, img8, st)
            print(f"{name} [{tag}]: {st.ms_render:.2f} ms (nearest {st.ms_trace_nearest:.2f}, shadow {st.ms_trace_shadow:.2f}, shade {st.ms_shade:.2f}), "
                  f"rays p/s/rl/rr {st.rays_primary}/{st.rays_shadow}/{st.rays_reflect}/{st.rays_refract}, "
                  f"tests entered {st.tests_triangle_executed / max(st.tests_triangle, 1):.4f} of {st.tests_triangle:.3e}, fix-ups {st.shadow_fixups}", flush=True)
        a = res["default"]
        for tag, _ in variants[1:]:
            b = res[tag]
            print(f"   default vs {tag}: differing floats =", int((a[0].view(np.uint32) != b[0].view(np.uint32)).sum()),
                  "| differing bytes =", int((a[1] != b[1]).sum()),
                  "| ray counts equal =", (a[2].rays_shadow, a[2].rays_reflect, a[2].rays_refract) == (b[2].rays_shadow, b[2].rays_reflect, b[2].rays_refract), flush=True)
    r.close()


if __name__ == "__main__":
    main()
