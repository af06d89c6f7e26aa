#!/usr/bin/env python
"""Where does the default mode (culling filters) differ from RTB200_F_EXACT_FP on config 3 at 4K?"""
import os, sys
from importlib import import_module
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
from oracle.portbind import PortOracle

rt = load_package(); synth = import_module("rtb200.synth")
port = PortOracle(rt.abi)
r = rt.Renderer(0)
ps = synth.soup(10000, 64, 4, seed=1234)
r.set_scene(ps)
W, H = 3840, 2160
for depth in (3, 0):
    fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=rt.F_FORCE_STREAM)
    fx = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=rt.F_FORCE_STREAM | rt.F_EXACT_FP)
    a, _, sa = r.render(fr); b, _, sb = r.render(fx)
    bad = np.argwhere((a.view(np.uint32) != b.view(np.uint32)).any(axis=2))
    print("depth", depth, "differing pixels:", bad.tolist(), "rays", (sa.rays_shadow, sa.rays_reflect, sa.rays_refract), (sb.rays_shadow, sb.rays_reflect, sb.rays_refract))
    for y, x in bad[:4]:
        want, _ = port.render_region(ps, fr, int(x), int(y), 1, 1, threads=1)
        print("  pixel", (x, y), "default", a[y, x], "exact", b[y, x], "oracle", want[0, 0])
        sig, hit = port.signatures(ps, rt.make_frame(ps.frame, W, H, max_bounces=depth), threads=16) if False else (None, None)
r.close()
