"""Steady-state cost of rtb200_set_scene + rtb200_build_bvh on the config-4 mesh (1 M triangles)."""
import os, sys, time
from importlib import import_module
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
rt = load_package(); synth = import_module("rtb200.synth")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
ps = synth.heightfield(n, 1000.0, 8, seed=4321)
r = rt.Renderer(0)
for i in range(4):
    t0 = time.perf_counter(); r.set_scene(ps); t1 = time.perf_counter(); info = r.build_bvh(); t2 = time.perf_counter()
    print(f"set_scene {1e3*(t1-t0):.3f} ms   build_bvh call {1e3*(t2-t1):.3f} ms   device build {info['build_ms']:.3f} ms", flush=True)
