"""Where the end-to-end time of config 2 goes (set_scene / render / read-back)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
rt = load_package()
ps = rt.PackedScene.load(os.path.join(ROOT, "tests/golden/packed/scene1.npz"))
r = rt.Renderer(0); r.set_scene(ps)
W, H = 1920, 1080
fr = rt.make_frame(ps.frame, W, H, max_bounces=2)
out8 = rt.pinned_empty((H, W, 4), np.uint8)
def timeit(fn, n=200):
    for _ in range(10): fn()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    return (time.perf_counter() - t0) / n * 1e3
print("set_scene            ms", timeit(lambda: r.set_scene(ps)))
print("render (no readback) ms", timeit(lambda: r.render(fr, want_f32=False, want_rgba8=False)))
st = [None]
def f():
    st[0] = r.render(fr, want_f32=False, want_rgba8=True, out_rgba8=out8)[2]
print("render + rgba8       ms", timeit(f), "d2h", st[0].ms_d2h, "render", st[0].ms_render, "launches", st[0].kernel_launches, "chunks", st[0].chunks)
pageable = np.empty((H, W, 4), np.uint8)
def g():
    st[0] = r.render(fr, want_f32=False, want_rgba8=True, out_rgba8=pageable)[2]
print("render + rgba8 pageable ms", timeit(g), "d2h", st[0].ms_d2h)
def h():
    r.set_scene(ps); f()
print("e2e                  ms", timeit(h))
import ctypes as C
L = rt.lib(); st1 = rt.abi.Stats(); pfr = C.byref(fr); pst = C.byref(st1); h = r.h
print("raw rtb200_render    ms", timeit(lambda: L.rtb200_render(h, pfr, None, None, pst)), "ms_render", st1.ms_render, "kernel", st1.ms_primary_kernel)
print("raw rtb200_render, no stats ms", timeit(lambda: L.rtb200_render(h, pfr, None, None, None)))
s = ps.c_struct(); psx = C.byref(s)
print("raw rtb200_set_scene ms", timeit(lambda: L.rtb200_set_scene(h, psx)))
print("python c_struct()    ms", timeit(lambda: ps.c_struct()))
