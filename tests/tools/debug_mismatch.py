"""Debugging aid (test infrastructure: it uses the oracle as the checker): where a streamed-path frame differs from
the oracle port.  usage: python tests/tools/debug_mismatch.py"""
import os, sys, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_package
from oracle.portbind import PortOracle
rt = load_package(); port = PortOracle(rt.abi); r = rt.Renderer(0)
for n in range(1, 10):
    ps = rt.PackedScene.load(os.path.join(ROOT, "tests/golden/packed", f"scene{n}.npz")); r.set_scene(ps)
    fr = rt.make_frame(ps.frame, 640, 480, max_bounces=4)
    got, got8, st = r.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads); want8, _ = port.quantise(want)
    sig, hit = port.signatures(ps, fr, threads=port.max_threads)
    d = np.abs(got8[..., :3].astype(int) - want8[..., :3].astype(int)).max(-1)
    ys, xs = np.nonzero(d > 1)
    print(f"scene{n}: bad={len(ys)}")
    for y, x in zip(ys, xs):
        for rad in (1, 2, 3, 5):
            blk = sig[max(0, y - rad):y + rad + 1, max(0, x - rad):x + rad + 1]
            if (blk != sig[y, x]).any():
                break
        else:
            rad = 99
        if rad > 1:
            print(f"   ({x},{y}) edge-radius={rad} hit={hit[y, x]} got={got[y, x]} want={want[y, x]}")
