"""Generates tests/golden/synth/ from the reference itself.  Run ONLY where /root/reference exists
(the build container):   python tests/golden/make_golden_synth.py

The bundled scenes (make_golden.py) hold at most 3 quadrics, 3 planes and 2 triangles.  The
benchmark configurations 3-5 of BASELINE.json are SYNTHETIC scenes (triangle soups with
ellipsoids, height fields), so the chain  reference -> oracle port -> CUDA path  needs pins on
scenes of that shape too.  Every scene below is produced by the seeded generators of
ray-tracing-gpu_b200/synth.py (the ones bench.py uses, smaller counts), fed to the compiled
reference (oracle/_ref, recursion restored) through its PUBLIC API -- CScene::AjouterSurface /
AjouterLumiere / AjusterPositionCamera ... (Scene.h:41-60), the way SURVEY.md 8c prescribes -- and
rendered by CScene::LancerRayons() (CPU branch).  Per scene:
  synth/<name>.npz   the INPUT arrays (layout of include/rtb200.h, so the fixture does not depend
                     on numpy's generator staying bit-stable), the camera block of the HOST
                     classes (the reference derives its own from position / look-at / up, so an
                     equal frame pins InitialiserCamera too), and the reference's OUTPUTS:
     rgb             float RGB frame W x H, depth D (row 0 = bottom)
     lattice         float RGB of a w x h region of the 2x2 sub-pixel lattice (the 4-spp samples
                     through the reference's per-ray entry CScene::ObtenirCouleur)
     idx, t, nrm     nearest hit of 4096 seeded rays (insertion order, distance, normal)
     filt            shadow filter of 4096 seeded (point, light) pairs
"""
import os
import sys
from importlib import import_module

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from __graft_entry__ import load_package  # noqa: E402
from oracle.refbind import RefRenderer  # noqa: E402

ARRAYS = ("tri_points", "tri_normal", "tri_material", "tri_order", "quad_coef", "quad_material", "quad_order",
          "plane_coef", "plane_material", "plane_order", "materials", "lights")

# name -> (generator call, W, H, depth)
SCENES = {
    # config-3 shape: soup + ellipsoids (20 % mirrors, 10 % glass, specular on everything) + ground plane
    "soup": (lambda s: s.soup(600, 24, 3, seed=2024, radius=45.0), 160, 96, 3),
    # config-4/5 shape: height field (10 % mirror triangles) + ground plane, 4 lights
    "heightfield": (lambda s: s.heightfield(5000, 300.0, 4, seed=11), 128, 96, 4),
    # quadric-heavy: many overlapping transparent / reflecting ellipsoids, few triangles
    "ellipsoids": (lambda s: s.soup(40, 96, 2, seed=99, radius=60.0), 128, 96, 4),
}


def main():
    rt = load_package()
    synth = import_module("rtb200.synth")
    os.makedirs(os.path.join(HERE, "synth"), exist_ok=True)
    ref = RefRenderer("rec")
    rng = np.random.default_rng(777)
    for name, (gen, W, H, depth) in SCENES.items():
        ps = gen(synth)
        cam = (ps.frame["camera_pos"], ps.frame["camera_look"], ps.frame["camera_up"])
        ref.new_scene(camera=cam, background=ps.frame["background"], max_bounces=depth)
        ref.add_packed(ps)
        rgb, _ = ref.render_full(W, H)
        # Pretraitement is not idempotent in the reference: a fresh scene for the per-ray entries
        ref.new_scene(camera=cam, background=ps.frame["background"], max_bounces=depth)
        ref.add_packed(ps)
        ref.prepare(W, H)
        lw, lh = 96, 64
        lx0, ly0 = (2 * W - lw) // 2, (2 * H - lh) // 2
        lattice, _ = ref.render_region(W, H, 2, lx0, ly0, lw, lh, threads=4)
        nq = 4096
        org = (np.asarray(cam[0], np.float32)[None, :] + rng.uniform(-40, 40, (nq, 3))).astype(np.float32)
        tgt = rng.uniform(-300, 300, (nq, 3)).astype(np.float32)
        dirs = tgt - org
        dirs = (dirs / np.linalg.norm(dirs, axis=1, keepdims=True)).astype(np.float32)
        idx, t, nrm = ref.nearest_hits(org, dirs)
        pts = rng.uniform(-250, 250, (nq, 3)).astype(np.float32)
        pts[:, 1] = rng.uniform(-60, 300, nq)
        lights = np.asarray(ps["lights"])[rng.integers(0, len(ps["lights"]), nq), 0:3].astype(np.float32)
        filt = ref.shadow_filters(pts, lights)
        out = {k: np.asarray(ps[k]) for k in ARRAYS}
        out.update({"frame_" + k: np.asarray(v) for k, v in ps.frame.items()})
        out.update(dict(W=W, H=H, depth=depth, rgb=rgb, lattice=lattice, lattice_rect=np.array([lx0, ly0, lw, lh]),
                        org=org, dirs=dirs, idx=idx, t=t, nrm=nrm, pts=pts, lights_q=lights, filt=filt))
        path = os.path.join(HERE, "synth", name + ".npz")
        np.savez_compressed(path, **out)
        hit = float((idx >= 0).mean())
        lit = float((filt.sum(axis=1) > 0).mean())
        print(f"{name}: T/Q/P = {ps.counts[:3]}, frame {W}x{H} d{depth}, ray hit rate {hit:.2f}, "
              f"unoccluded shadow queries {lit:.2f}, {os.path.getsize(path) / 1024:.0f} KiB")
    ref.free()


if __name__ == "__main__":
    main()
