"""Generates tests/golden/ from the reference itself.  Run ONLY where /root/reference exists
(the build container):   python tests/golden/make_golden.py

The reference ships no tests, golden vectors or fixtures, so every pin
is produced by its own compiled code (oracle/_ref, built from the unmodified sources by
oracle/Makefile):
  scenes/sceneN.dat      the nine bundled scene files with comment / blank lines dropped
                         (inputs of the host parser; the values are untouched)
  packed/sceneN.npz      post-Pretraitement coefficients, materials, lights and camera basis as
                         seen through the reference's getters (pins the host classes + packer)
  images/sceneN_dD.npz   float RGB frames of CScene::LancerRayons() (CPU branch), 160x120,
                         as shipped (D=0) and recursion restored at depth 2 and 4
  checksums.json         sum of all float channels (binary64 accumulator) + sha256 of the RGBA8
                         PPM of the 640x480 frames (the figures of BASELINE.md section 2)
  rays/sceneN.npz        4096 seeded rays per scene: nearest hit (order, t, normal) and shadow
                         filters for seeded (point, light) pairs
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_SCENES = "/root/reference/Projet-INF8702/Scenes"

from __graft_entry__ import load_package  # noqa: E402
from oracle.portbind import packed_from_ref_dump  # noqa: E402
from oracle.refbind import RefRenderer, quantise_rgba8  # noqa: E402


def strip_scene(src, dst):
    out = []
    with open(src, "r", encoding="latin-1") as f:
        for line in f:
            s = line.rstrip("\n")
            t = s.strip()
            if not t or t.startswith("*"):
                continue
            out.append(s.rstrip())
    with open(dst, "w", encoding="latin-1") as f:
        f.write("\n".join(out) + "\n")


def main():
    rt = load_package()
    for d in ("scenes", "packed", "images", "rays"):
        os.makedirs(os.path.join(HERE, d), exist_ok=True)
    ref = RefRenderer("rec")
    shipped = RefRenderer("shipped")
    sums = {}
    rng = np.random.default_rng(20240607)
    for n in range(1, 10):
        name = f"scene{n}"
        src = os.path.join(REF_SCENES, name + ".dat")
        strip_scene(src, os.path.join(HERE, "scenes", name + ".dat"))
        # packed scene + camera
        ref.load(src)
        ref.prepare(160, 120)
        ps = packed_from_ref_dump(rt.abi, ref.dump())
        ps.save(os.path.join(HERE, "packed", name + ".npz"))
        # seeded rays: origins around the camera, directions towards the scene volume
        cam = np.asarray(ps.frame["cam_pos"], np.float32)
        nq = 4096
        org = (cam[None, :] + rng.uniform(-30, 30, (nq, 3))).astype(np.float32)
        tgt = rng.uniform(-150, 150, (nq, 3)).astype(np.float32)
        dirs = tgt - org
        dirs = (dirs / np.linalg.norm(dirs, axis=1, keepdims=True)).astype(np.float32)
        idx, t, nrm = ref.nearest_hits(org, dirs)
        pts = rng.uniform(-120, 120, (nq, 3)).astype(np.float32)
        lights = np.asarray(ps["lights"])[rng.integers(0, len(ps["lights"]), nq), 0:3].astype(np.float32)
        filt = ref.shadow_filters(pts, lights)
        np.savez_compressed(os.path.join(HERE, "rays", name + ".npz"), org=org, dirs=dirs, idx=idx, t=t, nrm=nrm,
                            pts=pts, lights=lights, filt=filt)
        # frames
        for depth in (0, 2, 4):
            ref.load(src, max_bounces=depth)
            img, _ = ref.render_full(160, 120)
            np.savez_compressed(os.path.join(HERE, "images", f"{name}_d{depth}.npz"), rgb=img)
            ref.load(src, max_bounces=depth)
            big, _ = ref.render_full(640, 480)
            q = quantise_rgba8(big)[..., :3]
            ppm = b"P6\n640 480\n255\n" + q[::-1].tobytes()
            sums[f"{name}_d{depth}"] = {"sum": float(np.sum(big.astype(np.float64))),
                                        "sha256_ppm": hashlib.sha256(ppm).hexdigest()}
        # the as-shipped build must agree with the recursion-restored build at depth 0
        shipped.load(src)
        img_s, _ = shipped.render_full(160, 120)
        z = np.load(os.path.join(HERE, "images", f"{name}_d0.npz"))["rgb"]
        assert np.array_equal(img_s.view(np.uint32), z.view(np.uint32)), name
        print(name, "done")
    with open(os.path.join(HERE, "checksums.json"), "w") as f:
        json.dump(sums, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
