"""LBVH probe on a synthetic height field (BASELINE config 4 shape: 1 M triangles + ground plane,
8 lights, 3840x2160, depth 4): build time, frame time, Mrays/s, bytes moved by the traversal
(nodes x 64 B + leaf triangles x 48 B) against the HBM peak, and the equality census against the
brute-force kernels (same nearest hit / same shadow filter)."""
import argparse, json, os, sys
from importlib import import_module
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package

ap = argparse.ArgumentParser()
ap.add_argument("--tris", type=int, default=1000000); ap.add_argument("--extent", type=float, default=1000.0)
ap.add_argument("--lights", type=int, default=8); ap.add_argument("--seed", type=int, default=4321)
ap.add_argument("-x", type=int, default=3840); ap.add_argument("-y", type=int, default=2160)
ap.add_argument("--depth", type=int, default=4); ap.add_argument("--spp", type=int, default=1); ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--census", type=int, default=0, help="rays compared against brute force")
ap.add_argument("--crop", type=int, default=0, help="NxN-pixel frame rendered both ways")
ap.add_argument("--padc", type=str, default="", help="comma list of PADC values for the census")
a = ap.parse_args()
rt = load_package(); synth = import_module("rtb200.synth")
ps = synth.heightfield(a.tris, a.extent, a.lights, seed=a.seed)
r = rt.Renderer(0); r.set_scene(ps)
hbm = 6529.4
try:
    hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass

def census(n):
    rng = np.random.default_rng(99)
    cam = np.asarray(ps.frame["cam_pos"], np.float32)
    org = np.tile(cam, (n, 1)).astype(np.float32)
    tgt = np.stack([rng.uniform(-a.extent, a.extent, n), rng.uniform(-60, 60, n), rng.uniform(-a.extent, a.extent, n)], axis=1).astype(np.float32)
    d = tgt - org
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    ib, tb, _ = r.trace_nearest(org, d, flags=rt.F_FORCE_STREAM)
    iv, tv, _ = r.trace_nearest(org, d, flags=rt.F_USE_BVH)
    bad_near = int(((ib != iv) | (tb.view(np.uint32) != tv.view(np.uint32))).sum())
    # shadow queries from the hit points (grazing rays included) to every light
    hit = ib >= 0
    P = (org + d * tb[:, None])[hit].astype(np.float32)
    bad_sh, n_sh = 0, 0
    for l in range(len(ps["lights"])):
        Lp = np.tile(ps["lights"][l, 0:3], (len(P), 1)).astype(np.float32)
        fb = r.trace_shadow(P, Lp, flags=rt.F_FORCE_STREAM)
        fv = r.trace_shadow(P, Lp, flags=rt.F_USE_BVH)
        bad_sh += int((fb.view(np.uint32) != fv.view(np.uint32)).any(axis=1).sum()); n_sh += len(P)
    return {"rays": n, "hits": int(hit.sum()), "nearest_mismatch": bad_near, "shadow_queries": n_sh, "shadow_mismatch": bad_sh}

padcs = [float(v) for v in a.padc.split(",") if v] or [None]
for pc in padcs:
    if pc is not None:
        os.environ["RTB200_BVH_PADC"] = str(pc)
    info = r.build_bvh()
    out = {"padc": pc, "bvh": info}
    if a.census:
        out["census"] = census(a.census)
    if a.crop:
        fb = rt.make_frame(ps.frame, a.crop, a.crop, max_bounces=a.depth, flags=rt.F_FORCE_STREAM)
        fv = rt.make_frame(ps.frame, a.crop, a.crop, max_bounces=a.depth, flags=rt.F_USE_BVH)
        A, A8, sa = r.render(fb); B, B8, sb = r.render(fv)
        out["crop"] = {"size": a.crop, "float_mismatch": int((A.view(np.uint32) != B.view(np.uint32)).sum()), "rgba8_mismatch": int((A8 != B8).sum()),
                       "brute_ms": sa.ms_render, "bvh_ms": sb.ms_render, "rays": sb.rays}
    fr = rt.make_frame(ps.frame, a.x, a.y, max_bounces=a.depth, spp_grid=a.spp, flags=rt.F_USE_BVH)
    frames = []
    for i in range(a.frames):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
        byts = st.bvh_nodes_visited * 64.0 + st.tests_triangle * 48.0
        frames.append({"ms": st.ms_render, "Mrays": st.rays / 1e6, "Mrays_per_s": st.rays / st.ms_render / 1e3,
                       "rays": [st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract], "levels": st.depth_levels, "chunks": st.chunks,
                       "launches": st.kernel_launches, "nodes_per_ray": st.bvh_nodes_visited / max(st.rays, 1), "tri_tests_per_ray": st.tests_triangle / max(st.rays, 1),
                       "traversal_GBs": byts / st.ms_render / 1e6, "frac_of_hbm": byts / st.ms_render / 1e6 / hbm, "k_nearest_l0_ms": st.ms_primary_kernel})
    out["frames"] = frames
    print(json.dumps(out), flush=True)
