"""CPU estimate (numpy, no GPU, no oracle): how many (ray, triangle) tests the warp cones of k_trace_nearest
(rt_trace.cu) leave for the LEVEL-1 rays (reflected / refracted children of the primary hits) of BASELINE config 3,
as a function of the order of the child queue and of the number of cones a warp may hold.

On the B200 the secondary levels of config 3 cost 28 of the 32 ms of the nearest-hit stage -- as much as a plain scan:
a warp of 128 consecutive children mixes rays that left different triangles in different directions, its single cone
comes out "wide" and nothing is culled.  Counted here, on three 256x128 crops of the 4K frame:
  patch order          children in the order k_shade_a queues them today (16x16 patches, per warp: reflected, refracted)
  + sorted in the CTA  the CTA's children sorted by (kind, surface the parent hit)
  x 1 / 2 / 4 cones    the warp's rays split around the most distant directions (as k_trace_shadow does per light)

    python scripts/sim_secondary_grouping.py            # ~6 min on 8 cores
"""
import os
import sys
from importlib import import_module

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from __graft_entry__ import load_package  # noqa: E402
import test_bundle_cull_math as bm  # noqa: E402
import sim_shadow_grouping as sg  # noqa: E402

F = np.float32
EPS = 0.01


def nearest_ids(ps, o, d):
    """sim_shadow_grouping.nearest plus the index of the surface hit (triangles, then quadrics, then planes)."""
    n = len(o)
    T = len(ps["tri_points"])
    t, nrm = sg.nearest(ps, o, d)
    sid = np.full(n, -1)
    P = o + d * np.where(np.isfinite(t), t, 0)[:, None]
    # identify the surface by the stored normal / the surface equation at the hit point
    tn = np.asarray(ps["tri_normal"], np.float64)
    tp = np.asarray(ps["tri_points"], np.float64)
    hit = np.isfinite(t)
    for a in range(0, n, 2048):
        sel = np.arange(a, min(a + 2048, n))[hit[a:a + 2048]]
        if not len(sel):
            continue
        same_n = np.abs(nrm[sel][:, None, :] - tn[None]).max(-1) < 1e-12
        plane_d = np.abs(((P[sel][:, None, :] - tp[None, :, 0:3]) * tn[None]).sum(-1)) < 1e-6
        cand = same_n & plane_d
        has = cand.any(1)
        sid[sel[has]] = cand[has].argmax(1)
    rest = hit & (sid < 0)
    for qi, q in enumerate(np.asarray(ps["quad_coef"], np.float64)):
        A_, B_, C_, yz, xz, xy, lx, ly, lz, K = q
        M = np.array([[A_, xy / 2, xz / 2], [xy / 2, B_, yz / 2], [xz / 2, yz / 2, C_]]); lin = np.array([lx, ly, lz])
        val = ((P @ M) * P).sum(1) + P @ lin + K
        on = rest & (np.abs(val) < 1e-6)
        sid[on] = T + qi; rest &= ~on
    sid[rest] = T + len(ps["quad_coef"])            # the ground plane
    return t, nrm, sid


def split(o, d, n_cones):
    """Index sets of the warp's rays split around the most distant directions (1, 2 or 4 sets)."""
    sets = [np.arange(len(d))]
    while len(sets) < n_cones:
        out = []
        for s in sets:
            dn = d[s].astype(np.float64); dn /= np.linalg.norm(dn, axis=1, keepdims=True)
            pa = dn[np.argmax(1 - dn @ dn[0])]; pb = dn[np.argmax(1 - dn @ pa)]
            near_a = dn @ pa >= dn @ pb
            out += [s[near_a], s[~near_a]] if near_a.any() and (~near_a).any() else [s]
        if len(out) == len(sets):
            break
        sets = out
    return sets


def executed(o, d, order, c, r, w, n_cones):
    T = c.shape[1]
    ex = plain = wide_warps = warps = 0
    for a in range(0, len(order), 128):
        idx = order[a:a + 128]
        oo, dd = o[idx], d[idx]
        keep = np.zeros(T, bool); wide = False
        for s in split(oo, dd, n_cones):
            b = bm.bundle_of_rays(oo[None, s], dd[None, s])
            if b["wide"][0]:
                wide = True; break
            keep |= ~bm.culls(b, c, r, w, light=False)[0]
        keep &= w[0] > 0
        n = len(idx)
        ex += n * (T if wide else int(keep.sum())); plain += n * T
        wide_warps += wide; warps += 1
    return ex, plain, wide_warps, warps


def main():
    rt = load_package()
    synth = import_module("rtb200.synth")
    ps = synth.soup(10000, 64, 4, seed=1234)
    W, H = 3840, 2160
    tp = np.asarray(ps["tri_points"], F)
    c, r, w = bm.triangle_sphere(tp[None, :, 0:3], (tp[:, 3:6] - tp[:, 0:3])[None].astype(F), (tp[:, 6:9] - tp[:, 0:3])[None].astype(F))
    mats = np.asarray(ps["materials"], np.float64)
    mat_of = np.concatenate([np.asarray(ps["tri_material"]), np.asarray(ps["quad_material"]), np.asarray(ps["plane_material"])])
    tot = {}
    for (x0, y0) in ((1792, 1024), (512, 640), (2816, 1408)):
        cw, ch = 256, 128
        ys, xs = np.mgrid[y0:y0 + ch, x0:x0 + cw]
        o, d = sg.primary_rays(rt, ps, W, H, xs.ravel().astype(np.float64), ys.ravel().astype(np.float64))
        t, N, sid = nearest_ids(ps, o, d)
        hit = np.isfinite(t)
        P = o + d * np.where(hit, t, 0)[:, None]
        m = mats[mat_of[np.where(hit, sid, 0)]]
        kr, kt, mior = m[:, 7] * hit, m[:, 8] * hit, m[:, 9]
        refl = d - 2 * (d * N).sum(1)[:, None] * N
        ratio = 1.0 / np.where(mior > 0, mior, 1.0)
        Z = ratio[:, None] * (d - (d * N).sum(1)[:, None] * N)
        Tt = Z - np.sqrt(np.maximum(1 - (Z * Z).sum(1), 0))[:, None] * N
        refr = np.where(((Tt * N).sum(1) < 0)[:, None], Tt, refl)
        px, py = xs.ravel() - x0, ys.ravel() - y0
        cta = (py // 16) * (cw // 16) + px // 16
        tid = (py % 16) * 16 + px % 16
        warp = tid // 32
        want = [(kr > 0.01), (kt > 0.01)]
        ray = np.concatenate([np.nonzero(want[0])[0], np.nonzero(want[1])[0]])
        kind = np.concatenate([np.zeros(want[0].sum(), int), np.ones(want[1].sum(), int)])
        co = P[ray].astype(F)
        cd = np.where(kind[:, None] == 0, refl[ray], refr[ray]).astype(F)
        orders = {"patch order": np.lexsort((tid[ray], kind, warp[ray], cta[ray])),
                  "+ sorted in the CTA": np.lexsort((tid[ray], sid[ray], kind, cta[ray]))}
        print(f"crop ({x0},{y0}) {cw}x{ch}: {len(ray)} children ({int(want[0].sum())} reflected, {int(want[1].sum())} refracted)", flush=True)
        for oname, order in orders.items():
            for nc in (1, 2, 4):
                ex, plain, ww, nw = executed(co, cd, order, c, r, w, nc)
                name = f"{oname}, {nc} cone(s)"
                print(f"    {name:34s} executed {ex / max(plain, 1):7.4f} of the plain scan, {ww} of {nw} warps wide", flush=True)
                a = tot.setdefault(name, [0, 0]); a[0] += ex; a[1] += plain
    print("all crops:")
    for name, (ex, plain) in tot.items():
        print(f"    {name:34s} executed {ex / plain:7.4f} of the plain scan")


if __name__ == "__main__":
    main()
