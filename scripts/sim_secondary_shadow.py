"""CPU estimate (numpy, no GPU, no oracle): the shadow queries of the LEVEL-1 hits of BASELINE config 3 -- how many
(query, triangle) tests the per-light cones of k_trace_shadow leave
  cta order     as queued today: CTAs of 256 consecutive level-1 rays, each CTA's queries light by light
  sorted        all queries of the level sorted by (light, Morton cell of the hit point, 10 bits per axis)
with 2 cones per light (today) and with 4.  Level-1 hit points are scattered over the scene (a mirror triangle reflects
its patch of rays onto many surfaces), so the cones of consecutive queries are wide; a sort makes neighbours in the
queue neighbours in space.

    python scripts/sim_secondary_shadow.py            # ~8 min on 8 cores
"""
import os
import sys
from importlib import import_module

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for sub in ("", "tests", "scripts"):
    sys.path.insert(0, os.path.join(ROOT, sub))
from __graft_entry__ import load_package  # noqa: E402
import test_bundle_cull_math as bm  # noqa: E402
import sim_shadow_grouping as sg  # noqa: E402
import sim_secondary_grouping as s2  # noqa: E402

F = np.float32


def part1by2(v):
    v = v.astype(np.uint64) & 0x3ff
    v = (v | (v << 16)) & 0x30000ff
    v = (v | (v << 8)) & 0x300f00f
    v = (v | (v << 4)) & 0x30c30c3
    v = (v | (v << 2)) & 0x9249249
    return v


def light_cones_n(L, d, dist, n_cones):
    """As sim_shadow_grouping.light_cones, with up to n_cones parts (split only while a cone is wide / > 1.1 deg)."""
    sets = [np.arange(len(d))]
    def cone(s):
        b = bm.bundle_of_light(L[None], d[None, s], dist[None, s])
        return None if b["wide"][0] else b
    cones = [cone(sets[0])]
    while len(sets) < n_cones:
        new_sets, new_cones, changed = [], [], False
        for s, b in zip(sets, cones):
            if (b is None or b["sin_t"][0] > sg.SPLIT_SIN) and len(s) > 1:
                nd = -d[s].astype(np.float64)
                pa = nd[np.argmax(1 - nd @ nd[0])]; pb = nd[np.argmax(1 - nd @ pa)]
                near_a = nd @ pa >= nd @ pb
                if near_a.any() and (~near_a).any():
                    new_sets += [s[near_a], s[~near_a]]; new_cones += [cone(s[near_a]), cone(s[~near_a])]; changed = True
                    continue
            new_sets.append(s); new_cones.append(b)
        sets, cones = new_sets, new_cones
        if not changed:
            break
    return None if any(b is None for b in cones) else cones


def executed(groups, lights, c, r, w, n_cones):
    T = c.shape[1]
    ex = plain = 0
    for d, dist, lid in groups:
        n = len(d); plain += n * T
        keep = np.zeros(T, bool); wide = False
        for l in np.unique(lid):
            m = lid == l
            cones = light_cones_n(lights[l], d[m], dist[m], n_cones)
            if cones is None:
                wide = True; break
            keep |= sg.survivors(cones, c, r, w)
        ex += n * (T if wide else int(keep.sum()))
    return ex, plain


def main():
    rt = load_package()
    synth = import_module("rtb200.synth")
    ps = synth.soup(10000, 64, 4, seed=1234)
    W, H = 3840, 2160
    lights = np.asarray(ps["lights"])[:, 0:3].astype(F)
    tp = np.asarray(ps["tri_points"], F)
    c, r, w = bm.triangle_sphere(tp[None, :, 0:3], (tp[:, 3:6] - tp[:, 0:3])[None].astype(F), (tp[:, 6:9] - tp[:, 0:3])[None].astype(F))
    mats = np.asarray(ps["materials"], np.float64)
    mat_of = np.concatenate([np.asarray(ps["tri_material"]), np.asarray(ps["quad_material"]), np.asarray(ps["plane_material"])])
    tot = {}
    for (x0, y0) in ((1792, 1024), (2816, 1408)):
        cw, ch = 256, 128
        ys, xs = np.mgrid[y0:y0 + ch, x0:x0 + cw]
        o, d = sg.primary_rays(rt, ps, W, H, xs.ravel().astype(np.float64), ys.ravel().astype(np.float64))
        t, N, sid = s2.nearest_ids(ps, o, d)
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
        want = [(kr > 0.01), (kt > 0.01)]
        ray = np.concatenate([np.nonzero(want[0])[0], np.nonzero(want[1])[0]])
        kind = np.concatenate([np.zeros(want[0].sum(), int), np.ones(want[1].sum(), int)])
        order = np.lexsort((tid[ray], kind, tid[ray] // 32, cta[ray]))           # the level-1 queue as k_shade_a fills it
        co = P[ray][order].astype(np.float64)
        cd = np.where(kind[:, None] == 0, refl[ray], refr[ray])[order]
        cd /= np.linalg.norm(cd, axis=1, keepdims=True)
        t1, N1 = sg.nearest(ps, co, cd)                                            # level-1 hits
        hit1 = np.isfinite(t1)
        P1 = (co + cd * np.where(hit1, t1, 0)[:, None]).astype(F)
        Ld = (lights[None] - P1[:, None, :]).astype(F)
        face = hit1[:, None] & ((Ld.astype(np.float64) * N1[:, None, :]).sum(-1) > 0)
        dist = np.sqrt((Ld * Ld).sum(-1, dtype=F)).astype(F)
        dirs = (Ld / np.where(dist > 0, dist, 1)[..., None]).astype(F)
        ri, li = np.nonzero(face)
        cta1 = ri // 256
        lo, hi = P1[hit1].min(0), P1[hit1].max(0)
        cell = np.clip(((P1 - lo) / np.maximum(hi - lo, 1e-9) * 1023.0), 0, 1023).astype(np.int64)
        morton = part1by2(cell[:, 0]) | (part1by2(cell[:, 1]) << np.uint64(1)) | (part1by2(cell[:, 2]) << np.uint64(2))
        orders = {"cta order": np.lexsort((ri, li, cta1)), "sorted": np.lexsort((morton[ri], li))}
        print(f"crop ({x0},{y0}): {len(co)} level-1 rays, {int(hit1.sum())} hits, {len(ri)} shadow queries", flush=True)
        for oname, perm in orders.items():
            a_r, a_l = ri[perm], li[perm]
            groups = [(dirs[a_r[a:a + 128], a_l[a:a + 128]], dist[a_r[a:a + 128], a_l[a:a + 128]], a_l[a:a + 128]) for a in range(0, len(a_r), 128)]
            for nc in (2, 4):
                ex, plain = executed(groups, lights, c, r, w, nc)
                name = f"{oname}, {nc} cones per light"
                print(f"    {name:32s} executed {ex / max(plain, 1):7.4f} of the plain scan", flush=True)
                a = tot.setdefault(name, [0, 0]); a[0] += ex; a[1] += plain
    print("all crops:")
    for name, (ex, plain) in tot.items():
        print(f"    {name:32s} executed {ex / plain:7.4f} of the plain scan")


if __name__ == "__main__":
    main()
