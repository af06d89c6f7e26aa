"""CPU estimate (numpy, no GPU, no oracle): how many (shadow query, triangle) tests the per-light cones of
k_trace_shadow (rt_trace.cu) leave, as a function of HOW the queries are grouped into warps of 128.

Before (k_shade_a, RTB200_QUERY_GROUPING=0): the queries of one ray are contiguous, in light order, and the rays of a level-0 CTA are 256
consecutive pixels of a scanline; a warp of the shadow kernel then holds ~32 scanline pixels x 4 lights, keeps a
triangle if ANY light's cone needs it, and runs the kept triangle for all 128 queries.  Alternatives counted here:
  rows+lights : the CTA's queries ordered by light first (a warp = up to 128 scanline pixels of ONE light)
  tiles       : CTA = 16x16 pixels, queries in ray order (a warp = 16x2 pixels x 4 lights)
  tiles+lights: CTA = 16x16 pixels, ordered by light first (a warp = 16x8 pixels of ONE light) -- the default now
Scene / camera: BASELINE config 3 (synth.soup(10000, 64, 4, seed=1234), 3840x2160), crops of 256x128 pixels.
The cone maths is the numpy restatement tests/test_bundle_cull_math.py checks (including the two-cone split).

    python scripts/sim_shadow_grouping.py            # ~3 min on 8 cores
"""
import os
import sys
from importlib import import_module

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_package  # noqa: E402
import test_bundle_cull_math as bm  # noqa: E402

F = np.float32
EPS = 0.01
SPLIT_SIN = 0.02


def primary_rays(rt, ps, W, H, xs, ys):
    fr = rt.make_frame(ps.frame, W, H)
    U, V, N = (np.asarray(v, np.float64) for v in (fr.cam_u[:], fr.cam_v[:], fr.cam_n[:]))
    vx = (2.0 * xs / W - 1.0) * fr.half_w
    vy = (2.0 * ys / H - 1.0) * fr.half_h
    d = vx[:, None] * U + vy[:, None] * V - N
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.tile(np.asarray(fr.cam_pos[:], np.float64), (len(xs), 1)), d


def nearest(ps, o, d):
    """Geometric nearest hit in float64: (t, stored normal) per ray; t = inf for a miss."""
    n = len(o)
    best = np.full(n, np.inf); nrm = np.zeros((n, 3))
    tp = np.asarray(ps["tri_points"], np.float64)
    p0, e1, e2 = tp[:, 0:3], tp[:, 3:6] - tp[:, 0:3], tp[:, 6:9] - tp[:, 0:3]
    tn = np.asarray(ps["tri_normal"], np.float64)
    for a in range(0, n, 512):
        oo, dd = o[a:a + 512, None, :], d[a:a + 512, None, :]
        P = np.cross(dd, e2[None]); det = (e1[None] * P).sum(-1)
        ok = np.abs(det) >= EPS
        inv = 1.0 / np.where(ok, det, 1.0)
        S = oo - p0[None]; u = (S * P).sum(-1) * inv
        ok &= (u >= 0) & (u <= 1)
        Q = np.cross(S, e1[None]); v = (dd * Q).sum(-1) * inv
        ok &= (v >= 0) & (u + v <= 1)
        t = (e2[None] * Q).sum(-1) * inv
        t = np.where(ok & (t > EPS), t, np.inf)
        j = t.argmin(1); tj = t[np.arange(len(j)), j]
        upd = tj < best[a:a + 512]
        best[a:a + 512] = np.where(upd, tj, best[a:a + 512]); nrm[a:a + 512][upd] = tn[j[upd]]
    for q in np.asarray(ps["quad_coef"], np.float64):
        A_, B_, C_, yz, xz, xy, lx, ly, lz, K = q
        M = np.array([[A_, xy / 2, xz / 2], [xy / 2, B_, yz / 2], [xz / 2, yz / 2, C_]]); lin = np.array([lx, ly, lz])
        a = ((d @ M) * d).sum(1); b = 2 * ((o @ M) * d).sum(1) + d @ lin; c = ((o @ M) * o).sum(1) + o @ lin + K
        disc = b * b - 4 * a * c
        ok = (disc > 0) & (a != 0)
        sq = np.sqrt(np.where(ok, disc, 0)); a2 = np.where(ok, 2 * a, 1.0)
        t0, t1 = np.minimum((-b - sq) / a2, (-b + sq) / a2), np.maximum((-b - sq) / a2, (-b + sq) / a2)
        t = np.where(t0 < EPS, t1, t0)
        t = np.where(ok & (t > EPS), t, np.inf)
        upd = t < best
        Pp = o[upd] + d[upd] * t[upd, None]; g = 2 * Pp @ M + lin
        best[upd] = t[upd]; nrm[upd] = g / np.linalg.norm(g, axis=1, keepdims=True)
    for pc in np.asarray(ps["plane_coef"], np.float64):
        vd = d @ pc[:3]
        t = np.where(np.abs(vd) > EPS, -(o @ pc[:3] + pc[3]) / np.where(vd == 0, 1, vd), np.inf)
        t = np.where(t > EPS, t, np.inf)
        upd = t < best
        best[upd] = t[upd]; nrm[upd] = pc[:3]
    return best, nrm


def light_cones(L, d, dist):
    """One cone, or two after the split of k_trace_shadow; d [n, 3] binary32 towards the light."""
    def one(sel):
        b = bm.bundle_of_light(L[None], d[None, sel], dist[None, sel])
        return None if b["wide"][0] else b
    allq = np.ones(len(d), bool)
    b = one(allq)
    if b is None or b["sin_t"][0] > SPLIT_SIN:
        nd = -d.astype(np.float64)
        pa = nd[np.argmax(1 - nd @ nd[0])]
        pb = nd[np.argmax(1 - nd @ pa)]
        near_a = nd @ pa >= nd @ pb
        if near_a.any() and (~near_a).any():
            ba, bb = one(near_a), one(~near_a)
            if ba is not None and bb is not None:
                return [ba, bb]
        elif b is None:
            return None
    return None if b is None else [b]


def survivors(cones, c, r, w):
    keep = np.zeros(c.shape[1], bool)
    for b in cones:
        keep |= ~bm.culls(b, c, r, w, light=True)[0]
    return keep & (w[0] > 0)


def executed_tests(groups, lights, c, r, w):
    """groups: list of warps, each a list of (d, dist, light id) arrays.  Returns (executed, plain)."""
    T = c.shape[1]
    ex = plain = 0
    for d, dist, lid in groups:
        n = len(d); plain += n * T
        keep = np.zeros(T, bool); wide = False
        for l in np.unique(lid):
            m = lid == l
            cones = light_cones(lights[l], d[m], dist[m])
            if cones is None:
                wide = True; break
            keep |= survivors(cones, c, r, w)
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
    tot = {}
    for (x0, y0) in ((1792, 1024), (512, 640), (2816, 1408)):
        cw, ch = 256, 128
        ys, xs = np.mgrid[y0:y0 + ch, x0:x0 + cw]
        o, d = primary_rays(rt, ps, W, H, xs.ravel().astype(np.float64), ys.ravel().astype(np.float64))
        t, nrm = nearest(ps, o, d)
        hit = np.isfinite(t)
        P = (o + d * np.where(hit, t, 0)[:, None]).astype(F)
        Ld = (lights[None, :, :] - P[:, None, :]).astype(F)                     # [n, L, 3]
        face = hit[:, None] & ((Ld.astype(np.float64) * nrm[:, None, :]).sum(-1) > 0)
        dist = np.sqrt((Ld * Ld).sum(-1, dtype=F)).astype(F)
        dirs = (Ld / np.where(dist > 0, dist, 1)[..., None]).astype(F)
        px, py = xs.ravel() - x0, ys.ravel() - y0
        orders = {}
        for name, tiles, by_light in (("rows (before)", False, False), ("rows+lights", False, True), ("tiles", True, False), ("tiles+lights", True, True)):
            cta = (py // 16) * (cw // 16) + px // 16 if tiles else py * (cw // 256) + px // 256
            inner = (py % 16) * 16 + px % 16 if tiles else px % 256
            ri, li = np.nonzero(face)                                            # queries in (ray, light) order
            key = (cta[ri], li, inner[ri]) if by_light else (cta[ri], inner[ri], li)
            perm = np.lexsort(key[::-1])
            ri, li = ri[perm], li[perm]
            groups = [(dirs[ri[a:a + 128], li[a:a + 128]], dist[ri[a:a + 128], li[a:a + 128]], li[a:a + 128]) for a in range(0, len(ri), 128)]
            orders[name] = executed_tests(groups, lights, c, r, w)
        print(f"crop ({x0},{y0}) {cw}x{ch}: hit rate {hit.mean():.2f}, {int(face.sum())} shadow queries")
        for name, (ex, plain) in orders.items():
            print(f"    {name:14s} executed {ex / plain:7.4f} of the plain scan")
            a = tot.setdefault(name, [0, 0]); a[0] += ex; a[1] += plain
    print("all crops:")
    base = tot["rows (before)"][0]
    for name, (ex, plain) in tot.items():
        print(f"    {name:14s} executed {ex / plain:7.4f} of the plain scan   ({base / ex:4.2f}x fewer than before)")


if __name__ == "__main__":
    main()
