"""CPU estimate (numpy, no GPU, no oracle): how many spheres a level-0 shadow task of the culled scan (rt_cullscan.cu)
tests, as a function of the number of triangles per level-0 block -- 32 (as built), 16 or 8.

A task = the 128 shadow queries of ONE light of a 16 x 8-pixel patch (the queue order of k_shade_a).  Its cone(s) --
bundle_of_light, split in two when wider than 1.1 degrees -- walk the hierarchy; at level 0 a block's 32 spheres are
tested when the block's own sphere survived one level up.  Counted here, per task:
  blocks visited   level-1 entries (= level-0 blocks of B triangles) whose sphere the cones do not cull
  sphere tests     B x blocks visited (what the lanes execute at level 0; the rounds are that / 32)
  survivors        triangle spheres kept (the same for every B: what the ray-by-ray tests then see)
The block spheres are built as cull_prepare builds them (Morton order of the centroids, centre = centre of the members'
bounding box, radius = largest |c_i - centre| + r_i, largest w).  Levels above level 1 are ignored (a few rounds).
Scene / camera: BASELINE config 3 (synth.soup(10000, 64, 4, seed=1234), 3840x2160), three crops of 256x128 pixels.
The cone maths is the numpy restatement tests/test_bundle_cull_math.py checks.

    python scripts/sim_block_width.py            # ~2 min
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


def expand10(v):
    v = (v * 0x00010001) & 0xFF0000FF
    v = (v * 0x00000101) & 0x0F00F00F
    v = (v * 0x00000011) & 0xC30C30C3
    v = (v * 0x00000005) & 0x49249249
    return v


def morton_order(c):
    lo, hi = c.min(0), c.max(0)
    inv = 1023.0 / max((hi - lo).max(), 1e-9)                      # cubic cells
    q = np.clip((c - lo) * inv, 0, 1023).astype(np.uint64)
    key = (expand10(q[:, 0]) << np.uint64(2)) | (expand10(q[:, 1]) << np.uint64(1)) | expand10(q[:, 2])
    return np.argsort(key, kind="stable")


def block_spheres(c, r, w, B):
    n = len(c)
    nb = (n + B - 1) // B
    bc = np.zeros((nb, 3), F); br = np.zeros(nb, F); bw = np.zeros(nb, F)
    for b in range(nb):
        m = slice(b * B, min(n, (b + 1) * B))
        cc = ((c[m].min(0) + c[m].max(0)) * 0.5).astype(F)
        bc[b] = cc
        br[b] = (np.linalg.norm(c[m] - cc, axis=1) + r[m]).max() * 1.0001
        bw[b] = w[m].max()
    return bc, br, bw


def main():
    rt = load_package()
    synth = import_module("rtb200.synth")
    ps = synth.soup(10000, 64, 4, seed=1234)
    W, H = 3840, 2160
    lights = np.asarray(ps["lights"])[:, 0:3].astype(F)
    tp = np.asarray(ps["tri_points"], F)
    c, r, w = bm.triangle_sphere(tp[:, 0:3], (tp[:, 3:6] - tp[:, 0:3]).astype(F), (tp[:, 6:9] - tp[:, 0:3]).astype(F))
    order = morton_order(c.astype(np.float64))
    c, r, w = c[order], r[order], w[order]
    widths = (32, 16, 8)
    blocks = {B: block_spheres(c, r, w, B) for B in widths}
    for B in widths:
        print(f"blocks of {B:2d}: {len(blocks[B][1]):4d} blocks, median radius {np.median(blocks[B][1]):6.1f} (triangle spheres: {np.median(r):5.1f})")
    tot = {B: [0, 0] for B in widths}
    n_tasks = n_wide = surv = 0
    for (x0, y0) in ((1792, 1024), (512, 640), (2816, 1408)):
        cw, ch = 256, 128
        ys, xs = np.mgrid[y0:y0 + ch, x0:x0 + cw]
        o, d = sg.primary_rays(rt, ps, W, H, xs.ravel().astype(np.float64), ys.ravel().astype(np.float64))
        t, nrm = sg.nearest(ps, o, d)
        hit = np.isfinite(t)
        P = (o + d * np.where(hit, t, 0)[:, None]).astype(F)
        Ld = (lights[None, :, :] - P[:, None, :]).astype(F)
        face = hit[:, None] & ((Ld.astype(np.float64) * nrm[:, None, :]).sum(-1) > 0)
        dist = np.sqrt((Ld * Ld).sum(-1, dtype=F)).astype(F)
        dirs = (Ld / np.where(dist > 0, dist, 1)[..., None]).astype(F)
        px, py = xs.ravel() - x0, ys.ravel() - y0
        cta = (py // 16) * (cw // 16) + px // 16
        inner = (py % 16) * 16 + px % 16
        ri, li = np.nonzero(face)
        perm = np.lexsort((inner[ri], li, cta[ri]))
        ri, li = ri[perm], li[perm]
        for a in range(0, len(ri), 128):
            rr, ll = ri[a:a + 128], li[a:a + 128]
            keep_t = np.zeros(len(c), bool)
            keep_b = {B: np.zeros(len(blocks[B][1]), bool) for B in widths}
            wide = False
            for l in np.unique(ll):
                m = ll == l
                cones = sg.light_cones(lights[l], dirs[rr[m], l], dist[rr[m], l])
                if cones is None or max(float(b["sin_t"][0]) for b in cones) > 0.25:      # the lane kernels take these
                    wide = True
                    break
                keep_t |= sg.survivors(cones, c[None], r[None], w[None])
                for B in widths:
                    bc, br, bw = blocks[B]
                    keep_b[B] |= sg.survivors(cones, bc[None], br[None], bw[None])
            n_tasks += 1
            if wide:
                n_wide += 1
                continue
            surv += int(keep_t.sum())
            for B in widths:
                tot[B][0] += int(keep_b[B].sum())
                tot[B][1] += B * int(keep_b[B].sum())
    n = n_tasks - n_wide
    print(f"{n_tasks} tasks of 128 queries, {n_wide} with a cone wider than sin 0.25 (lane kernels); per remaining task:")
    print(f"    triangle spheres kept (survivors): {surv / n:7.1f}")
    for B in widths:
        print(f"    blocks of {B:2d}: {tot[B][0] / n:7.1f} blocks visited, {tot[B][1] / n:8.1f} sphere tests at level 0 "
              f"({tot[B][1] / max(tot[32][1], 1):4.2f} x blocks of 32), {tot[B][1] / n / 32:6.1f} rounds of 32 lanes")


if __name__ == "__main__":
    main()
