"""CPU: the margins of the triangle culling filter (ray-tracing-gpu_b200/csrc/rt_cull.cuh) are conservative.

The filter evaluates CTriangle::Intersection (Triangle.cpp:127-172) in fused arithmetic and may only reject a
(ray, triangle) pair the reference's own rounding sequence cannot accept.  Both evaluations are restated here in
numpy -- the reference with one binary32 rounding per operation, the fused one as exact product + sum in binary64
rounded once to binary32 -- with the constants PARSED from rt_cull.cuh, and compared on adversarial pairs: triangles
seen up to 1e-6 rad from edge-on, rays aimed exactly at edges and vertices, distances 1 .. 3000, edge lengths
0.3 .. 300, directions that are not unit vectors.  No hit the reference accepts may be rejected; the first version
of the filter (fixed margins of 1e-3 on u, v) is kept as a control and must FAIL on the same pairs."""
import os
import re

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HDR = open(os.path.join(HERE, "..", "ray-tracing-gpu_b200", "csrc", "rt_cull.cuh")).read()


def _const(name):
    m = re.search(r"#define\s+%s\s+\(?([0-9.eE+-]+)f(?:\s*\*\s*CULL_EPS24\))?" % name, HDR)
    assert m, name
    return float(m.group(1))


F = np.float32
EPS32 = F(_const("CULL_EPS24"))
assert EPS32 == F(2.0 ** -24)
RT_EPS = F(0.01)
CULL_DET, CULL_UV = F(_const("CULL_DET")), F(_const("CULL_UV"))
KU, KD, KV, KT = (F(_const(k)) * EPS32 for k in ("CULL_KU", "CULL_KD", "CULL_KV", "CULL_KT"))


def fma(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F)

def cross_ref(a, b):   # fl(fl(ay bz) - fl(az by)) ...
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2], a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], 1).astype(F)

def dot_ref(a, b):     # (x x' + y y') + z z'
    return ((a[:, 0] * b[:, 0] + a[:, 1] * b[:, 1]).astype(F) + a[:, 2] * b[:, 2]).astype(F)

def cross_fused(a, b):  # fma(ay, bz, -fl(az by))
    return np.stack([fma(a[:, 1], b[:, 2], -(a[:, 2] * b[:, 1])), fma(a[:, 2], b[:, 0], -(a[:, 0] * b[:, 2])), fma(a[:, 0], b[:, 1], -(a[:, 1] * b[:, 0]))], 1)

def dot_fused(a, b):    # fma(z, z', fma(y, y', x x'))
    return fma(a[:, 2], b[:, 2], fma(a[:, 1], b[:, 1], (a[:, 0] * b[:, 0]).astype(F)))

def reference(o, d, p0, e1, e2):
    P = cross_ref(d, e2); det = dot_ref(e1, P)
    ok = ~(np.abs(det) < RT_EPS)
    with np.errstate(all="ignore"):
        inv = (F(1) / det).astype(F)
        S = (o - p0).astype(F)
        u = (dot_ref(S, P) * inv).astype(F)
        ok &= ~((u < 0) | (u > 1))
        Q = cross_ref(S, e1)
        v = (dot_ref(d, Q) * inv).astype(F)
        ok &= ~((v < 0) | ((u + v).astype(F) > 1))
        t = (dot_ref(e2, Q) * inv).astype(F)
    return ok, t

def filt(o, d, p0, e1, e2, tmax):
    dmax = np.abs(d).max(1).astype(F)
    a1 = np.abs(e1).max(1).astype(F); a2 = np.abs(e2).max(1).astype(F)
    a2d = (a2 * dmax).astype(F)
    k_num = (KU * a2d).astype(F); e_det = ((KD * a1).astype(F) * a2d).astype(F)
    k_v = ((KV * a1).astype(F) * dmax).astype(F); k_t = ((KT * a1).astype(F) * a2).astype(F)
    P = cross_fused(d, e2); det = dot_fused(e1, P)
    S = (o - p0).astype(F)
    un = dot_fused(S, P)
    s1 = ((np.abs(S[:, 0]) + np.abs(S[:, 1])).astype(F) + np.abs(S[:, 2])).astype(F)
    ad = np.abs(det)
    E = fma(s1, k_num, e_det)
    p1 = ((ad + e_det).astype(F) >= CULL_DET) & (np.abs(fma(F(-0.5) * np.ones_like(det), det, un)) <= fma((F(0.5) + CULL_UV) * np.ones_like(ad), ad, E))
    sg = np.where(det < 0, F(-1), F(1)).astype(F)
    un2 = (un * sg).astype(F)
    Q = cross_fused(S, e1)
    vn = (dot_fused(d, Q) * sg).astype(F); tn = (dot_fused(e2, Q) * sg).astype(F)
    e_num = (s1 * k_num).astype(F); e_v = (s1 * k_v).astype(F); e_t = (s1 * k_t).astype(F)
    with np.errstate(all="ignore"):
        c1 = vn >= -e_v
        c2 = (un2 + vn).astype(F) <= fma(ad, (F(1) + CULL_UV) * np.ones_like(ad), ((e_det + e_num).astype(F) + e_v).astype(F))
        c3 = tn >= fma(F(RT_EPS * F(0.99)) * np.ones_like(ad), ad, -(fma(RT_EPS * np.ones_like(ad), e_det, e_t)))
        c4 = tn <= fma(tmax, (ad + e_det).astype(F), e_t)
    p2 = np.where(ad > F(2) * e_det, c1 & c2 & c3 & c4, True)
    return p1, p2, (c1, c2, c3, c4), ad, e_det

def old_filter(o, d, p0, e1, e2, tmax):
    P = cross_fused(d, e2); det = dot_fused(e1, P); S = (o - p0).astype(F)
    with np.errstate(all="ignore"):
        rd = (F(1) / det).astype(F)
        u = (dot_fused(S, P) * rd).astype(F)
        p1 = (np.abs(det) >= CULL_DET) & (np.abs(u - F(0.5)) <= F(0.5) + CULL_UV)
        Q = cross_fused(S, e1)
        v = (dot_fused(d, Q) * rd).astype(F); t = (dot_fused(e2, Q) * rd).astype(F)
        p2 = (v >= -CULL_UV) & ((u + v).astype(F) <= F(1) + CULL_UV) & (t > RT_EPS * F(0.99)) & (t < tmax)
    return p1, p2

def gen(n, rng, scale, graze):
    p0 = rng.uniform(-400, 400, (n, 3))
    e1 = rng.uniform(-1, 1, (n, 3)) * scale[:, None]; e2 = rng.uniform(-1, 1, (n, 3)) * scale[:, None]
    # target on / near the border or inside
    u = rng.uniform(-0.05, 1.05, n); v = rng.uniform(-0.05, 1.05, n) * (1 - np.clip(u, 0, 1))
    snap = rng.integers(0, 4, n)
    u = np.where(snap == 0, 0.0, u); v = np.where(snap == 1, 0.0, v); v = np.where(snap == 2, 1 - u, v)
    tgt = p0 + u[:, None] * e1 + v[:, None] * e2
    nrm = np.cross(e1, e2); nrm /= np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-30)
    inpl = e1 * rng.uniform(-1, 1, (n, 1)) + e2 * rng.uniform(-1, 1, (n, 1)); inpl /= np.maximum(np.linalg.norm(inpl, axis=1, keepdims=True), 1e-30)
    ang = np.where(graze, 10.0 ** rng.uniform(-6, -1, n), rng.uniform(0.05, 1.5, n))
    dirn = inpl * np.cos(ang)[:, None] + nrm * np.sin(ang)[:, None] * rng.choice([-1, 1], (n, 1))
    dist = 10.0 ** rng.uniform(0, 3.5, n)
    o = tgt - dirn * dist[:, None]
    dlen = np.where(rng.uniform(0, 1, n) < 0.1, rng.choice([0.25, 3.0], n), 1.0)
    return o.astype(F), (dirn * dlen[:, None]).astype(F), p0.astype(F), e1.astype(F), e2.astype(F)


def test_error_scaled_margins_reject_no_accepted_hit():
    rng = np.random.default_rng(1)
    tot = acc = new_bad = old_bad = cand_new = cand_old = 0
    for _ in range(6):
        n = 400000
        scale = 10.0 ** rng.uniform(-0.5, 2.5, n)
        o, d, p0, e1, e2 = gen(n, rng, scale, rng.uniform(0, 1, n) < 0.6)
        ok, t = reference(o, d, p0, e1, e2)
        hit = ok & (t > RT_EPS)
        tmax = np.where(hit, (t * F(1.001) + F(1e-3)).astype(F), F(3e38)).astype(F)     # the tightest bound a caller passes
        p1, p2, _, _, _ = filt(o, d, p0, e1, e2, tmax)
        q1, q2 = old_filter(o, d, p0, e1, e2, tmax)
        tot += n
        acc += int(hit.sum())
        new_bad += int((hit & ~(p1 & p2)).sum())
        old_bad += int((hit & ~(q1 & q2)).sum())
        cand_new += int(p1.sum())
        cand_old += int(q1.sum())
    print(f"{tot} pairs, {acc} accepted by the reference; rejected by the filter: {new_bad}; by the fixed-margin version: {old_bad}; "
          f"phase-1 candidates {cand_new / cand_old:.3f} x")
    assert acc > tot // 4
    assert new_bad == 0
    assert old_bad > 1000                     # the control: fixed margins are not conservative on these pairs
    assert cand_new < 1.15 * cand_old         # and the new margins stay tight


def test_margins_on_ordinary_pairs_cost_nothing():
    """Triangles of config-3 size seen at ordinary angles: the candidate rate of phase 1 is that of the fixed margins."""
    rng = np.random.default_rng(2)
    n = 400000
    o, d, p0, e1, e2 = gen(n, rng, np.full(n, 15.0), np.zeros(n, bool))
    tmax = np.full(n, F(3e38), F)
    p1, _, _, _, _ = filt(o, d, p0, e1, e2, tmax)
    q1, _ = old_filter(o, d, p0, e1, e2, tmax)
    assert (q1 & ~p1).sum() == 0 and p1.sum() <= 1.01 * q1.sum()
