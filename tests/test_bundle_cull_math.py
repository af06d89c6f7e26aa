"""CPU: the warp-bundle culling of the tiled kernels (ray-tracing-gpu_b200/csrc/rt_trace.cu, "warp bundles") is
conservative -- restated in numpy and attacked.

A warp bounds its 128 rays by a cone (bundle_of_rays: mean origin / direction, apex pushed back by rho / sin(theta);
bundle_of_light: apex = the light) and skips every triangle whose inflated bounding sphere (triangle_sphere + pad)
misses the cone (bundle_culls / light_cone_culls).  Contract: a triangle ANY ray of the bundle is accepted on by the
reference's rounding sequence (CTriangle::Intersection, Triangle.cpp:127-172, with the caller's t > EPSILON / t < dist)
is never culled.  Triangles are placed so that a ray of the bundle -- preferably the outermost one -- passes over their
border, at every orientation down to 1e-5 rad from edge-on; bundles span cones of 3e-4 .. 0.3 rad, origins spread over
0 .. 30 units, coordinates up to 5e4.  The same functions must still cull: triangles away from the cone are rejected."""
import numpy as np

from test_filter_margins import F, RT_EPS, reference

PADK = F(1.0) * F(5.9604645e-8) / RT_EPS          # BUNDLE_PADK


def f32(x):
    return np.asarray(x, F)


def dot(a, b):
    return (a * b).sum(-1, dtype=F)


def triangle_sphere(p0, e1, e2):
    a0 = ((e1 + e2) * F(-1.0 / 3.0)).astype(F)
    c = (p0 - a0).astype(F)
    a1, a2 = (a0 + e1).astype(F), (a0 + e2).astype(F)
    r2 = np.maximum(dot(a0, a0), np.maximum(dot(a1, a1), dot(a2, a2)))
    return c, (np.sqrt(r2) * F(1.00001)).astype(F), (np.sqrt(dot(e1, e1) * dot(e2, e2)) * F(1.00001)).astype(F)


def bundle_of_rays(o, d):
    """o, d: [B, R, 3] float32.  Returns dict of [B] arrays (state 1 = cone, 2 = wide)."""
    B, R, _ = o.shape
    l2 = dot(d, d)
    dn = (d * (F(1) / np.sqrt(l2))[..., None]).astype(F)
    n = F(R)
    m = (o.sum(1, dtype=F) / n).astype(F)
    sd = dn.sum(1, dtype=F)
    len2 = dot(sd, sd)
    wide = ~(len2 > F(0.25) * n * n)
    u = (sd / np.sqrt(np.maximum(len2, F(1e-30)))[:, None]).astype(F)
    q = (o - m[:, None, :]).astype(F)
    rho2 = dot(q, q).max(1)
    x = np.cross(u[:, None, :], dn).astype(F)
    sin2 = dot(x, x).max(1)
    wide |= ~(dot(u[:, None, :], dn) > F(0.5)).all(1)
    sin_e = np.maximum(np.sqrt(sin2) * F(1.0001) + F(2e-6), F(1e-4)).astype(F)
    wide |= ~(sin_e < F(0.7))
    m1 = np.abs(m).sum(1, dtype=F)
    rho_e = (np.sqrt(rho2) * F(1.0001) + (F(2e-6) * m1 + F(1e-6))).astype(F)
    shift = (rho_e / sin_e * F(1.00001)).astype(F)
    apex = (m - u * shift[:, None]).astype(F)
    slack = (F(4e-7) * (m1 + shift)).astype(F)
    return dict(apex=apex, u=u, sin_t=sin_e, cos2=(F(1) - sin_e * sin_e).astype(F), far=np.full(B, F(3e38), F),
                dadd=((shift + rho_e) * F(1.0001) + slack).astype(F),
                padk=(PADK * np.maximum(F(1), np.sqrt(l2.max(1)) * F(1.00001))).astype(F), slack=slack, wide=wide)


def bundle_of_light(L, d, dist):
    """L: [B, 3]; d: [B, R, 3] unit directions TOWARDS the light; dist: [B, R]."""
    B, R, _ = d.shape
    sd = (-d).sum(1, dtype=F)
    n = F(R)
    len2 = dot(sd, sd)
    wide = ~(len2 > F(0.25) * n * n)
    u = (sd / np.sqrt(np.maximum(len2, F(1e-30)))[:, None]).astype(F)
    x = np.cross(u[:, None, :], d).astype(F)
    sin2 = dot(x, x).max(1)
    wide |= ~(-dot(u[:, None, :], d) > F(0.5)).all(1)
    sin_e = (np.sqrt(sin2) * F(1.0002) + F(4e-6)).astype(F)
    wide |= ~(sin_e < F(0.7))
    far = (dist.max(1) * F(1.0001) + F(1e-4)).astype(F)
    slack = (F(4e-7) * np.abs(L).sum(1, dtype=F)).astype(F)
    return dict(apex=L, u=u, sin_t=sin_e, cos2=(F(1) - sin_e * sin_e).astype(F), far=far, dadd=(far + slack).astype(F),
                padk=np.full(B, PADK * F(1.0001), F), slack=slack, wide=wide)


def culls(b, c, r, w, light):
    """b: bundle dict of [B]; c: [B, K, 3]; r, w: [B, K].  True = culled."""
    v = (c - b["apex"][:, None, :]).astype(F)
    s = dot(v, b["u"][:, None, :])
    vv = dot(v, v)
    dist = (np.sqrt(vv) * F(1.00001)).astype(F)
    D = (dist + b["dadd"][:, None] + r).astype(F)
    pad = np.maximum(b["padk"][:, None] * D * w, F(3.8e-6) * D).astype(F)
    rp = ((r + pad) * F(1.0001) + (F(1e-5) + b["slack"][:, None])).astype(F)
    if light:
        perp2 = (vv - s * s - F(4e-6) * vv).astype(F)
        h = (s * b["sin_t"][:, None] + rp).astype(F)
    else:
        x = np.cross(v, b["u"][:, None, :]).astype(F)
        perp2 = dot(x, x)
        h = (s * b["sin_t"][:, None] + rp + F(2e-6) * dist).astype(F)
    fr = (b["far"][:, None] + rp).astype(F)
    with np.errstate(over="ignore"):
        return ((h < 0) | (perp2 * b["cos2"][:, None] > h * h) | (vv > fr * fr)) & ~b["wide"][:, None]


def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def cone_dirs(rng, axis, theta, R):
    """R unit directions within `theta` of `axis` ([B, 3], [B]); the first 8 lie ON the boundary."""
    B = len(axis)
    t1 = unit(np.cross(axis, rng.normal(size=(B, 3))))
    t2 = np.cross(axis, t1)
    phi = rng.uniform(0, 2 * np.pi, (B, R))
    a = theta[:, None] * np.sqrt(rng.uniform(0, 1, (B, R)))
    a[:, :8] = theta[:, None]
    return unit(axis[:, None, :] * np.cos(a)[..., None] + (t1[:, None, :] * np.cos(phi)[..., None] + t2[:, None, :] * np.sin(phi)[..., None]) * np.sin(a)[..., None])


def triangles_on_rays(rng, o, d, tmin, tmax, K):
    """K triangles per bundle, each with the point o_i + t d_i of one ray (outermost rays preferred) on its border or
    inside, at orientations down to 1e-5 rad from edge-on."""
    B, R, _ = o.shape
    i = np.where(rng.uniform(0, 1, (B, K)) < 0.5, rng.integers(0, 8, (B, K)), rng.integers(0, R, (B, K)))
    bi = np.arange(B)[:, None]
    oi, di = o[bi, i].astype(np.float64), unit(d[bi, i].astype(np.float64))
    t = tmin[bi, i] + (tmax[bi, i] - tmin[bi, i]) * rng.uniform(0.02, 0.98, (B, K))
    X = oi + di * t[..., None]
    scale = 10.0 ** rng.uniform(-0.5, 2.3, (B, K))
    ang = np.where(rng.uniform(0, 1, (B, K)) < 0.4, 10.0 ** rng.uniform(-5, -1, (B, K)), rng.uniform(0.05, 1.5, (B, K)))
    t1 = unit(np.cross(di, rng.normal(size=(B, K, 3))))
    nrm = unit(t1 * np.cos(ang)[..., None] + di * np.sin(ang)[..., None])      # angle between ray and plane = ang
    b1 = unit(np.cross(nrm, rng.normal(size=(B, K, 3))))
    b2 = np.cross(nrm, b1)
    e1 = (b1 * rng.uniform(0.3, 1, (B, K, 1)) + b2 * rng.uniform(-0.3, 0.3, (B, K, 1))) * scale[..., None]
    e2 = (b2 * rng.uniform(0.3, 1, (B, K, 1)) + b1 * rng.uniform(-0.3, 0.3, (B, K, 1))) * scale[..., None]
    uu = rng.uniform(0, 1, (B, K)); vv = rng.uniform(0, 1, (B, K)) * (1 - uu)
    snap = rng.integers(0, 5, (B, K))
    uu = np.where(snap == 0, 0.0, uu); vv = np.where(snap == 1, 0.0, vv); vv = np.where(snap == 2, 1 - uu, vv)
    uu = np.where(snap == 3, rng.integers(0, 2, (B, K)), uu); vv = np.where(snap == 3, np.where(uu == 1, 0, rng.integers(0, 2, (B, K))), vv)
    p0 = X - uu[..., None] * e1 - vv[..., None] * e2
    return f32(p0), f32(e1), f32(e2)


def accepted_by_any_ray(o, d, p0, e1, e2, dist=None):
    B, R, _ = o.shape
    K = p0.shape[1]
    hit = np.zeros((B, K), bool)
    for k in range(K):
        ok, t = reference(o.reshape(-1, 3), d.reshape(-1, 3), np.repeat(p0[:, k], R, 0), np.repeat(e1[:, k], R, 0), np.repeat(e2[:, k], R, 0))
        acc = ok & (t > RT_EPS)
        if dist is not None:
            acc &= t < dist.reshape(-1)
        hit[:, k] = acc.reshape(B, R).any(1)
    return hit


def _far_triangles(rng, centre, B, K):
    c = centre[:, None, :] + unit(rng.normal(size=(B, K, 3))) * rng.uniform(300, 3000, (B, K, 1))
    e1 = rng.uniform(-1, 1, (B, K, 3)) * 15.0; e2 = rng.uniform(-1, 1, (B, K, 3)) * 15.0
    return f32(c), f32(e1), f32(e2)


def test_ray_bundles_never_cull_an_accepted_triangle():
    rng = np.random.default_rng(11)
    B, R, K = 1500, 128, 24
    centre = rng.uniform(-1, 1, (B, 3)) * 10.0 ** rng.uniform(0, 4.7, (B, 1))
    axis = unit(rng.normal(size=(B, 3)))
    theta = 10.0 ** rng.uniform(-3.5, -0.55, B)
    rho = np.where(rng.uniform(0, 1, B) < 0.4, 0.0, 10.0 ** rng.uniform(-1, 1.5, B))         # 40 %: a common origin
    o = centre[:, None, :] + unit(rng.normal(size=(B, R, 3))) * (rho[:, None, None] * rng.uniform(0, 1, (B, R, 1)) ** (1 / 3))
    dlen = np.where(rng.uniform(0, 1, (B, 1, 1)) < 0.1, rng.choice([0.25, 3.0], (B, 1, 1)), 1.0)
    d = cone_dirs(rng, axis, theta, R) * dlen
    o, d = f32(o), f32(d)
    b = bundle_of_rays(o, d)
    assert (~b["wide"]).mean() > 0.9
    tmin = np.full((B, R), 1.0) / dlen[:, :, 0]; tmax = np.full((B, R), 2500.0) / dlen[:, :, 0]
    p0, e1, e2 = triangles_on_rays(rng, o, d, tmin, tmax, K)
    hit = accepted_by_any_ray(o, d, p0, e1, e2)
    c, r, w = triangle_sphere(p0, e1, e2)
    cull = culls(b, c, r, w, light=False)
    assert hit.mean() > 0.5
    assert not (hit & cull).any(), int((hit & cull).sum())
    # and it still culls: small triangles scattered around the bundle's origin, mostly off the cone
    pf, f1, f2 = _far_triangles(rng, centre, B, K)
    cf, rf, wf = triangle_sphere(pf, f1, f2)
    away = culls(b, cf, rf, wf, light=False)
    hit_f = accepted_by_any_ray(o, d, pf, f1, f2)
    assert not (hit_f & away).any()
    tight = (theta < 0.02) & ~b["wide"]
    assert away[tight].mean() > 0.95


def test_light_cones_never_cull_an_occluder():
    rng = np.random.default_rng(12)
    B, R, K = 1500, 128, 24
    L = rng.uniform(-1, 1, (B, 3)) * 10.0 ** rng.uniform(1, 4.5, (B, 1))
    m = L + unit(rng.normal(size=(B, 3))) * 10.0 ** rng.uniform(1.5, 3.3, (B, 1))
    rho = 10.0 ** rng.uniform(-1, 1.7, B)
    pts = m[:, None, :] + unit(rng.normal(size=(B, R, 3))) * (rho[:, None, None] * rng.uniform(0, 1, (B, R, 1)) ** (1 / 3))
    pts, L32 = f32(pts), f32(L)
    Ld = (L32[:, None, :] - pts).astype(F)                       # Scene.cpp:1847-1848 in binary32
    dist = np.sqrt(dot(Ld, Ld)).astype(F)
    d = (Ld * (F(1) / dist)[..., None]).astype(F)
    b = bundle_of_light(L32, d, dist)
    assert (~b["wide"]).mean() > 0.8
    tmin = np.full((B, R), 0.02); tmax = dist.astype(np.float64)
    p0, e1, e2 = triangles_on_rays(rng, pts, d, tmin, tmax, K)
    occ = accepted_by_any_ray(pts, d, p0, e1, e2, dist=dist)
    c, r, w = triangle_sphere(p0, e1, e2)
    cull = culls(b, c, r, w, light=True)
    assert occ.mean() > 0.5
    assert not (occ & cull).any(), int((occ & cull).sum())
    pf, f1, f2 = _far_triangles(rng, m, B, K)
    cf, rf, wf = triangle_sphere(pf, f1, f2)
    away = culls(b, cf, rf, wf, light=True)
    occ_f = accepted_by_any_ray(pts, d, pf, f1, f2, dist=dist)
    assert not (occ_f & away).any()
    tight = (b["sin_t"] < 0.05) & ~b["wide"]
    assert away[tight].mean() > 0.9


# ---- the lane walk (rt_cullscan.cu: cs_lane_ray / cs_lane_query / cs_ray_misses): ONE ray against a sphere --------
def lane_ray(o, d):
    """nearest-hit ray: half line from o along d.  o, d: [B, 3]."""
    l2 = dot(d, d)
    u = (d * (F(1) / np.sqrt(l2))[:, None]).astype(F)
    slack = (F(4e-7) * np.abs(o).sum(1, dtype=F) + F(1e-6)).astype(F)
    return dict(o=o, u=u, far=np.full(len(o), F(3e38), F), dadd=slack, slack=slack,
                padk=(PADK * np.maximum(F(1), np.sqrt(l2) * F(1.00001))).astype(F))


def lane_query(L, d, dist):
    """shadow query seen from the light: half line from L along -d, hits within `far` of L; the query's own origin --
    what the pad scales with -- lies within dadd = far of L."""
    far = (dist * F(1.0001) + F(1e-4)).astype(F)
    slack = (F(4e-7) * np.abs(L).sum(1, dtype=F) + F(1e-6)).astype(F)
    return dict(o=L, u=(-d).astype(F), far=far, dadd=(far + slack).astype(F), slack=slack, padk=np.full(len(L), PADK * F(1.0001), F))


def ray_misses(r, c, rad, w):
    """r: dict of [B]; c: [B, K, 3]; rad, w: [B, K].  True = the inflated sphere cannot hold a reported hit of the ray."""
    v = (c - r["o"][:, None, :]).astype(F)
    s = dot(v, r["u"][:, None, :])
    vv = dot(v, v)
    dist = (np.sqrt(vv) * F(1.00001)).astype(F)
    D = (dist + r["dadd"][:, None] + rad).astype(F)
    pad = np.maximum(r["padk"][:, None] * D * w, F(3.8e-6) * D).astype(F)
    rp = ((rad + pad) * F(1.0001) + (F(1e-5) + r["slack"][:, None] + F(2e-6) * dist)).astype(F)
    perp2 = (vv - s * s - F(4e-6) * vv).astype(F)
    with np.errstate(over="ignore"):
        return (perp2 > rp * rp) | (s < -rp) | (s > r["far"][:, None] + rp)


def test_lane_walk_never_skips_an_accepted_triangle():
    """One ray per group of lanes (the bundles no cone fits): a triangle the reference accepts the ray on -- placed with
    the ray over its border, down to 1e-5 rad from edge-on, coordinates up to 5e4 -- is never behind a sphere test that
    says "miss" (without the pad and the slacks the same attack finds ~3 700 violations in 240 000 placements); and the
    test still rejects what lies off the ray."""
    rng = np.random.default_rng(13)
    B, K = 40000, 12
    # ---- nearest-hit rays: any origin, any direction, |d| in {0.25, 1, 3}
    o = rng.uniform(-1, 1, (B, 3)) * 10.0 ** rng.uniform(0, 4.7, (B, 1))
    dlen = np.where(rng.uniform(0, 1, (B, 1)) < 0.1, rng.choice([0.25, 3.0], (B, 1)), 1.0)
    d = unit(rng.normal(size=(B, 3))) * dlen
    o1, d1 = f32(o)[:, None, :], f32(d)[:, None, :]
    tmin, tmax = np.full((B, 1), 0.02) / dlen, np.full((B, 1), 3000.0) / dlen
    rep = lambda a: np.repeat(a, 8, axis=1)                     # (the helper picks among >= 8 rays: eight copies of the one)
    p0, e1, e2 = triangles_on_rays(rng, rep(o1), rep(d1), rep(tmin), rep(tmax), K)
    hit = accepted_by_any_ray(o1, d1, p0, e1, e2)
    c, rad, w = triangle_sphere(p0, e1, e2)
    miss = ray_misses(lane_ray(o1[:, 0], d1[:, 0]), c, rad, w)
    assert hit.mean() > 0.3                                    # (a lone ray over a border: accepted about half of the time)
    assert not (hit & miss).any(), int((hit & miss).sum())
    pf, f1, f2 = _far_triangles(rng, o, B, K)
    cf, rf, wf = triangle_sphere(pf, f1, f2)
    away = ray_misses(lane_ray(o1[:, 0], d1[:, 0]), cf, rf, wf)
    assert not (accepted_by_any_ray(o1, d1, pf, f1, f2) & away).any()
    assert away.mean() > 0.95
    # ---- shadow queries: hit point -> light, Scene.cpp:1847-1848 in binary32
    L = rng.uniform(-1, 1, (B, 3)) * 10.0 ** rng.uniform(1, 4.5, (B, 1))
    pts = L + unit(rng.normal(size=(B, 3))) * 10.0 ** rng.uniform(0.5, 3.5, (B, 1))
    pts, L32 = f32(pts), f32(L)
    Ld = (L32 - pts).astype(F)
    dist = np.sqrt(dot(Ld, Ld)).astype(F)
    dq = (Ld * (F(1) / dist)[:, None]).astype(F)
    p1, d1 = pts[:, None, :], dq[:, None, :]
    p0, e1, e2 = triangles_on_rays(rng, rep(p1), rep(d1), np.full((B, 8), 0.02), rep(dist[:, None].astype(np.float64)), K)
    occ = accepted_by_any_ray(p1, d1, p0, e1, e2, dist=dist[:, None])
    c, rad, w = triangle_sphere(p0, e1, e2)
    miss = ray_misses(lane_query(L32, dq, dist), c, rad, w)
    assert occ.mean() > 0.3
    assert not (occ & miss).any(), int((occ & miss).sum())
    # beyond the light and behind the hit point nothing occludes, and the test knows
    beyond = f32(L + unit(L - pts.astype(np.float64)) * rng.uniform(40, 400, (B, 1)))[:, None, :]
    e = f32(rng.uniform(-1, 1, (B, 1, 3)) * 5.0)
    cb, rb, wb = triangle_sphere(beyond, e, np.roll(e, 1, axis=2))
    assert ray_misses(lane_query(L32, dq, dist), cb, rb, wb).mean() > 0.99
