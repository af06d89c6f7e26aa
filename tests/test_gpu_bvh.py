"""GPU (-m gpu): the GPU-built LBVH must return the SAME nearest hit (order and distance bits) and
the SAME shadow filter as the brute-force linear scan (the scan it has to
agree with is Scene.cpp:1710-1715 / 1850-1858), on random, regular and adversarial scenes."""
from importlib import import_module

import numpy as np
import pytest

from conftest import bits_equal

pytestmark = pytest.mark.gpu

F = np.float32


@pytest.fixture(scope="module")
def synth(rt):
    return import_module("rtb200.synth")


def _camera_rays(ps, rng, n, jitter=30.0, lo=(-400, -40, -400), hi=(400, 360, 400)):
    cam = np.asarray(ps.frame["cam_pos"], F)
    org = (cam + rng.uniform(-jitter, jitter, (n, 3))).astype(F)
    tgt = rng.uniform(lo, hi, (n, 3)).astype(F)
    d = tgt - org
    return org, (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(F)


def _compare_queries(rt, r, ps, org, d, pts, lights):
    idx_b, t_b, n_b = r.trace_nearest(org, d, flags=rt.F_FORCE_STREAM)
    idx_v, t_v, n_v = r.trace_nearest(org, d, flags=rt.F_USE_BVH)
    assert np.array_equal(idx_b, idx_v)
    assert bits_equal(t_b, t_v) and bits_equal(n_b, n_v)
    f_b = r.trace_shadow(pts, lights, flags=rt.F_FORCE_STREAM)
    f_v = r.trace_shadow(pts, lights, flags=rt.F_USE_BVH)
    assert bits_equal(f_b, f_v)
    return idx_b, f_b


def test_bvh_requires_build(rt, renderer, synth):
    ps = synth.soup(200, 2, 1, seed=1)
    renderer.set_scene(ps)                       # a new scene invalidates any earlier LBVH
    fr = rt.make_frame(ps.frame, 32, 32, max_bounces=1, flags=rt.F_USE_BVH)
    with pytest.raises(rt.RtError) as e:
        renderer.render(fr)
    assert e.value.status == -6


def test_soup_queries_equal_brute_force(rt, renderer, port, synth):
    ps = synth.soup(6000, 16, 2, seed=11)
    renderer.set_scene(ps)
    info = renderer.build_bvh()
    assert info["triangles"] == 6000 and info["nodes"] == 6000 + 16 - 1     # the 16 ellipsoids are tree primitives too
    rng = np.random.default_rng(4)
    n = 65536
    org, d = _camera_rays(ps, rng, n)
    pts = rng.uniform(-300, 300, (n, 3)).astype(F)
    lights = np.tile(ps["lights"][0, 0:3], (n, 1)).astype(F)
    idx, filt = _compare_queries(rt, renderer, ps, org, d, pts, lights)
    assert (idx >= 0).mean() > 0.3 and (filt == 0).any() and (filt == 1).any()
    # and against the CPU oracle (the restated linear scan) on a subset
    m = 8192
    widx, wt, _ = port.nearest(ps, org[:m], d[:m], threads=port.max_threads)
    assert np.array_equal(idx[:m], widx)


def test_heightfield_frame_equals_brute_force(rt, renderer, port, synth):
    ps = synth.heightfield(30000, 400.0, 3, seed=21)
    renderer.set_scene(ps)
    renderer.build_bvh()
    W, H = 192, 108
    fr_b = rt.make_frame(ps.frame, W, H, max_bounces=3, flags=rt.F_FORCE_STREAM)
    fr_v = rt.make_frame(ps.frame, W, H, max_bounces=3, flags=rt.F_USE_BVH)
    a, a8, st_a = renderer.render(fr_b)
    b, b8, st_b = renderer.render(fr_v)
    assert st_a.path == 1 and st_b.path == 2
    assert bits_equal(a, b) and np.array_equal(a8, b8)
    assert (st_a.rays_shadow, st_a.rays_reflect, st_a.rays_refract) == (st_b.rays_shadow, st_b.rays_reflect, st_b.rays_refract)
    assert st_b.bvh_nodes_visited > 0 and 0 < st_b.tests_triangle < st_a.tests_triangle / 50
    want, _ = port.render(ps, fr_b, threads=port.max_threads)
    assert np.allclose(b, want, rtol=0, atol=4e-6)


def test_supersampled_bvh_frame(rt, renderer, synth):
    ps = synth.heightfield(8000, 300.0, 2, seed=5)
    renderer.set_scene(ps)
    renderer.build_bvh()
    a, _, _ = renderer.render(rt.make_frame(ps.frame, 64, 48, max_bounces=2, spp_grid=4, flags=rt.F_FORCE_STREAM))
    b, _, st = renderer.render(rt.make_frame(ps.frame, 64, 48, max_bounces=2, spp_grid=4, flags=rt.F_USE_BVH))
    assert bits_equal(a, b) and st.rays_primary == 64 * 48 * 16


def test_coplanar_duplicates_lowest_order_wins(rt, renderer, synth):
    """Adversarial: every triangle exists 3 times (identical vertices); the linear scan keeps the
    first one (strict <, Scene.cpp:1713), so must the LBVH, whatever order it visits them in."""
    base = synth.soup(1500, 0, 1, seed=3)
    tri = np.concatenate([base["tri_points"]] * 3)
    rng = np.random.default_rng(8)
    perm = rng.permutation(len(tri))
    tri = tri[perm]
    T = len(tri)
    arrays = {k: base[k] for k in ("quad_coef", "quad_material", "plane_coef", "plane_material", "materials", "lights")}
    arrays.update({"tri_points": tri, "tri_normal": synth.face_normals(tri),
                   "tri_material": rng.integers(0, len(base["materials"]), T).astype(np.int32),
                   "tri_order": np.arange(T, dtype=np.int32), "quad_order": np.zeros(0, np.int32),
                   "plane_order": np.array([T], np.int32)})
    ps = rt.PackedScene(arrays, base.frame)
    renderer.set_scene(ps)
    renderer.build_bvh()
    org, d = _camera_rays(ps, rng, 40000)
    pts = rng.uniform(-300, 300, (40000, 3)).astype(F)
    lights = np.tile(ps["lights"][0, 0:3], (40000, 1)).astype(F)
    idx, _ = _compare_queries(rt, renderer, ps, org, d, pts, lights)
    hit = idx[(idx >= 0) & (idx < T)]
    assert len(hit) > 1000
    # a winner is always the lowest-order copy of its geometry
    first_copy = {}
    for o, src in enumerate(perm % 1500):
        first_copy.setdefault(int(src), o)
    assert all(first_copy[int(perm[i] % 1500)] == int(i) for i in hit[:5000])


@pytest.mark.parametrize("n_tris", [1, 3, 4, 5, 8, 9, 17])
def test_tiny_scenes(rt, renderer, synth, n_tris):
    ps = synth.soup(n_tris, 2, 1, seed=40 + n_tris, radius=120.0)
    renderer.set_scene(ps)
    renderer.build_bvh()
    rng = np.random.default_rng(n_tris)
    org, d = _camera_rays(ps, rng, 20000)
    pts = rng.uniform(-300, 300, (20000, 3)).astype(F)
    lights = np.tile(ps["lights"][0, 0:3], (20000, 1)).astype(F)
    _compare_queries(rt, renderer, ps, org, d, pts, lights)


def test_stacked_transparent_occluders_keep_scan_order(rt, renderer, synth):
    """Shadow rays through several transparent triangles: the product of colour*kt must be taken
    in insertion order (Scene.cpp:1856), not in traversal order."""
    rng = np.random.default_rng(12)
    n_layers = 6
    tri = []
    for i in range(n_layers):                 # big horizontal triangles stacked at y = 10, 20, ...
        y = 10.0 * (i + 1)
        tri.append([-500, y, -500, 500, y, -500, 0, y, 700])
    tri = np.asarray(tri, F)
    tri = tri[rng.permutation(n_layers)]
    mats = np.zeros((n_layers + 1, 10), F)
    mats[:, 0:3] = rng.uniform(0.3, 0.97, (n_layers + 1, 3))
    mats[:, 3], mats[:, 4] = 0.2, 0.8
    mats[:n_layers, 8], mats[:n_layers, 9] = rng.uniform(0.5, 0.95, n_layers), 1.3
    # enough extra small triangles that the tree has internal nodes
    extra = synth.soup(300, 0, 1, seed=2)["tri_points"]
    extra[:, 1::3] += 400.0
    allt = np.concatenate([tri, extra]).astype(F)
    T = len(allt)
    tm = np.concatenate([np.arange(n_layers), np.full(len(extra), n_layers)]).astype(np.int32)
    arrays = {"tri_points": allt, "tri_normal": synth.face_normals(allt), "tri_material": tm,
              "tri_order": np.arange(T, dtype=np.int32),
              "quad_coef": np.zeros((0, 10), F), "quad_material": np.zeros(0, np.int32), "quad_order": np.zeros(0, np.int32),
              "plane_coef": np.array([[0, 1, 0, 45.0]], F), "plane_material": np.array([n_layers], np.int32),
              "plane_order": np.array([T], np.int32), "materials": mats,
              "lights": np.array([[0, 300, 0, 1, 1, 1, 1.0]], F)}
    base = synth.soup(10, 0, 1, seed=1)
    ps = rt.PackedScene(arrays, base.frame)
    renderer.set_scene(ps)
    renderer.build_bvh()
    n = 4096
    pts = np.stack([rng.uniform(-100, 100, n), np.full(n, 1.0), rng.uniform(-100, 100, n)], axis=1).astype(F)
    lights = np.tile(np.array([0, 300, 0], F), (n, 1))
    f_b = renderer.trace_shadow(pts, lights, flags=rt.F_FORCE_STREAM)
    f_v = renderer.trace_shadow(pts, lights, flags=rt.F_USE_BVH)
    assert bits_equal(f_b, f_v)
    assert ((f_b > 0) & (f_b < 0.7)).all()       # all six layers filtered the light


def _quadric_scene(rt, synth, rng, n_ell, n_cyl, n_tris, scale=1.0):
    """Ellipsoids / spheres of very different sizes (tree primitives), axis-aligned cylinders and one
    rotated ellipsoid (side list), optional triangles, ground plane; a quarter of the materials transparent."""
    base = synth.soup(max(n_tris, 1), 0, 2, seed=int(rng.integers(1 << 30)))
    ctr = rng.uniform(-300, 300, (n_ell, 3)) * scale
    ctr[:, 1] = rng.uniform(0, 250, n_ell) * scale
    rad = np.exp(rng.uniform(np.log(0.5), np.log(60.0), (n_ell, 3))) * scale
    rad[: n_ell // 3] = rad[: n_ell // 3, :1]                              # a third are spheres
    quads = [synth._ellipsoid(ctr[i], rad[i]) for i in range(n_ell)]
    for _ in range(n_cyl):                                                  # x^2 + z^2 = r^2 around (cx, cz): unbounded in y
        cx, cz, r = rng.uniform(-300, 300) * scale, rng.uniform(-300, 300) * scale, rng.uniform(3, 20) * scale
        quads.append(np.array([1, 0, 1, 0, 0, 0, -2 * cx, 0, -2 * cz, cx * cx + cz * cz - r * r], F))
    rot = synth._ellipsoid(np.array([40.0, 90.0, -30.0]) * scale, np.array([30.0, 12.0, 20.0]) * scale)
    rot[5] = 0.25 * rot[0]                                                  # an xy term: not axis-aligned any more
    quads.append(rot)
    quads = np.stack(quads).astype(F)
    Q = len(quads)
    tri = base["tri_points"][:n_tris] * scale
    T = len(tri)
    mats = np.zeros((32, 10), F)
    mats[:, 0:3] = rng.uniform(0.2, 1.0, (32, 3)); mats[:, 3], mats[:, 4], mats[:, 5], mats[:, 6] = 0.2, 0.8, 0.3, 12.0
    mats[:8, 8], mats[:8, 9] = rng.uniform(0.4, 0.9, 8), 1.3               # transparent: coloured shadow filters
    slots = rng.permutation(T + Q + 1)
    arrays = {"tri_points": tri.astype(F), "tri_normal": synth.face_normals(tri.astype(F)) if T else np.zeros((0, 3), F),
              "tri_material": rng.integers(0, 32, T).astype(np.int32), "tri_order": np.sort(slots[:T]).astype(np.int32),
              "quad_coef": quads, "quad_material": rng.integers(0, 32, Q).astype(np.int32),
              "quad_order": np.sort(slots[T:T + Q]).astype(np.int32),
              "plane_coef": np.array([[0, 1, 0, 45.0 * scale]], F), "plane_material": np.array([20], np.int32),
              "plane_order": slots[T + Q:].astype(np.int32), "materials": mats,
              "lights": np.array([[100 * scale, 600 * scale, 200 * scale, 1, 1, 1, 0.6], [-300 * scale, 500 * scale, -100 * scale, 1, 0.9, 0.8, 0.5]], F)}
    return rt.PackedScene(arrays, base.frame), ctr, rad


def _grazing_rays(rng, ctr, rad, n, far):
    """Rays aimed at the silhouette of sphere-like quadrics (closest approach = radius * (1 +- 1e-5..1e-2))."""
    k = rng.integers(0, len(ctr), n)
    c, r = ctr[k], rad[k, 0]
    org = c + rng.normal(size=(n, 3)) * far
    to_c = c - org
    to_c /= np.linalg.norm(to_c, axis=1, keepdims=True)
    side = np.cross(to_c, rng.normal(size=(n, 3)))
    side /= np.linalg.norm(side, axis=1, keepdims=True)
    eps = rng.choice([-1e-2, -1e-4, -1e-5, 1e-5, 1e-4, 1e-2], n)
    tgt = c + side * (r * (1 + eps))[:, None]
    d = tgt - org
    return org.astype(F), (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(F)


@pytest.mark.parametrize("n_tris,scale,far", [(0, 1.0, 400.0), (3000, 1.0, 400.0), (500, 0.02, 30.0), (0, 1.0, 20000.0)])
def test_bounded_quadrics_in_tree_equal_linear_scan(rt, renderer, synth, n_tris, scale, far):
    """ellipsoids / spheres as LBVH primitives (inflated boxes) must give the linear
    scan's winners and filters bit for bit -- random rays, rays grazing the silhouettes, origins far
    outside the scene (refit for a larger R), tiny scenes, quadric-only scenes."""
    rng = np.random.default_rng(1000 + n_tris)
    ps, ctr, rad = _quadric_scene(rt, synth, rng, 300, 6, n_tris, scale)
    renderer.set_scene(ps)
    info = renderer.build_bvh()
    assert info["nodes"] == n_tris + 300 - 1                                # cylinders and the rotated one stay outside
    n = 60000
    org1, d1 = _camera_rays(ps, rng, n // 2, jitter=200.0 * scale, lo=tuple(-350 * scale for _ in range(3)), hi=tuple(350 * scale for _ in range(3)))
    org2, d2 = _grazing_rays(rng, ctr[:100], rad[:100], n // 2, far * scale if far < 1000 else far)
    org, d = np.concatenate([org1, org2]), np.concatenate([d1, d2])
    pts = (ctr[rng.integers(0, len(ctr), n)] + rng.normal(size=(n, 3)) * 40 * scale).astype(F)
    lights = np.tile(ps["lights"][0, 0:3], (n, 1)).astype(F)
    idx, filt = _compare_queries(rt, renderer, ps, org, d, pts, lights)
    T = n_tris
    hit_quadric = np.isin(idx, ps["quad_order"])
    assert hit_quadric.mean() > 0.1 and ((filt > 0) & (filt < 1)).any()
    fr_b = rt.make_frame(ps.frame, 160, 90, max_bounces=2, flags=rt.F_FORCE_STREAM)
    fr_v = rt.make_frame(ps.frame, 160, 90, max_bounces=2, flags=rt.F_USE_BVH)
    a, a8, st_a = renderer.render(fr_b)
    b, b8, st_b = renderer.render(fr_v)
    assert bits_equal(a, b) and np.array_equal(a8, b8)
    if far < 1000 and scale == 1.0:   # (origins 20000 away, or a camera 10^5 radii away from a tiny scene: the boxes
                                      #  are inflated until little is culled -- still exact)
        assert st_b.tests_quadric < st_a.tests_quadric / 2                  # the tree really culls them


def test_quadrics_can_be_kept_out_of_the_tree(rt, renderer, synth):
    import os
    rng = np.random.default_rng(9)
    ps, _, _ = _quadric_scene(rt, synth, rng, 50, 2, 200)
    renderer.set_scene(ps)
    os.environ["RTB200_BVH_NO_QUADRICS"] = "1"
    try:
        info = renderer.build_bvh()
    finally:
        del os.environ["RTB200_BVH_NO_QUADRICS"]
    assert info["nodes"] == 199
    fr = rt.make_frame(ps.frame, 96, 54, max_bounces=2, flags=rt.F_USE_BVH)
    a, a8, _ = renderer.render(fr)
    renderer.build_bvh()
    b, b8, _ = renderer.render(fr)
    assert bits_equal(a, b) and np.array_equal(a8, b8)


def test_rays_that_start_beyond_the_padded_radius(rt, renderer, port, synth):
    """The boxes are padded for origins within R of the mesh (R covers the camera).  Rays that start FARTHER out -- the
    reflections off a mirror wall 4000 units away, the shadow queries of ground points near the horizon -- take the walk
    with boxes grown for their own origin (rt_bvh.cu, "far rays"; rtb200_stats.bvh_out_of_radius counts them): same
    frame as the linear scan, same as the oracle."""
    hf = synth.heightfield(20000, 300.0, 3, seed=17, p_mirror=0.2)
    a = {k: np.array(hf[k]) for k in hf.a}
    T = len(a["tri_order"])
    mats = a["materials"]
    ground = len(mats) - 1
    mats[ground, 7] = 0.5                                               # the ground reflects
    wall = np.array([[0.9, 0.9, 0.95, 0.1, 0.3, 0, 0, 0.9, 0, 0]], F)   # a mirror wall at x = -4000
    a["materials"] = np.concatenate([mats, wall]).astype(F)
    a["plane_coef"] = np.concatenate([a["plane_coef"], np.array([[1.0, 0.0, 0.0, 4000.0]], F)]).astype(F)
    a["plane_material"] = np.concatenate([a["plane_material"], [len(mats)]]).astype(np.int32)
    a["plane_order"] = np.concatenate([a["plane_order"], [T + 1]]).astype(np.int32)
    hostbind = import_module("rtb200.hostbind")
    frame = hostbind.camera_frame((-700.0, 160.0, 40.0), (-4000.0, 60.0, 0.0), (0.0, 1.0, 0.0))
    frame["background"] = np.array([0.1, 0.2, 0.5], F)
    ps = rt.PackedScene(a, frame)
    renderer.set_scene(ps)
    info = renderer.build_bvh()
    W, H = 320, 200
    fr_b = rt.make_frame(ps.frame, W, H, max_bounces=3, flags=rt.F_USE_BVH)
    got, got8, st = renderer.render(fr_b)
    assert st.path == 2 and st.bvh_out_of_radius > 1000, st.bvh_out_of_radius
    assert st.rays_reflect > W * H // 2                                  # the wall fills the view: its reflections start 3300+ units out
    ref, ref8, st_s = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=3, flags=rt.F_FORCE_STREAM))
    assert (st.rays_shadow, st.rays_reflect) == (st_s.rays_shadow, st_s.rays_reflect)
    assert bits_equal(got, ref) and np.array_equal(got8, ref8)
    x0, y0, w, h = 120, 70, 64, 40                                       # the mirrored terrain is in the middle of the view
    want, _ = port.render_region(ps, fr_b, x0, y0, w, h, threads=port.max_threads)
    crop = got[y0:y0 + h, x0:x0 + w]
    assert np.allclose(crop, want, rtol=0, atol=4e-6), float(np.abs(crop - want).max())
    assert len(np.unique(want.reshape(-1, 3), axis=0)) > 200             # (a real picture, not one flat colour)
    assert info["pad_radius"] < 3000.0                                   # the wall really lies beyond the padded radius
