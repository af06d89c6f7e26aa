"""GPU (-m gpu): the streamed (tiled brute-force, TMA-staged) kernels against the oracle.
The default mode culls with fused arithmetic but every accepted distance comes from the
reference's rounding sequence, so frames must be bit-identical to the CPU oracle."""
import os
from importlib import import_module

import numpy as np
import pytest

from conftest import GOLDEN, bits_equal, classify_mismatches, golden_image, packed_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def synth(rt):
    return import_module("rtb200.synth")


@pytest.mark.parametrize("n", range(1, 10))
def test_bundled_scenes_through_streamed_kernels(rt, renderer, n):
    ps = packed_scene(rt, n)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 160, 120, max_bounces=4, flags=rt.F_FORCE_STREAM)
    got, _, st = renderer.render(fr)
    assert st.path == 1
    want = golden_image(n, 4)
    if n == 4:
        assert np.allclose(got, want, rtol=0, atol=2e-6)
    else:
        assert bits_equal(got, want)


@pytest.mark.parametrize("flags_name", ["default", "exact"])
def test_soup_matches_oracle_bitwise(rt, renderer, port, synth, flags_name):
    ps = synth.soup(3000, 24, 4, seed=77)
    renderer.set_scene(ps)
    flags = rt.F_FORCE_STREAM | (rt.F_EXACT_FP if flags_name == "exact" else 0)
    fr = rt.make_frame(ps.frame, 256, 144, max_bounces=3, flags=flags)
    got, got8, st = renderer.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    # powf enters through the specular term: compare with a few-ulp tolerance on the colour,
    # and exactly on everything that is decided by geometry (ray counts)
    c = port.ray_counts()
    assert (st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract) == (c["primary"], c["shadow"], c["reflect"], c["refract"])
    assert np.allclose(got, want, rtol=0, atol=4e-6)
    want8, _ = port.quantise(want)
    assert (np.abs(got8.astype(int) - want8.astype(int)) <= 1).all()
    T, Q, P, _ = ps.counts
    n_near = st.rays_primary + st.rays_reflect + st.rays_refract
    assert st.tests_triangle >= n_near * T and st.tests_triangle <= (n_near + st.rays_shadow) * T
    assert st.tests_quadric >= n_near * Q and st.tests_plane >= n_near * P


def test_soup_ray_queries(rt, renderer, port, synth):
    ps = synth.soup(5000, 32, 2, seed=5)
    renderer.set_scene(ps)
    rng = np.random.default_rng(3)
    n = 20000
    org = (np.array([0, 150, 900], np.float32) + rng.uniform(-50, 50, (n, 3))).astype(np.float32)
    tgt = rng.uniform(-400, 400, (n, 3)).astype(np.float32)
    d = tgt - org
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    idx, t, nrm = renderer.trace_nearest(org, d, flags=rt.F_FORCE_STREAM)
    widx, wt, wn = port.nearest(ps, org, d, threads=port.max_threads)
    assert np.array_equal(idx, widx)
    assert bits_equal(t, wt) and bits_equal(nrm, wn)
    assert (idx >= 0).mean() > 0.3
    pts = rng.uniform(-300, 300, (n, 3)).astype(np.float32)
    lights = np.tile(ps["lights"][0, 0:3], (n, 1)).astype(np.float32)
    filt = renderer.trace_shadow(pts, lights, flags=rt.F_FORCE_STREAM)
    wf = port.shadow(ps, pts, lights, threads=port.max_threads)
    # the product is the reference's, in insertion order (Scene.cpp:1856): queries that met >= 3 occluders out of order are redone
    assert bits_equal(filt, wf)


def test_soup_fused_mode_within_tolerance(rt, renderer, port, synth):
    ps = synth.soup(3000, 24, 4, seed=77)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 256, 144, max_bounces=3, flags=rt.F_FORCE_STREAM | rt.F_FUSED_FP)
    got, got8, st = renderer.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    want8, _ = port.quantise(want)
    sig, _ = port.signatures(ps, fr, threads=port.max_threads)
    ok, n_bad, n_off = classify_mismatches(got8, want8, sig, radius=2)
    print(f"soup fused: within +-1 {100 * ok:.4f} %, bad {n_bad}, off-edge {n_off}")
    # opt-in fused mode on a low-resolution soup: every pixel is close to some triangle edge, so
    # rounding flips are frequent; they must all sit on silhouettes / shadow terminators
    assert ok >= 0.99 and n_off == 0


def test_heightfield_and_supersampling_streamed(rt, renderer, port, synth):
    ps = synth.heightfield(4000, 300.0, 3, seed=9)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 96, 64, max_bounces=2, spp_grid=2, flags=rt.F_FORCE_STREAM)
    got, _, st = renderer.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    assert np.allclose(got, want, rtol=0, atol=4e-6)
    assert st.rays_primary == 96 * 64 * 4


# ---- warp-bundle culling (rt_trace.cu): whole triangles skipped per warp must change no bit -----------------
def _frames_equal_exact(rt, renderer, ps, W, H, depth, spp=1):
    a, a8, sa = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp, flags=rt.F_FORCE_STREAM))
    b, b8, sb = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp, flags=rt.F_FORCE_STREAM | rt.F_EXACT_FP))
    assert (sa.rays_shadow, sa.rays_reflect, sa.rays_refract) == (sb.rays_shadow, sb.rays_reflect, sb.rays_refract)
    assert bits_equal(a, b) and np.array_equal(a8, b8)
    return sa


def test_bundle_culling_large_and_far_scenes(rt, renderer, synth):
    """Triangles 10x larger than config 3's (the pad grows with |e1||e2|), the same soup 50 000 units away from the
    world origin (rounding of apex and centres), a height field seen from inside (rays in every direction)."""
    big = synth.soup(800, 8, 3, seed=41, radius=140.0)
    renderer.set_scene(big)
    st = _frames_equal_exact(rt, renderer, big, 384, 216, 3)
    assert 0 < st.tests_triangle_executed < st.tests_triangle
    far = synth.soup(1500, 8, 2, seed=42)
    shift = np.array([50000.0, -20000.0, 30000.0], np.float32)
    arr = {k: np.array(far[k]) for k in far.a}
    arr["tri_points"] = (arr["tri_points"].reshape(-1, 3, 3) + shift).reshape(-1, 9).astype(np.float32)
    arr["lights"][:, 0:3] += shift
    q = arr["quad_coef"].astype(np.float64)          # re-centre the ellipsoids: x -> x - shift
    ctr = -q[:, 6:9] / (2 * q[:, 0:3]) + shift
    rhs = (q[:, 0:3] * (ctr - shift) ** 2).sum(1) - q[:, 9]
    q[:, 6:9] = -2 * q[:, 0:3] * ctr
    q[:, 9] = (q[:, 0:3] * ctr * ctr).sum(1) - rhs
    arr["quad_coef"] = q.astype(np.float32)
    arr["plane_coef"][:, 3] -= arr["plane_coef"][:, 0:3] @ shift
    frame = dict(far.frame)
    frame["cam_pos"] = (np.asarray(frame["cam_pos"], np.float32) + shift).astype(np.float32)
    moved = rt.PackedScene(arr, frame)
    renderer.set_scene(moved)
    _frames_equal_exact(rt, renderer, moved, 320, 180, 2)
    hf = synth.heightfield(6000, 300.0, 4, seed=43, camera=((0.0, 40.0, 0.0), (200.0, 0.0, 150.0), (0.0, 1.0, 0.0)))
    renderer.set_scene(hf)
    _frames_equal_exact(rt, renderer, hf, 320, 180, 3, spp=2)


def test_query_grouping_changes_no_bit(rt, renderer, synth, monkeypatch):
    """Queue order of the tiled path (rt_stream.cu: level-0 rays by 16x16 patches, each CTA's shadow queries light by
    light; RTB200_QUERY_GROUPING=0 restores ray order): a permutation of the queues, so frames, bytes and ray counts
    are equal -- ragged 16x16 patches (251 x 131 at 2x2 samples), three levels with both kinds of children, 8 lights."""
    cases = ((synth.soup(600, 24, 3, seed=2024, radius=45.0), 251, 131, 3, 2),
             (synth.heightfield(3000, 250.0, 8, seed=45), 250, 150, 2, 1))
    for n, (ps, W, H, depth, spp) in enumerate(cases):
        renderer.set_scene(ps)
        fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp, flags=rt.F_FORCE_STREAM)
        monkeypatch.setenv("RTB200_QUERY_GROUPING", "0")
        a, a8, sa = renderer.render(fr)
        monkeypatch.delenv("RTB200_QUERY_GROUPING")
        b, b8, sb = renderer.render(fr)
        assert bits_equal(a, b) and np.array_equal(a8, b8)
        assert (sa.rays_shadow, sa.rays_reflect, sa.rays_refract) == (sb.rays_shadow, sb.rays_reflect, sb.rays_refract)
        assert sa.rays_shadow > 0
        if n == 0:          # measured on this scene: 7.5e7 -> 3.8e7 executed tests (profiles/r1_c3_ab_query_grouping.log)
            assert sb.tests_triangle_executed < sa.tests_triangle_executed


def test_bundle_culling_ray_queries(rt, renderer, synth):
    """Coherent bundles through the ray-level entry: a common origin and a narrow fan (cones of a fraction of a
    degree), directions that are not unit vectors, parallel rays from spread origins, NaN / zero rays in between."""
    ps = synth.soup(4000, 16, 2, seed=44)
    renderer.set_scene(ps)
    rng = np.random.default_rng(6)
    n = 16384
    cam = np.asarray(ps.frame["cam_pos"], np.float32)
    tgt = (np.array([30.0, 120.0, -50.0]) + rng.uniform(-25, 25, (n, 3))).astype(np.float32)
    tgt = tgt[np.argsort(tgt[:, 0] // 4 * 1000 + tgt[:, 1] // 4)]            # neighbours in the list are neighbours in space
    d = tgt - cam
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    o = np.tile(cam, (n, 1))
    cases = {"fan": (o, d), "scaled": (o, (d * rng.choice([0.25, 1.0, 3.0], (n, 1))).astype(np.float32))}
    op = (tgt + np.array([0, 0, 900.0], np.float32)).astype(np.float32)
    cases["parallel"] = (op, np.tile(np.array([0.0, 0.0, -1.0], np.float32), (n, 1)))
    bad_o, bad_d = o.copy(), d.copy()
    bad_d[5::97] = 0.0
    bad_d[11::89, 1] = np.nan
    bad_o[13::101, 0] = np.inf
    cases["bad"] = (bad_o, bad_d)
    for name, (oo, dd) in cases.items():
        i0, t0, n0 = renderer.trace_nearest(oo, dd, flags=rt.F_FORCE_STREAM)
        i1, t1, n1 = renderer.trace_nearest(oo, dd, flags=rt.F_FORCE_STREAM | rt.F_EXACT_FP)
        assert np.array_equal(i0, i1), name
        assert bits_equal(t0, t1) and bits_equal(n0, n1), name
    assert (i0 >= 0).any()


def test_bundle_culling_many_lights(rt, renderer, synth):
    """More than 8 lights: the shadow kernel keeps no cones and scans; 8 lights: the largest cone table."""
    for n_lights in (8, 11):
        ps = synth.heightfield(3000, 250.0, n_lights, seed=45)
        renderer.set_scene(ps)
        st = _frames_equal_exact(rt, renderer, ps, 256, 144, 2)
        assert st.rays_shadow > 0


def test_bundle_culling_rays_at_triangle_borders(rt, renderer, synth):
    """The adversarial case for a conservative cull: tight bundles (128 consecutive rays aimed at ONE or two
    triangles) whose rays pass exactly over edges and vertices, or miss them by 1e-6 .. 1e-2 of an edge length --
    where CTriangle::Intersection accepts or rejects by rounding.  Common origin (a cone from a point) and parallel
    rays from spread origins (apex far behind).
      * warp-bundle culling against the plain scan (RTB200_NO_BUNDLE_CULL=1): bit for bit, no exception;
      * the default mode against the filter-free exact mode: the per-ray filter of rt_cull.cuh compares in numerator
        space with error-scaled margins (its first version, fixed margins of 1e-3, lost 5 of these 192 000 answers on
        a triangle seen 1.7e-4 rad from edge-on): bit for bit as well."""
    ps = synth.soup(1500, 0, 1, seed=46, radius=20.0)
    renderer.set_scene(ps)
    rng = np.random.default_rng(8)
    tp = ps["tri_points"].reshape(-1, 3, 3).astype(np.float64)
    T, per = len(tp), 64
    tri = np.repeat(np.arange(T), per)
    # barycentric targets on the border: vertices, edge points, and small steps across the border
    kind = rng.integers(0, 3, T * per)
    a = rng.uniform(0, 1, T * per)
    u = np.where(kind == 0, a, np.where(kind == 1, 0.0, a))
    v = np.where(kind == 0, 0.0, np.where(kind == 1, a, 1.0 - a))
    vert = rng.uniform(0, 1, T * per) < 0.15
    u = np.where(vert, rng.integers(0, 2, T * per), u).astype(np.float64)
    v = np.where(vert, np.where(u == 1.0, 0.0, rng.integers(0, 2, T * per)), v).astype(np.float64)
    step = rng.choice([0.0, 1e-6, -1e-6, 1e-4, -1e-4, 1e-2, -1e-2], T * per)
    u, v = u + step * rng.uniform(-1, 1, T * per), v + step * rng.uniform(-1, 1, T * per)
    p0, e1, e2 = tp[tri, 0], tp[tri, 1] - tp[tri, 0], tp[tri, 2] - tp[tri, 0]
    tgt = p0 + u[:, None] * e1 + v[:, None] * e2
    cam = np.asarray(ps.frame["cam_pos"], np.float64)
    d = tgt - cam
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    cases = {"from the camera": (np.tile(cam, (len(tgt), 1)), d)}
    axis = np.array([0.3, -0.2, -0.93]); axis /= np.linalg.norm(axis)
    cases["parallel"] = (tgt - 700.0 * axis, np.tile(axis, (len(tgt), 1)))
    n_hit = n_filter = 0
    try:
        for name, (o, dd) in cases.items():
            o32, d32 = o.astype(np.float32), dd.astype(np.float32)
            pts = (o + dd * 1400.0).astype(np.float32)
            os.environ["RTB200_NO_BUNDLE_CULL"] = "0"
            i0, t0, n0 = renderer.trace_nearest(o32, d32, flags=rt.F_FORCE_STREAM)
            f0 = renderer.trace_shadow(pts, o32, flags=rt.F_FORCE_STREAM)
            os.environ["RTB200_NO_BUNDLE_CULL"] = "1"
            ip, tpl, npl = renderer.trace_nearest(o32, d32, flags=rt.F_FORCE_STREAM)
            fp = renderer.trace_shadow(pts, o32, flags=rt.F_FORCE_STREAM)
            assert np.array_equal(i0, ip) and bits_equal(t0, tpl) and bits_equal(n0, npl) and bits_equal(f0, fp), name
            i1, t1, _ = renderer.trace_nearest(o32, d32, flags=rt.F_FORCE_STREAM | rt.F_EXACT_FP)
            bad = np.flatnonzero((i0 != i1) | (t0.view(np.uint32) != t1.view(np.uint32)))
            assert len(bad) == 0, (name, len(bad), bad[:5].tolist())
            n_filter += len(bad)
            n_hit += int((i0 >= 0).sum())
    finally:
        os.environ["RTB200_NO_BUNDLE_CULL"] = "0"
    print(f"border rays: {n_hit} hits of {2 * T * per}, {n_filter} answers differ from the exact mode")
    assert n_hit > T * per and n_filter == 0


def _layered_scene(rt, synth, rng, n_extra=400):
    """Transparent layers of ALL THREE surface types between the ground and the light, inserted in shuffled order:
    big horizontal triangles, flat ellipsoids and horizontal planes.  A shadow ray from below crosses up to nine of them,
    so the filter product has many factors and Scene.cpp:1856 fixes their order (insertion order)."""
    F = np.float32
    layers = []                                   # (kind, geometry)
    for i in range(3):
        y = 12.0 * (i + 1)
        layers.append(("tri", [-600, y, -600, 0, y, 800, 600, y, -600]))      # normal up: a ray coming down enters (a normal along the ray would send it back up)
    for i in range(3):
        c, a = np.array([0.0, 50.0 + 14.0 * i, 0.0]), np.array([500.0, 3.0, 500.0])
        q = 1.0 / (a * a)
        layers.append(("quad", [q[0], q[1], q[2], 0, 0, 0, *(-2.0 * c * q), float(np.sum(c * c * q) - 1.0)]))
    for i in range(3):
        layers.append(("plane", [0.0, 1.0, 0.0, -(100.0 + 9.0 * i)]))       # y = 100, 109, 118
    order = rng.permutation(len(layers))
    extra = synth.soup(n_extra, 0, 1, seed=3)["tri_points"].copy()
    extra[:, 1::3] += 600.0                                                # a cloud far above everything (fills the tiles / the tree)
    tri, tri_o, tri_m, quad, quad_o, quad_m, pl, pl_o, pl_m = [], [], [], [], [], [], [], [], []
    for slot, li in enumerate(order):
        kind, g = layers[li]
        if kind == "tri":
            tri.append(g); tri_o.append(slot); tri_m.append(li)
        elif kind == "quad":
            quad.append(g); quad_o.append(slot); quad_m.append(li)
        else:
            pl.append(g); pl_o.append(slot); pl_m.append(li)
    n_l = len(layers)
    ground_slot = n_l
    pl.append([0.0, 1.0, 0.0, 45.0]); pl_o.append(ground_slot); pl_m.append(n_l)
    tri = np.concatenate([np.asarray(tri, F).reshape(-1, 9), extra]).astype(F)
    tri_o += list(range(n_l + 1, n_l + 1 + len(extra)))
    tri_m += [n_l] * len(extra)
    mats = np.zeros((n_l + 1, 10), F)
    mats[:, 0:3] = rng.uniform(0.55, 0.99, (n_l + 1, 3))
    mats[:, 3], mats[:, 4] = 0.2, 0.8
    # transparent; every layer its own index of refraction: a ray whose ior EQUALS the surface's is treated as leaving
    # (Scene.cpp:1800, normal flipped) and would turn back at these upward-facing layers
    mats[:n_l, 8], mats[:n_l, 9] = rng.uniform(0.8, 0.99, n_l), 1.05 + 0.03 * np.arange(n_l)
    arrays = {"tri_points": tri, "tri_normal": synth.face_normals(tri), "tri_material": np.asarray(tri_m, np.int32),
              "tri_order": np.asarray(tri_o, np.int32),
              "quad_coef": np.asarray(quad, F), "quad_material": np.asarray(quad_m, np.int32), "quad_order": np.asarray(quad_o, np.int32),
              "plane_coef": np.asarray(pl, F), "plane_material": np.asarray(pl_m, np.int32), "plane_order": np.asarray(pl_o, np.int32),
              "materials": mats, "lights": np.array([[0, 400, 0, 1, 1, 1, 1.0], [150, 380, -80, 1, 1, 1, 0.5]], F)}
    base = synth.soup(10, 0, 1, seed=1)
    frame = dict(base.frame)
    return rt.PackedScene(arrays, frame)


def test_shadow_product_in_insertion_order(rt, renderer, port, synth):
    """Scene.cpp:1850-1858 multiplies the filter in m_Surfaces order.  Nine transparent layers of mixed types in shuffled
    insertion order: every path (plain tiled scan, culled scan inside a frame, LBVH) must return the oracle's bits;
    the paths that do not meet the occluders in insertion order redo such queries (rtb200_stats.shadow_fixups)."""
    rng = np.random.default_rng(21)
    ps = _layered_scene(rt, synth, rng)
    renderer.set_scene(ps)
    renderer.build_bvh()
    n = 6000
    pts = np.stack([rng.uniform(-120, 120, n), np.full(n, 1.0), rng.uniform(-120, 120, n)], axis=1).astype(np.float32)
    lights = np.tile(ps["lights"][0, 0:3], (n, 1)).astype(np.float32)
    want = port.shadow(ps, pts, lights, threads=port.max_threads)
    assert ((want > 0) & (want < 0.9)).all()                       # every query is filtered by several layers
    for flags in (rt.F_FORCE_STREAM, rt.F_FORCE_STREAM | rt.F_EXACT_FP, rt.F_USE_BVH, rt.F_FORCE_RESIDENT):
        got = renderer.trace_shadow(pts, lights, flags=flags)
        assert bits_equal(got, want), flags
    # inside a frame: the camera looks down at the ground through all the layers
    F = np.float32
    hostbind = import_module("rtb200.hostbind")
    frame = hostbind.camera_frame((0.0, 300.0, 10.0), (0.0, 0.0, 0.0), (0.0, 0.0, -1.0))
    frame["background"] = np.zeros(3, F)
    ps2 = rt.PackedScene(ps.a, frame)
    renderer.set_scene(ps2)
    renderer.build_bvh()
    fr = rt.make_frame(ps2.frame, 192, 128, max_bounces=12, flags=rt.F_FORCE_STREAM)     # deep enough to reach the ground
    wantf, _ = port.render(ps2, fr, threads=port.max_threads)
    for flags in (rt.F_FORCE_STREAM, rt.F_USE_BVH, rt.F_FORCE_RESIDENT):
        gotf, _, st = renderer.render(rt.make_frame(ps2.frame, 192, 128, max_bounces=12, flags=flags))
        assert st.depth_levels >= 9
        assert bits_equal(gotf, wantf), flags
        if flags != rt.F_FORCE_RESIDENT:
            assert st.shadow_fixups > 0


def test_lane_group_kernels_alone_give_the_reference_frame(rt, renderer, synth, port, monkeypatch):
    """RTB200_LANE_SIN = RTB200_LANE_SIN_SHADOW = 0: every bundle of the culled scan except the primary rays' goes to
    the lane kernels (a ray walked by a group of 8 lanes, rt_cullscan.cu cs_lane_walk).  Nearest hits (best of the
    lanes' bests), shadow products (product of the lanes' products; three or more factors -> fix-up list) and the
    side lists shared out over the lanes must give the oracle's frame bit for bit -- on a soup with quadrics, mirrors
    and glass, and on the layered scene whose shadow rays cross up to nine transparent surfaces of all three kinds."""
    ps = synth.soup(3000, 16, 3, seed=11)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 240, 144, max_bounces=3, flags=rt.F_FORCE_STREAM)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    c = port.ray_counts()
    base, _, _ = renderer.render(fr)                                   # the thresholds as shipped
    monkeypatch.setenv("RTB200_LANE_SIN", "0")
    monkeypatch.setenv("RTB200_LANE_SIN_SHADOW", "0")
    got, _, st = renderer.render(fr)
    assert st.rays_reflect + st.rays_refract > 300 and st.rays_shadow > 10000
    assert (st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract) == (c["primary"], c["shadow"], c["reflect"], c["refract"])
    assert bits_equal(got, base)                                       # same device, same powf: bit for bit
    assert np.allclose(got, want, rtol=0, atol=4e-6)                   # (powf of the specular term: host against device)
    rng = np.random.default_rng(21)
    pl = _layered_scene(rt, synth, rng)
    hostbind = import_module("rtb200.hostbind")
    frame = hostbind.camera_frame((0.0, 300.0, 10.0), (0.0, 0.0, 0.0), (0.0, 0.0, -1.0))
    frame["background"] = np.zeros(3, np.float32)
    ps2 = rt.PackedScene(pl.a, frame)
    renderer.set_scene(ps2)
    fr2 = rt.make_frame(ps2.frame, 160, 96, max_bounces=12, flags=rt.F_FORCE_STREAM)
    want2, _ = port.render(ps2, fr2, threads=port.max_threads)
    got2, _, st2 = renderer.render(fr2)
    assert st2.depth_levels >= 9 and st2.shadow_fixups > 0
    assert bits_equal(got2, want2)
    # the ray-level query: 128 consecutive rays that fly apart in every direction -- no cone fits, the lane kernel takes them
    renderer.set_scene(ps)
    n = 20000
    org = (np.asarray(ps.frame["cam_pos"], np.float32) * 0.3 + rng.uniform(-300, 300, (n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    gi, gt, gn = renderer.trace_nearest(org, d, flags=rt.F_FORCE_STREAM)
    wi, wt, wn = port.nearest(ps, org, d, threads=port.max_threads)
    assert np.array_equal(gi, wi) and bits_equal(gt, wt) and bits_equal(gn, wn)
    assert (wi >= 0).mean() > 0.2


def test_row_chunks_change_nothing(rt, renderer, synth, monkeypatch):
    """RTB200_CHUNK_SAMPLES: the streamed and LBVH paths render a frame larger than the chunk size in row chunks (own
    queues, counters and read-back per chunk).  A supersampled, ragged frame cut into many chunks must equal the frame
    rendered in one piece, bit for bit, floats and RGBA8."""
    soup = synth.soup(1200, 8, 3, seed=4)
    hf = synth.heightfield(6000, 300.0, 3, seed=9)
    for ps, flags, spp in ((soup, rt.F_FORCE_STREAM, 1), (soup, rt.F_FORCE_STREAM, 2), (hf, rt.F_USE_BVH, 2)):
        renderer.set_scene(ps)
        if flags & rt.F_USE_BVH:
            renderer.build_bvh()
        fr = rt.make_frame(ps.frame, 333, 301, max_bounces=3, spp_grid=spp, flags=flags)
        monkeypatch.delenv("RTB200_CHUNK_SAMPLES", raising=False)
        want, want8, st1 = renderer.render(fr)
        assert st1.chunks == 1
        monkeypatch.setenv("RTB200_CHUNK_SAMPLES", "30000")
        got, got8, st = renderer.render(fr)
        assert st.chunks >= 4
        assert bits_equal(got, want) and np.array_equal(got8, want8)
        assert (st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract) == (st1.rays_primary, st1.rays_shadow, st1.rays_reflect, st1.rays_refract)
