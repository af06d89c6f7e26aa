"""GPU (-m gpu): the CUDA path through the C ABI against the CPU oracle and the golden frames.

Bars: the DEFAULT mode keeps the reference's rounding sequence and must be bit-identical to the
oracle's float frame on every bundled scene except where powf enters (scene4: <= 4 ulp of the
Phong term) -- which is the north-star gate (+-1 on >= 99.9 % of pixels) with room to spare.
The opt-in fully fused mode (RTB200_F_FUSED_FP) is held to the north-star tolerance on scenes
1-8; scene9 (two 0.99 mirrors + self-shadowing cylinders) is chaotic under ANY change of
rounding -- the CPU oracle recompiled with -ffp-contract=fast differs from itself on 1.03 % of
its pixels -- so there the count is reported and bounded at 2 %."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, bits_equal, classify_mismatches, golden_image, packed_scene

pytestmark = pytest.mark.gpu
SCENES = range(1, 10)


@pytest.mark.parametrize("n", SCENES)
@pytest.mark.parametrize("depth", [0, 2, 4])
def test_default_mode_bitwise_vs_golden(rt, renderer, n, depth):
    ps = packed_scene(rt, n)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 160, 120, max_bounces=depth)
    got, got8, st = renderer.render(fr)
    want = golden_image(n, depth)
    if n == 4:      # specular scene: powf (CUDA <= 4 ulp vs glibc) enters the Phong term
        assert np.allclose(got, want, rtol=0, atol=2e-6)
    else:
        nd = int((got.view(np.uint32) != want.view(np.uint32)).sum())
        assert nd == 0, f"{nd} differing floats"
    assert st.rays_primary == 160 * 120 and st.nan_pixels == 0


@pytest.mark.parametrize("n", SCENES)
def test_fused_mode_tolerance_vs_oracle(rt, renderer, port, n):
    ps = packed_scene(rt, n)
    renderer.set_scene(ps)
    W, H, depth = 640, 480, 4
    fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=rt.F_FUSED_FP)
    got, got8, st = renderer.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    want8, _ = port.quantise(want)
    sig, _ = port.signatures(ps, fr, threads=port.max_threads)
    ok, n_bad, n_unexplained = classify_mismatches(got8, want8, sig, radius=3)
    print(f"scene{n} fused: within +-1: {100 * ok:.4f} %, mismatching pixels {n_bad}, "
          f"further than 3 px from a silhouette/terminator {n_unexplained}, max |float diff| {np.abs(got - want).max():.3g}")
    if n == 9:
        assert ok >= 0.98
    else:
        assert ok >= 0.999
        assert n_unexplained == 0
    # ray counts are those of the reference's recursion
    c = port.ray_counts()
    assert st.rays_primary == c["primary"]
    assert abs(int(st.rays_shadow) - c["shadow"]) <= max(4, c["shadow"] // 2000)
    assert abs(int(st.rays_reflect) - c["reflect"]) <= max(4, c["reflect"] // 2000)
    assert abs(int(st.rays_refract) - c["refract"]) <= max(4, c["refract"] // 2000)


@pytest.mark.parametrize("n", SCENES)
def test_ray_queries_exact(rt, renderer, n):
    ps = packed_scene(rt, n)
    renderer.set_scene(ps)
    z = np.load(os.path.join(GOLDEN, "rays", f"scene{n}.npz"))
    idx, t, nrm = renderer.trace_nearest(z["org"], z["dirs"])
    assert np.array_equal(idx, z["idx"])
    assert bits_equal(t, z["t"]) and bits_equal(nrm, z["nrm"])
    filt = renderer.trace_shadow(z["pts"], z["lights"])
    assert bits_equal(filt, z["filt"])
    # fused mode: same winners except for rays within rounding of an edge
    idx2, t2, _ = renderer.trace_nearest(z["org"], z["dirs"], flags=rt.F_FUSED_FP)
    assert (idx2 != z["idx"]).mean() < 2e-3
    same = idx2 == z["idx"]
    assert np.allclose(t2[same], z["t"][same], rtol=1e-4, atol=1e-3)


def test_default_bounces_20_and_odd_sizes(rt, renderer, port):
    """Reference default m_NbRebondsMax = 20 (Scene.cpp:68); sizes that are not multiples of 16
    (the GLSL path silently dropped edge pixels, Scene.cpp:1308)."""
    ps = packed_scene(rt, 9)                       # two facing mirrors: long reflection chains
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 237, 131)
    assert fr.max_bounces == 20
    got, _, st = renderer.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    assert bits_equal(got, want)
    assert st.depth_levels >= 10


def test_supersampling_exact(rt, renderer, port):
    ps = packed_scene(rt, 7)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 96, 72, max_bounces=3, spp_grid=4)
    got, _, st = renderer.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    assert bits_equal(got, want)
    assert st.rays_primary == 96 * 72 * 16


def test_quantiser_and_device_resident_frame(rt, renderer, port):
    ps = packed_scene(rt, 5)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 200, 150, max_bounces=2)
    _, _, st = renderer.render(fr, want_f32=False, want_rgba8=False)     # stays on the device
    f, q = renderer.read_frame(150, 200)
    q_want, _ = port.quantise(f)
    assert np.array_equal(q, q_want)
    assert st.ms_d2h == 0.0 and st.ms_render > 0.0


def test_full_size_1080p_properties(rt, renderer, port):
    """BASELINE config 2 at full size: size-independent checks (the oracle takes ~1 s here, so
    the whole frame is compared as well)."""
    ps = packed_scene(rt, 1)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 1920, 1080, max_bounces=2)
    got, got8, st = renderer.render(fr)
    want, _ = port.render(ps, fr, threads=port.max_threads)
    assert bits_equal(got, want)
    # idempotence: a second render of the same frame is bit-identical
    again, _, _ = renderer.render(fr)
    assert bits_equal(again, got)
    # a crop of a frame equals the same pixels rendered as the lattice region by the oracle
    reg, _ = port.render_region(ps, fr, 900, 500, 64, 32, threads=4)
    assert bits_equal(got[500:532, 900:964], reg)


def test_warp_vote_culling_changes_no_bit(rt, renderer, port):
    """Resident kernels: culling by bounding-sphere / fused-filter warp votes (default) against the
    same kernels with every test evaluated (RTB200_F_EXACT_FP) and against the oracle: 400 small
    ellipsoids + 300 triangles seen from far away, so that most rays graze many bounding spheres."""
    from importlib import import_module
    synth = import_module("rtb200.synth")
    ps = synth.soup(300, 400, 3, seed=31, radius=25.0)
    renderer.set_scene(ps)
    W, H = 320, 200
    a, _, sa = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=3, flags=rt.F_FORCE_RESIDENT))
    b, _, sb = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=3, flags=rt.F_FORCE_RESIDENT | rt.F_EXACT_FP))
    assert sa.path == 0 and sb.path == 0
    assert bits_equal(a, b)
    assert (sa.rays_shadow, sa.rays_reflect, sa.rays_refract) == (sb.rays_shadow, sb.rays_reflect, sb.rays_refract)
    want, _ = port.render(ps, rt.make_frame(ps.frame, W, H, max_bounces=3), threads=port.max_threads)
    assert np.allclose(a, want, rtol=0, atol=4e-6)
    # near-tangent rays: aim at the silhouettes of the bounding spheres
    rng = np.random.default_rng(5)
    q = ps["quad_coef"]
    ctr = -q[:, 6:9] / (2 * q[:, 0:3])
    rad = np.sqrt((q[:, 0:3] * ctr * ctr).sum(1, keepdims=True) - q[:, 9:10]) / np.sqrt(q[:, 0:3])   # semi-axes
    n = 40000
    pick = rng.integers(0, len(q), n)
    off = rng.normal(size=(n, 3)); off /= np.linalg.norm(off, axis=1, keepdims=True)
    tgt = ctr[pick] + off * rad[pick] * rng.uniform(0.9, 1.1, (n, 1))
    org = (np.asarray(ps.frame["cam_pos"], np.float32) + rng.uniform(-20, 20, (n, 3))).astype(np.float32)
    d = tgt - org
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    i0, t0, n0 = renderer.trace_nearest(org, d, flags=rt.F_FORCE_RESIDENT)
    i1, t1, n1 = renderer.trace_nearest(org, d, flags=rt.F_FORCE_RESIDENT | rt.F_EXACT_FP)
    assert np.array_equal(i0, i1) and bits_equal(t0, t1) and bits_equal(n0, n1)
    wi, wt, _ = port.nearest(ps, org, d, threads=port.max_threads)
    assert np.array_equal(i0, wi) and bits_equal(t0, wt)
    lights = np.tile(ps["lights"][0, 0:3], (n, 1)).astype(np.float32)
    pts = (tgt + rng.normal(size=(n, 3)) * 30).astype(np.float32)
    f0 = renderer.trace_shadow(pts, lights, flags=rt.F_FORCE_RESIDENT)
    f1 = renderer.trace_shadow(pts, lights, flags=rt.F_FORCE_RESIDENT | rt.F_EXACT_FP)
    assert bits_equal(f0, f1)


def _with_pack_min(value, fn):
    os.environ["RTB200_PACK_ON_DEVICE_MIN"] = str(value)
    try:
        return fn()
    finally:
        del os.environ["RTB200_PACK_ON_DEVICE_MIN"]


def test_device_packing_equals_host_packing(rt, renderer):
    """Large meshes are packed by a kernel from the raw arrays (rt_pack.cu), small scenes by the host
    loop: same device scene either way -- frames and ray queries bit-equal on the resident path
    (unified list in shared memory), the streamed path (tile arrays) and the LBVH."""
    from importlib import import_module
    synth = import_module("rtb200.synth")
    NEVER, ALWAYS = 1 << 30, 1
    # resident kernels: bundled scene 2 (12 triangles + plane, 3 lights)
    ps = packed_scene(rt, 2)
    fr = rt.make_frame(ps.frame, 200, 120, max_bounces=3)
    frames = []
    for pack_min in (NEVER, ALWAYS):
        _with_pack_min(pack_min, lambda: renderer.set_scene(ps))
        frames.append(renderer.render(fr))
    assert frames[0][2].path == 0
    assert bits_equal(frames[0][0], frames[1][0]) and np.array_equal(frames[0][1], frames[1][1])
    # streamed + LBVH: soup with quadrics and planes between the triangles (orders interleaved)
    base = synth.soup(3000, 12, 2, seed=77)
    rng = np.random.default_rng(5)
    T, Q, P, _ = base.counts
    slots = rng.permutation(T + Q + P)                      # insertion order: the three types interleaved
    arrays = {k: base[k] for k in base.a}
    arrays["tri_order"] = np.sort(slots[:T]).astype(np.int32)
    arrays["quad_order"] = np.sort(slots[T:T + Q]).astype(np.int32)
    arrays["plane_order"] = np.sort(slots[T + Q:]).astype(np.int32)
    ps = rt.PackedScene(arrays, base.frame)
    org = (np.asarray(ps.frame["cam_pos"], np.float32) + rng.uniform(-20, 20, (20000, 3))).astype(np.float32)
    d = rng.normal(size=(20000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    got = []
    for pack_min in (NEVER, ALWAYS):
        _with_pack_min(pack_min, lambda: renderer.set_scene(ps))
        renderer.build_bvh()
        a = renderer.render(rt.make_frame(ps.frame, 160, 90, max_bounces=2, flags=rt.F_FORCE_STREAM))
        b = renderer.render(rt.make_frame(ps.frame, 160, 90, max_bounces=2, flags=rt.F_USE_BVH))
        q = renderer.trace_nearest(org, d, flags=rt.F_FORCE_STREAM)
        got.append((a, b, q))
    for i in range(2):
        assert bits_equal(got[0][i][0], got[1][i][0]) and np.array_equal(got[0][i][1], got[1][i][1])
    assert np.array_equal(got[0][2][0], got[1][2][0]) and bits_equal(got[0][2][1], got[1][2][1]) and bits_equal(got[0][2][2], got[1][2][2])
    assert bits_equal(got[0][0][0], got[0][1][0])           # and the LBVH frame equals the brute-force frame


@pytest.mark.parametrize("pack_min", [1 << 30, 1])
def test_invalid_orders_rejected(rt, renderer, pack_min):
    from importlib import import_module
    synth = import_module("rtb200.synth")
    base = synth.soup(50, 2, 1, seed=3)
    for key, bad in (("tri_order", lambda a: np.concatenate([a[:1], a[:-1]])),       # a duplicate, not increasing
                     ("tri_material", lambda a: np.where(np.arange(len(a)) == 7, 10 ** 6, a)),
                     ("plane_order", lambda a: a + 5)):                              # out of range / not a permutation
        arrays = {k: base[k].copy() for k in base.a}
        arrays[key] = bad(arrays[key]).astype(np.int32)
        ps = rt.PackedScene(arrays, base.frame)
        with pytest.raises(rt.RtError):
            _with_pack_min(pack_min, lambda: renderer.set_scene(ps))
