"""GPU (-m gpu): BASELINE configs 3, 4 and 5 at their FULL sizes (the scenes and frames bench.py
times), checked through properties that do not need a full CPU frame (hours):
  * a small crop of the full-size frame against the oracle rendering exactly those lattice samples
    (the reference's per-ray entry CScene::ObtenirCouleur, Scene.cpp:1705, restated in oracle/),
  * the LBVH against the brute-force scan on seeded primary rays of the full-size camera,
  * idempotence (a second render is bit-identical), exact ray-count identities, the RGBA8 frame
    equal to the quantised float frame, 16-spp pixels equal to the mean of their 4x4 lattice samples.
The specular term is powf (CUDA <= 4 ulp against glibc), hence ATOL on colours; everything decided
by geometry (hit ids, distances, ray counts) must be exact."""
from importlib import import_module

import numpy as np
import pytest

from conftest import bits_equal

pytestmark = pytest.mark.gpu

ATOL = 4e-6
F = np.float32


@pytest.fixture(scope="module")
def synth(rt):
    return import_module("rtb200.synth")


def _pixel_rays(rt, ps, W, H, xs, ys):
    """Primary rays of pixels (xs, ys) of a W x H frame, Scene.cpp:1543-1552 in binary32."""
    fr = rt.make_frame(ps.frame, W, H)
    U, V, N = (np.asarray(v, F) for v in (fr.cam_u[:], fr.cam_v[:], fr.cam_n[:]))
    inv_w, inv_h = F(1.0) / F(W), F(1.0) / F(H)
    vx = ((F(2.0) * xs.astype(F) * inv_w).astype(F) - F(1.0)).astype(F) * F(fr.half_w)
    vy = ((F(2.0) * ys.astype(F) * inv_h).astype(F) - F(1.0)).astype(F) * F(fr.half_h)
    d = vx[:, None] * U[None, :] + vy[:, None] * V[None, :] - N[None, :]
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(F)
    o = np.tile(np.asarray(fr.cam_pos[:], F), (len(xs), 1))
    return o, d


def test_config3_full_size(rt, renderer, port, synth):
    """10 000 triangles + 64 quadrics + plane, 4 lights, 3840x2160, depth 3, tiled brute force."""
    ps = synth.soup(10000, 64, 4, seed=1234)
    renderer.set_scene(ps)
    W, H = 3840, 2160
    fr = rt.make_frame(ps.frame, W, H, max_bounces=3, flags=rt.F_FORCE_STREAM)
    got, got8, st = renderer.render(fr)
    assert st.path == 1 and st.rays_primary == W * H and st.nan_pixels == 0
    T, Q, P, L = ps.counts
    n_near = st.rays_primary + st.rays_reflect + st.rays_refract
    assert st.rays_shadow <= n_near * L and st.rays_reflect > 0 and st.rays_refract > 0
    assert st.tests_plane >= n_near * P and st.tests_quadric >= n_near * Q
    q, _ = port.quantise(got)
    assert np.array_equal(q, got8)
    # two crops against the oracle: image centre (dense geometry) and a corner region
    for x0, y0, w, h in ((1888, 1062, 64, 36), (40, 300, 48, 24)):
        want, _ = port.render_region(ps, fr, x0, y0, w, h, threads=port.max_threads)
        crop = got[y0:y0 + h, x0:x0 + w]
        assert np.allclose(crop, want, rtol=0, atol=ATOL), float(np.abs(crop - want).max())
    again, again8, st2 = renderer.render(fr)
    assert bits_equal(again, got) and np.array_equal(again8, got8)
    assert (st2.rays_shadow, st2.rays_reflect, st2.rays_refract) == (st.rays_shadow, st.rays_reflect, st.rays_refract)
    # warp-bundle culling skips most (ray, triangle) tests and changes no bit
    assert 0 < st.tests_triangle_executed < 0.25 * st.tests_triangle
    _count_against_exact_mode(rt, renderer, ps, W, H, 3, rt.F_FORCE_STREAM, got, got8)


def _count_against_exact_mode(rt, renderer, ps, W, H, depth, flags, got, got8):
    """The whole frame against RTB200_F_EXACT_FP (the reference's rounding sequence for EVERY test, no culling
    filter of any kind): the filters of the default mode are conservative (rt_cull.cuh: numerator-space comparisons
    with error-scaled margins; tests/test_filter_margins.py), so not one float may differ.  (With the first version's
    fixed margins config 3 at 3840x2160 had ONE pixel of 8.3 M off by 1.6e-4.)"""
    exact, exact8, stx = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=flags | rt.F_EXACT_FP))
    differ = (got.view(np.uint32) != exact.view(np.uint32)).any(axis=2)
    n = int(differ.sum())
    worst = float(np.abs(got - exact).max())
    print(f"default vs EXACT_FP at {W}x{H}: {n} differing pixel(s) of {W * H}, max |delta| = {worst:.3g}")
    assert n == 0 and np.array_equal(got8, exact8)


def test_config4_full_size(rt, renderer, port, synth):
    """1 000 000-triangle height field + plane, 8 lights, 3840x2160, depth 4, GPU-built LBVH."""
    ps = synth.heightfield(1000000, 1000.0, 8, seed=4321)
    renderer.set_scene(ps)
    info = renderer.build_bvh()
    assert info["triangles"] == 1000000
    W, H = 3840, 2160
    fr = rt.make_frame(ps.frame, W, H, max_bounces=4, flags=rt.F_USE_BVH)
    got, got8, st = renderer.render(fr)
    assert st.path == 2 and st.rays_primary == W * H and st.nan_pixels == 0
    assert st.rays_reflect > 0 and st.rays_refract == 0 and st.bvh_nodes_visited > 0
    q, _ = port.quantise(got)
    assert np.array_equal(q, got8)
    # LBVH == brute-force scan on 32 768 seeded primary rays of this camera (+ shadow queries from their hits)
    rng = np.random.default_rng(44)
    n = 32768
    xs, ys = rng.integers(0, W, n), rng.integers(0, H, n)
    o, d = _pixel_rays(rt, ps, W, H, xs, ys)
    ib, tb, nb = renderer.trace_nearest(o, d, flags=rt.F_FORCE_STREAM)
    iv, tv, nv = renderer.trace_nearest(o, d, flags=rt.F_USE_BVH)
    assert np.array_equal(ib, iv) and bits_equal(tb, tv) and bits_equal(nb, nv)
    assert (ib >= 0).mean() > 0.9
    hit = ib >= 0
    pts = (o + d * tb[:, None]).astype(F)[hit][:16384]
    lights = ps["lights"][rng.integers(0, 8, len(pts)), 0:3].astype(F)
    fb = renderer.trace_shadow(pts, lights, flags=rt.F_FORCE_STREAM)
    fv = renderer.trace_shadow(pts, lights, flags=rt.F_USE_BVH)
    assert bits_equal(fb, fv)
    # the oracle (linear scan over 1 M triangles per ray) on 96 of those rays and on a 6 x 4 crop of the frame
    m = 96
    wi, wt, _ = port.nearest(ps, o[:m], d[:m], threads=port.max_threads)
    assert np.array_equal(iv[:m], wi) and bits_equal(tv[:m], wt)
    x0, y0, w, h = 1917, 700, 6, 4
    want, _ = port.render_region(ps, fr, x0, y0, w, h, threads=port.max_threads)
    crop = got[y0:y0 + h, x0:x0 + w]
    assert np.allclose(crop, want, rtol=0, atol=ATOL), float(np.abs(crop - want).max())
    again, _, _ = renderer.render(fr)
    assert bits_equal(again, got)
    _count_against_exact_mode(rt, renderer, ps, W, H, 4, rt.F_USE_BVH, got, got8)


def test_config5_full_size(rt, renderer, port, synth):
    """100 000-triangle height field, 4 lights, depth 2, 8192x8192 at 16 spp (4x4 grid), LBVH."""
    ps = synth.heightfield(100000, 350.0, 4, seed=5)
    renderer.set_scene(ps)
    renderer.build_bvh()
    W = H = 8192
    fr = rt.make_frame(ps.frame, W, H, max_bounces=2, spp_grid=4, flags=rt.F_USE_BVH)
    _, got8, st = renderer.render(fr, want_f32=False)
    assert st.path == 2 and st.rays_primary == W * H * 16 and st.nan_pixels == 0
    f32, q8 = renderer.read_frame(H, W)                 # the float frame stays on the device until asked for
    assert np.array_equal(q8, got8)
    del q8
    # 6 x 6-pixel crops: the oracle renders their 24 x 24 lattice samples, pixel = mean of 4 x 4
    for x0, y0 in ((4093, 3000), (600, 5200)):
        w = h = 6
        lat, _ = port.render_region(ps, fr, 4 * x0, 4 * y0, 4 * w, 4 * h, threads=port.max_threads)
        want = lat.reshape(h, 4, w, 4, 3).astype(np.float64).mean(axis=(1, 3)).astype(F)
        crop = f32[y0:y0 + h, x0:x0 + w]
        assert np.allclose(crop, want, rtol=0, atol=2 * ATOL), float(np.abs(crop - want).max())
        want8, _ = port.quantise(want)
        crop8 = got8[y0:y0 + h, x0:x0 + w]
        assert (np.abs(crop8[..., :3].astype(int) - want8[..., :3].astype(int)) <= 1).all()
    del f32
    _, again8, st2 = renderer.render(fr, want_f32=False)
    assert np.array_equal(again8, got8)
    assert (st2.rays_shadow, st2.rays_reflect) == (st.rays_shadow, st.rays_reflect)
