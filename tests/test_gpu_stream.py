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
    # the product is taken type by type (triangles, quadrics, planes) instead of in insertion
    # order: identical when at most one factor is not 0/1, otherwise within an ulp or two
    assert np.allclose(filt, wf, rtol=3e-7, atol=0)
    assert np.array_equal(filt == 0, wf == 0)


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
