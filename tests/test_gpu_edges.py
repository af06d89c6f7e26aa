"""GPU (-m gpu): edge cases against the C oracle, through all three paths (scene-resident fused
kernels, TMA-tiled brute force, LBVH): empty scenes, scenes without lights, 1-pixel and 1-row images,
degenerate triangles, depth 0 with mirrors, non-finite coefficients (NaN pixels are quantised to 0
and counted), supersampling on a tiny odd image."""
from importlib import import_module

import numpy as np
import pytest

from conftest import bits_equal

pytestmark = pytest.mark.gpu

F = np.float32
CAMERA = {"cam_pos": [0.0, 60.0, 300.0], "cam_u": [1.0, 0.0, 0.0], "cam_v": [0.0, 1.0, 0.0], "cam_n": [0.0, 0.0, 1.0],
          "half_h": 0.41421357, "background": [0.1, 0.2, 0.6], "max_bounces": 3, "min_energy": 0.01, "scene_ior": 1.0}


def _scene(rt, tri=(), quads=(), planes=(), mats=None, lights=()):
    tri = np.asarray(tri, F).reshape(-1, 9)
    quads = np.asarray(quads, F).reshape(-1, 10)
    planes = np.asarray(planes, F).reshape(-1, 4)
    T, Q, P = len(tri), len(quads), len(planes)
    synth = import_module("rtb200.synth")
    if mats is None:
        mats = np.array([[0.8, 0.3, 0.2, 0.2, 0.8, 0.4, 20.0, 0, 0, 0], [0.2, 0.7, 0.9, 0.2, 0.8, 0.2, 8.0, 0.5, 0, 0],
                         [0.9, 0.9, 0.5, 0.2, 0.8, 0.0, 1.0, 0, 0.7, 1.3]], F)
    n = T + Q + P
    arrays = {"tri_points": tri, "tri_normal": synth.face_normals(tri) if T else np.zeros((0, 3), F),
              "tri_material": (np.arange(T) % len(mats)).astype(np.int32), "tri_order": np.arange(T, dtype=np.int32),
              "quad_coef": quads, "quad_material": ((np.arange(Q) + 1) % len(mats)).astype(np.int32),
              "quad_order": np.arange(T, T + Q, dtype=np.int32),
              "plane_coef": planes, "plane_material": ((np.arange(P) + 2) % len(mats)).astype(np.int32),
              "plane_order": np.arange(T + Q, n, dtype=np.int32),
              "materials": np.asarray(mats, F).reshape(-1, 10) if n else np.zeros((0, 10), F),
              "lights": np.asarray(lights, F).reshape(-1, 7)}
    return rt.PackedScene(arrays, dict(CAMERA))


def _all_paths_equal_oracle(rt, renderer, port, ps, W, H, exact_bits=True, **kw):
    want, _ = port.render(ps, rt.make_frame(ps.frame, W, H, **kw), threads=2)
    want8, _ = port.quantise(want)
    renderer.set_scene(ps)
    renderer.build_bvh()
    out = []
    for flags in (0, rt.F_FORCE_STREAM, rt.F_USE_BVH):
        got, got8, st = renderer.render(rt.make_frame(ps.frame, W, H, flags=flags, **kw))
        assert got.shape == (H, W, 3) and got8.shape == (H, W, 4)
        if exact_bits:
            assert bits_equal(got, want), flags
        else:                                     # powf: <= 4 ulp, NaN positions must agree
            assert np.array_equal(np.isnan(got), np.isnan(want)), flags
            assert np.allclose(np.nan_to_num(got), np.nan_to_num(want), rtol=0, atol=4e-6), flags
        assert np.array_equal(got8[..., :3], want8[..., :3]), flags
        out.append((got, got8, st))
    return out


LIGHT = [[100.0, 400.0, 200.0, 1.0, 1.0, 1.0, 0.9]]
SPHERE = [1.0 / 900, 1.0 / 900, 1.0 / 900, 0, 0, 0, 0, -2 * 40.0 / 900, 0, 40.0 * 40.0 / 900 - 1.0]   # r = 30 at (0, 40, 0)
GROUND = [0.0, 1.0, 0.0, 5.0]


def test_empty_scene_is_background(rt, renderer, port):
    ps = _scene(rt)
    for got, got8, st in _all_paths_equal_oracle(rt, renderer, port, ps, 33, 17):
        assert np.all(got == np.asarray(CAMERA["background"], F)) and st.rays_shadow == 0
        assert np.all(got8[..., :3] == np.array([26, 51, 153], np.uint8))          # floor(c * 255 + 0.5)


def test_no_lights_is_ambient_only(rt, renderer, port):
    ps = _scene(rt, quads=[SPHERE], planes=[GROUND], lights=[])
    for _, _, st in _all_paths_equal_oracle(rt, renderer, port, ps, 64, 40):
        assert st.rays_shadow == 0


@pytest.mark.parametrize("W,H", [(1, 1), (1, 7), (17, 1), (2, 3), (31, 33)])
def test_tiny_and_odd_images(rt, renderer, port, W, H):
    tri = [[-80, 0, -40, 80, 0, -40, 0, 120, -60]]
    ps = _scene(rt, tri=tri, quads=[SPHERE], planes=[GROUND], lights=LIGHT)
    _all_paths_equal_oracle(rt, renderer, port, ps, W, H, exact_bits=False)


def test_degenerate_triangles_never_hit(rt, renderer, port):
    tri = [[0, 10, 0, 0, 10, 0, 0, 10, 0],            # a point
           [-50, 20, 0, 50, 20, 0, 0, 20, 0],         # collinear
           [-80, 0, -40, 80, 0, -40, 0, 120, -60]]    # a real one behind them
    ps = _scene(rt, tri=tri, planes=[GROUND], lights=LIGHT)
    renderer.set_scene(ps)
    org = np.tile(np.asarray(CAMERA["cam_pos"], F), (4096, 1))
    rng = np.random.default_rng(0)
    d = (rng.uniform([-60, 0, 0], [60, 60, 0], (4096, 3)) - org).astype(F)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    idx, _, _ = renderer.trace_nearest(org, d)
    assert not np.isin(idx, [0, 1]).any() and (idx == 2).any()
    _all_paths_equal_oracle(rt, renderer, port, ps, 48, 32, exact_bits=False)


def test_depth_zero_ignores_mirrors(rt, renderer, port):
    ps = _scene(rt, quads=[SPHERE], planes=[GROUND], lights=LIGHT)
    for _, _, st in _all_paths_equal_oracle(rt, renderer, port, ps, 64, 48, exact_bits=False, max_bounces=0):
        assert st.rays_reflect == 0 and st.rays_refract == 0 and st.depth_levels == 1
    for _, _, st in _all_paths_equal_oracle(rt, renderer, port, ps, 64, 48, exact_bits=False, max_bounces=2):
        assert st.rays_reflect > 0


def test_non_finite_coefficients_give_counted_nan_pixels(rt, renderer, port):
    bad = list(SPHERE)
    bad[9] = float("nan")                                                           # K = NaN: every test against it is NaN
    ps = _scene(rt, quads=[bad], planes=[GROUND], lights=LIGHT)
    _all_paths_equal_oracle(rt, renderer, port, ps, 40, 30, exact_bits=False)
    inf_light = [[float("inf"), 400.0, 200.0, 1.0, 1.0, 1.0, 0.9]]
    ps = _scene(rt, quads=[SPHERE], planes=[GROUND], lights=inf_light)
    outs = _all_paths_equal_oracle(rt, renderer, port, ps, 40, 30, exact_bits=False)
    want_nan = int(np.isnan(outs[0][0]).sum())
    for got, got8, st in outs:
        assert st.nan_pixels == want_nan
        assert np.all(got8[..., :3][np.isnan(got)] == 0)


def test_supersampling_on_a_tiny_odd_image(rt, renderer, port):
    ps = _scene(rt, quads=[SPHERE], planes=[GROUND], lights=LIGHT)
    _all_paths_equal_oracle(rt, renderer, port, ps, 5, 3, exact_bits=False, spp_grid=3)
