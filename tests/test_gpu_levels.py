"""GPU (-m gpu): the level loop without host round trips (resident path with secondary rays: every
level kernel reads its input count from the device, queues sized by a guess) must produce exactly
the frame of the synchronous level loop (RTB200_SYNC_LEVELS=1) -- on a first frame (no history), on
a repeated frame (guess from history) and when the guess is too small (frame redone, flagged)."""
import os

import numpy as np
import pytest

from conftest import bits_equal, packed_scene

pytestmark = pytest.mark.gpu


def _sync_frame(renderer, fr):
    os.environ["RTB200_SYNC_LEVELS"] = "1"
    try:
        return renderer.render(fr)
    finally:
        del os.environ["RTB200_SYNC_LEVELS"]


@pytest.mark.parametrize("scene", [5, 6, 7, 8, 9])
def test_async_levels_equal_sync_levels(rt, renderer, scene):
    ps = packed_scene(rt, scene)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 320, 200)                 # depth / energy cut-off of the scene file
    want, want8, st_s = _sync_frame(renderer, fr)
    assert st_s.depth_levels > 1 and st_s.rays_reflect + st_s.rays_refract > 0
    renderer.set_scene(ps)                                  # forget the history: first-frame guess
    got1, got18, st1 = renderer.render(fr)
    got2, got28, st2 = renderer.render(fr)                  # guess from the frame before
    for got, got8, st in ((got1, got18, st1), (got2, got28, st2)):
        assert bits_equal(got, want) and np.array_equal(got8, want8)
        assert (st.rays_shadow, st.rays_reflect, st.rays_refract) == (st_s.rays_shadow, st_s.rays_reflect, st_s.rays_refract)
        assert st.depth_levels == st_s.depth_levels
    assert st2.level_loop_redone == 0
    assert st2.kernel_launches <= st1.kernel_launches       # the history trims the levels that stay empty


def test_async_levels_supersampled_and_chunked(rt, renderer):
    ps = packed_scene(rt, 9)
    renderer.set_scene(ps)
    fr = rt.make_frame(ps.frame, 96, 64, max_bounces=6, spp_grid=3)
    want, want8, _ = _sync_frame(renderer, fr)
    renderer.set_scene(ps)
    for _ in range(2):
        got, got8, st = renderer.render(fr)
        assert bits_equal(got, want) and np.array_equal(got8, want8)


def test_too_small_guess_is_redone_and_flagged(rt, renderer):
    """History from a frame that sees (almost) nothing, then the camera turns towards the mirrors:
    the guessed queues overflow, the frame is redone with exact capacities and says so."""
    ps = packed_scene(rt, 7)
    renderer.set_scene(ps)
    W, H = 320, 200
    away = rt.make_frame(ps.frame, W, H)
    for i in range(3):
        away.cam_n[i] = -away.cam_n[i]                      # look the other way (basis stays orthonormal)
        away.cam_u[i] = -away.cam_u[i]
    toward = rt.make_frame(ps.frame, W, H)
    want, want8, st_s = _sync_frame(renderer, toward)
    renderer.set_scene(ps)
    _, _, st_a = renderer.render(away)
    got, got8, st = renderer.render(toward)
    if st_a.rays_reflect + st_a.rays_refract < 4096 < st_s.rays_reflect + st_s.rays_refract:
        assert st.level_loop_redone == 1
    assert bits_equal(got, want) and np.array_equal(got8, want8)
    assert (st.rays_shadow, st.rays_reflect, st.rays_refract) == (st_s.rays_shadow, st_s.rays_reflect, st_s.rays_refract)
    # the redone frame left a history: the next one is quick again and still right
    got, got8, st = renderer.render(toward)
    assert st.level_loop_redone == 0
    assert bits_equal(got, want) and np.array_equal(got8, want8)
