"""GPU (-m gpu): rtb200_render_sequence (a camera path as one pipelined submission) must deliver
exactly the frames rtb200_render delivers one by one -- pipelined form (scene1: no secondary rays)
and frame-by-frame fallback (scene7: mirrors + glass)."""
import numpy as np
import pytest

from conftest import packed_scene

pytestmark = pytest.mark.gpu


def _path(rt, ps, n, W, H, depth, flags=0):
    frames = []
    for i in range(n):
        fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, flags=flags)
        fr.cam_pos[0] += 3.0 * i                      # the camera slides sideways; basis unchanged
        fr.cam_pos[1] += 1.5 * i
        frames.append(fr)
    return frames


@pytest.mark.parametrize("scene,depth", [(1, 2), (4, 2), (7, 3)])
def test_sequence_equals_single_frames(rt, renderer, scene, depth):
    ps = packed_scene(rt, scene)
    renderer.set_scene(ps)
    W, H, n = 320, 200, 7
    frames = _path(rt, ps, n, W, H, depth)
    want, stats1 = [], []
    for fr in frames:
        _, q, st = renderer.render(fr, want_f32=False)
        want.append(q.copy()); stats1.append(st)
    out = rt.pinned_empty((n, H, W, 4), np.uint8)
    out[:] = 0
    got, stats = renderer.render_sequence(frames, out_rgba8=out)
    for i in range(n):
        assert np.array_equal(got[i], want[i]), i
        assert stats[i].rays_primary == W * H and stats[i].rays_shadow == stats1[i].rays_shadow
        if scene != 7:      # beyond level 0 the queue order (atomics) decides which rays share a warp's early exit
            assert stats[i].tests_quadric == stats1[i].tests_quadric
    assert not np.array_equal(want[0], want[-1])     # the camera really moved
    # without a host destination the last frame stays on the device
    renderer.render_sequence(frames, out_rgba8=None)
    if scene != 7:
        _, q = renderer.read_frame(H, W, want_f32=False)
        assert np.array_equal(q, want[-1])


def test_sequence_rejects_mixed_sizes(rt, renderer):
    ps = packed_scene(rt, 1)
    renderer.set_scene(ps)
    a = rt.make_frame(ps.frame, 64, 48, max_bounces=2)
    b = rt.make_frame(ps.frame, 64, 50, max_bounces=2)
    with pytest.raises(rt.RtError):
        renderer.render_sequence([a, b])
