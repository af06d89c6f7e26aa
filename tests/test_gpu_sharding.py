"""GPU (-m gpu): band sharding WITHOUT a communicator (rtb200_set_sharding) and RTB200_F_HOST_BANDS -- every
"rank" writes its own interleaved 16-row bands into the caller's full host frame.  Several contexts on ONE
device stand in for the GPUs of a box (the band arithmetic and the strided copies are the same); the frame they
assemble must be the frame one context renders, byte for byte."""
from importlib import import_module

import numpy as np
import pytest

from conftest import bits_equal, packed_scene

pytestmark = pytest.mark.gpu


def _assemble(rt, ps, fr_flags, W, H, depth, n_ranks, spp=1, bvh=False, pinned=True):
    out8 = rt.pinned_empty((H, W, 4), np.uint8) if pinned else np.empty((H, W, 4), np.uint8)
    outf = np.empty((H, W, 3), np.float32)
    out8[:] = 7
    outf[:] = -1.0
    rows = 0
    for rank in range(n_ranks):
        r = rt.Renderer(0)
        r.set_sharding(rank, n_ranks)
        r.set_scene(ps)
        if bvh:
            r.build_bvh()
        fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp, flags=fr_flags | rt.F_HOST_BANDS)
        _, _, st = r.render(fr, out_f32=outf, out_rgba8=out8)
        rows += st.rays_primary // (W * spp * spp)
        r.close()
    assert rows == H
    return outf, out8


@pytest.mark.parametrize("scene,depth,W,H,n_ranks", [(1, 2, 640, 480, 2), (7, 4, 333, 250, 3), (9, 3, 1280, 1000, 4), (5, 2, 200, 37, 8)])
def test_host_bands_resident(rt, renderer, scene, depth, W, H, n_ranks):
    ps = packed_scene(rt, scene)
    renderer.set_scene(ps)
    want, want8, _ = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=depth))
    got, got8 = _assemble(rt, ps, 0, W, H, depth, n_ranks)
    assert bits_equal(got, want) and np.array_equal(got8, want8)


def test_host_bands_streamed_and_lbvh(rt, renderer):
    synth = import_module("rtb200.synth")
    ps = synth.soup(700, 12, 3, seed=3)
    renderer.set_scene(ps)
    W, H = 1024, 600                                 # above the size where the read-back is overlapped chunk by chunk
    want, want8, _ = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=2, flags=rt.F_FORCE_STREAM))
    got, got8 = _assemble(rt, ps, rt.F_FORCE_STREAM, W, H, 2, 4)
    assert bits_equal(got, want) and np.array_equal(got8, want8)
    hf = synth.heightfield(20000, 300.0, 3, seed=9)
    renderer.set_scene(hf)
    renderer.build_bvh()
    want, want8, _ = renderer.render(rt.make_frame(hf.frame, 512, 300, max_bounces=2, spp_grid=2, flags=rt.F_USE_BVH))
    got, got8 = _assemble(rt, hf, rt.F_USE_BVH, 512, 300, 2, 2, spp=2, bvh=True, pinned=False)
    assert bits_equal(got, want) and np.array_equal(got8, want8)


def test_sharding_without_flag_gives_compact_rows(rt, renderer):
    ps = packed_scene(rt, 4)
    renderer.set_scene(ps)
    W, H = 320, 100
    want, want8, _ = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=2))
    r = rt.Renderer(0)
    r.set_sharding(1, 3)
    r.set_scene(ps)
    rows = rt.band_map(H, 16, 3, 1)
    out8 = np.zeros((H, W, 4), np.uint8)
    r.render(rt.make_frame(ps.frame, W, H, max_bounces=2), want_f32=False, out_rgba8=out8)
    assert np.array_equal(out8[:len(rows)], want8[rows])
    with pytest.raises(rt.RtError):
        r.set_sharding(3, 3)
    r.close()


def test_max_bounces_is_bounded(rt, renderer):
    ps = packed_scene(rt, 7)
    renderer.set_scene(ps)
    with pytest.raises(rt.RtError):
        renderer.render(rt.make_frame(ps.frame, 64, 48, max_bounces=1000000))
