"""GPU (-m gpu), collected LAST on purpose: rtb200_trace_nearest_static -- the
nearest-hit ray query with its scratch in a static __device__ array, for callers that may not call cudaMalloc -- must
return what rtb200_trace_nearest returns (bit for bit: same kernel, same rounding), on two calls in a row (the second
one finds the first one's contents in the shared array), through the C ABI.  No counterpart in the
reference."""
import numpy as np
import pytest

from conftest import bits_equal, packed_scene

pytestmark = pytest.mark.gpu


def _rays(rt, ps, n, seed):
    rng = np.random.default_rng(seed)
    cam = np.asarray(ps.frame["cam_pos"], np.float32)
    o = (cam + rng.uniform(-20, 20, (n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d[:, 2] -= 1.5                                    # mostly towards the scene
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return o, d


def test_static_scratch_query_equals_the_allocating_one(rt):
    r = rt.Renderer(0)                                # a context of its own: independent of what ran before
    try:
        assert rt.lib().rtb200_abi_version() == 1
        ps = packed_scene(rt, 9)                      # planes + quadrics + triangles, mirrors and glass
        r.set_scene(ps)
        # two calls in a row, the second smaller than the first: stale entries of call 1 lie behind call 2's
        for n, seed in ((50000, 1), (1237, 2)):
            o, d = _rays(rt, ps, n, seed)
            want_i, want_t, want_n = r.trace_nearest(o, d)
            got_i, got_t, got_n = r.trace_nearest_static(o, d)
            assert np.array_equal(got_i, want_i)
            assert bits_equal(got_t, want_t) and bits_equal(got_n, want_n)
            assert (want_i >= 0).mean() > 0.2 and (want_i < 0).any()
        # another scene in the same process: the array is shared, the scene is the context's
        ps = packed_scene(rt, 5)
        r.set_scene(ps)
        o, d = _rays(rt, ps, 4096, 3)
        for flags in (0, rt.F_EXACT_FP, rt.F_FUSED_FP):
            want = r.trace_nearest(o, d, flags=flags)
            got = r.trace_nearest_static(o, d, flags=flags)
            assert np.array_equal(got[0], want[0]) and bits_equal(got[1], want[1]) and bits_equal(got[2], want[2])
        # limits: more rays than the static array holds, and paths whose kernels keep scratch of their own
        o, d = _rays(rt, ps, 65537, 4)
        with pytest.raises(rt.RtError) as e:
            r.trace_nearest_static(o, d)
        assert "RTB200_STATIC_QUERY_MAX" in str(e.value)
        with pytest.raises(rt.RtError):
            r.trace_nearest_static(o[:16], d[:16], flags=rt.F_FORCE_STREAM)
        gi, gt, gn = r.trace_nearest_static(o[:65536], d[:65536])            # exactly the maximum
        wi, wt, wn = r.trace_nearest(o[:65536], d[:65536])
        assert np.array_equal(gi, wi) and bits_equal(gt, wt) and bits_equal(gn, wn)
        assert r.trace_nearest_static(o[:0], d[:0])[0].shape == (0,)
    finally:
        r.close()
