"""GPU (-m gpu): randomised scenes through EVERY path -- scene-resident kernels, plain tiled scan (exact mode), culled scan
(default streamed path), LBVH -- against the oracle: triangles, spheres / ellipsoids / cylinders and planes in SHUFFLED
insertion order, opaque / mirror / glass / transparent materials, 1-5 lights, cameras inside and outside the cloud,
with and without supersampling.  Everything geometry decides (ray counts, float frames up to the powf term) must agree."""
from importlib import import_module

import numpy as np
import pytest

from conftest import bits_equal

pytestmark = pytest.mark.gpu
F = np.float32


def _scene(rt, synth, hostbind, seed):
    rng = np.random.default_rng(seed)
    T, Q, P = int(rng.integers(20, 400)), int(rng.integers(0, 12)), int(rng.integers(1, 3))
    ext = float(rng.choice([60.0, 250.0]))
    c = rng.uniform(-ext, ext, (T, 3)); c[:, 1] = rng.uniform(0, ext, T)
    tri = (c[:, None, :] + rng.uniform(2, 0.25 * ext) * rng.uniform(-1, 1, (T, 3, 3))).reshape(T, 9).astype(F)
    quads = []
    for _ in range(Q):
        ctr = rng.uniform(-ext, ext, 3); ctr[1] = rng.uniform(5, ext)
        a = rng.uniform(0.05 * ext, 0.3 * ext, 3)
        q = 1.0 / (a * a)
        if rng.uniform() < 0.25:
            q[int(rng.integers(0, 3))] = 0.0                      # a cylinder along one axis (unbounded: side list)
        lin = -2.0 * ctr * q
        quads.append([q[0], q[1], q[2], 0, 0, 0, lin[0], lin[1], lin[2], float(np.sum(ctr * ctr * q) - 1.0)])
    quads = np.asarray(quads, F).reshape(Q, 10)
    planes = [[0.0, 1.0, 0.0, 5.0]] + ([[1.0, 0.0, 0.2, 2.5 * ext]] if P > 1 else [])
    planes = np.asarray(planes, F)
    N = T + Q + P
    order = rng.permutation(N)
    kinds = np.concatenate([np.zeros(T, int), np.ones(Q, int), np.full(P, 2)])
    slot = np.empty(N, np.int64); slot[order] = np.arange(N)      # surface j (type-major numbering) sits at insertion index slot[j]
    mats = np.zeros((N, 10), F)
    mats[:, 0:3] = rng.uniform(0.2, 1.0, (N, 3)); mats[:, 3] = 0.2; mats[:, 4] = 0.8
    mats[:, 5], mats[:, 6] = rng.choice([0.0, 0.4], N), 20.0
    kind = rng.uniform(0, 1, N)
    mats[kind < 0.2, 7] = 0.5                                                   # mirrors
    glass = (kind >= 0.2) & (kind < 0.4)
    mats[glass, 8] = rng.uniform(0.4, 0.9, int(glass.sum())); mats[glass, 9] = rng.choice([1.0, 1.3, 1.5], int(glass.sum()))

    def pick(k):
        idx = np.nonzero(kinds == k)[0]
        o = slot[idx]
        s = np.argsort(o)
        return idx[s], o[s].astype(np.int32)
    ti, to = pick(0); qi, qo = pick(1); pi, po = pick(2)
    arrays = {"tri_points": tri[ti], "tri_normal": synth.face_normals(tri[ti]) if T else np.zeros((0, 3), F),
              "tri_material": ti.astype(np.int32), "tri_order": to,
              "quad_coef": quads[qi - T] if Q else np.zeros((0, 10), F), "quad_material": qi.astype(np.int32), "quad_order": qo,
              "plane_coef": planes[pi - T - Q], "plane_material": pi.astype(np.int32), "plane_order": po,
              "materials": mats, "lights": np.zeros((0, 7), F)}
    L = int(rng.integers(1, 6))
    lights = np.zeros((L, 7), F)
    lights[:, 0:3] = rng.uniform(-ext, ext, (L, 3)); lights[:, 1] = rng.uniform(ext, 3 * ext, L)
    lights[:, 3:6] = rng.uniform(0.5, 1.0, (L, 3)); lights[:, 6] = 1.2 / L
    arrays["lights"] = lights
    cam = rng.uniform(-ext, ext, 3); cam[1] = rng.uniform(10, 1.5 * ext)
    if rng.uniform() < 0.5:
        cam *= 2.5
    frame = hostbind.camera_frame(tuple(cam), (0.0, 0.3 * ext, 0.0), (0.0, 1.0, 0.0))
    frame["background"] = np.array([0.1, 0.1, 0.3], F)
    return rt.PackedScene(arrays, frame)


@pytest.mark.parametrize("seed", range(12))
def test_random_scene_all_paths_equal_the_oracle(rt, renderer, port, seed):
    synth = import_module("rtb200.synth")
    hostbind = import_module("rtb200.hostbind")
    ps = _scene(rt, synth, hostbind, 1000 + seed)
    ok, why = rt.validate_scene(ps)
    assert ok, why
    W, H, depth, spp = 112, 72, int(3 + seed % 3), 2 if seed % 4 == 0 else 1
    renderer.set_scene(ps)
    renderer.build_bvh()
    fr0 = rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp)
    lat, _ = port.render_region(ps, fr0, 0, 0, W * spp, H * spp, threads=port.max_threads)
    want = lat if spp == 1 else lat.reshape(H, spp, W, spp, 3).astype(np.float64).mean(axis=(1, 3)).astype(F)
    c = port.ray_counts()
    specular = bool((ps["materials"][:, 5] > 0).any())
    frames = {}
    for name, flags in (("resident", rt.F_FORCE_RESIDENT), ("culled scan", rt.F_FORCE_STREAM), ("plain scan, exact", rt.F_FORCE_STREAM | rt.F_EXACT_FP),
                        ("lbvh", rt.F_USE_BVH)):
        got, got8, st = renderer.render(rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=spp, flags=flags))
        assert (st.rays_primary, st.rays_shadow, st.rays_reflect, st.rays_refract) == (c["primary"], c["shadow"], c["reflect"], c["refract"]), name
        if specular or spp > 1:
            assert np.allclose(got, want, rtol=0, atol=8e-6), (name, float(np.abs(got - want).max()))
        else:
            assert bits_equal(got, want), (name, int((got.view(np.uint32) != want.view(np.uint32)).sum()))
        frames[name] = got
    # the four CUDA paths share powf: among themselves they agree bit for bit
    for name in ("culled scan", "plain scan, exact", "lbvh"):
        assert bits_equal(frames[name], frames["resident"]), name
