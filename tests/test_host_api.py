"""The fresh C++ host scene API (parser, camera, Pretraitement, SoA packer) against the
reference: post-Pretraitement coefficients, materials, lights and camera basis of all nine
bundled scenes must equal -- bit for bit -- what the reference's own classes produce
(tests/golden/packed/*.npz, dumped through the reference's getters)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, REF_DIR, bits_equal, golden_image, packed_scene

KEYS = ("tri_points", "tri_normal", "quad_coef", "plane_coef", "materials", "lights")
IKEYS = ("tri_material", "tri_order", "quad_material", "quad_order", "plane_material", "plane_order")


@pytest.fixture(scope="module")
def host(rt):
    from importlib import import_module
    hb = import_module("rtb200.hostbind")
    yield hb
    hb.free()


def _same_scene(ps, want, f, W, H):
    for k in KEYS:
        assert ps[k].shape == want[k].shape, k
        assert bits_equal(ps[k], want[k]), k
    for k in IKEYS:
        assert np.array_equal(ps[k], want[k]), k
    for k in ("cam_pos", "cam_u", "cam_v", "cam_n", "background"):
        assert bits_equal(ps.frame[k], want.frame[k]), k
    assert np.float32(ps.frame["half_h"]) == np.float32(want.frame["half_h"])
    assert np.float32(f.half_w) == (np.float32(W) / np.float32(H)) * np.float32(want.frame["half_h"])
    assert f.max_bounces == 20 and np.float32(f.min_energy) == np.float32(0.01) and f.scene_ior == 1.0


@pytest.mark.parametrize("n", range(1, 10))
def test_parser_and_pretraitement_match_reference(rt, host, n):
    ps, f = host.load_and_pack(os.path.join(GOLDEN, "scenes", f"scene{n}.dat"), 160, 120)
    _same_scene(ps, packed_scene(rt, n), f, 160, 120)


@pytest.mark.skipif(not os.path.exists(os.path.join(REF_DIR, "Scenes")), reason="reference not present")
@pytest.mark.parametrize("n", range(1, 10))
def test_original_scene_files_parse_identically(rt, host, n):
    """The unstripped bundled files (comments, indented '* ---' lines, 'instance:' keys)."""
    ps, f = host.load_and_pack(os.path.join(REF_DIR, "Scenes", f"scene{n}.dat"), 640, 480)
    _same_scene(ps, packed_scene(rt, n), f, 640, 480)


def test_parser_hardening(rt, host, tmp_path):
    """CRLF line endings and lines longer than the reference's 80-byte buffer."""
    src = open(os.path.join(GOLDEN, "scenes", "scene1.dat")).read()
    long_comment = "* " + "x" * 200 + "\n"
    p = tmp_path / "crlf.dat"
    p.write_bytes((long_comment + src).replace("\n", "\r\n").encode())
    ps, f = host.load_and_pack(str(p), 160, 120)
    _same_scene(ps, packed_scene(rt, 1), f, 160, 120)
    with pytest.raises(FileNotFoundError):
        host.load_and_pack(str(tmp_path / "missing.dat"), 16, 16)


def test_host_pack_feeds_oracle_to_golden_frame(rt, host, port):
    """host classes -> packed scene -> C oracle == golden frame of the reference renderer."""
    for n in (3, 9):
        ps, _ = host.load_and_pack(os.path.join(GOLDEN, "scenes", f"scene{n}.dat"), 160, 120)
        fr = rt.make_frame(ps.frame, 160, 120, max_bounces=4)
        out, _ = port.render(ps, fr, threads=4)
        assert bits_equal(out, golden_image(n, 4))


def test_cli_fails_loudly_without_gpu(rt, host):
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    r = subprocess.run([host.CLI_PATH, os.path.join(GOLDEN, "scenes", "scene1.dat"), "-x", "64", "-y", "48"],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 7, 9])
def test_dropin_lancer_rayons_on_gpu(rt, host, n):
    """CScene::TraiterFichierDeScene + LancerRayons (the reference's call sequence, Main.cpp:87-181)
    through CUDA == golden frame of the reference CPU renderer, bit for bit."""
    host.lib().rtscene_load(os.fsencode(os.path.join(GOLDEN, "scenes", f"scene{n}.dat")), 4)
    f, q, st = host.render_loaded(160, 120)
    assert bits_equal(f, golden_image(n, 4))
    assert st.rays_primary == 160 * 120 and q[..., 3].min() == 255


@pytest.mark.gpu
def test_cli_renders_ppm(rt, host, tmp_path):
    out = tmp_path / "s7.ppm"
    r = subprocess.run([host.CLI_PATH, os.path.join(GOLDEN, "scenes", "scene7.dat"), "-x", "160", "-y", "120",
                        "--depth", "4", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    data = out.read_bytes()
    assert data.startswith(b"P6\n160 120\n255\n")
    img = np.frombuffer(data[len(b"P6\n160 120\n255\n"):], np.uint8).reshape(120, 160, 3)[::-1]
    want = golden_image(7, 4)
    q = np.floor(np.clip(want, 0, 1) * np.float32(255) + np.float32(0.5)).astype(np.uint8)
    assert np.array_equal(img, q)


@pytest.mark.gpu
def test_cli_gpus_option_assembles_the_same_image(rt, host, tmp_path):
    """rt_cli --gpus g: one context + thread per GPU, band sharding without a communicator, every context writes its
    bands into the host frame (RTB200_F_HOST_BANDS).  With fewer devices than g the contexts share devices: the image
    must be the one --gpus 1 writes, byte for byte (ragged height: 250 rows = 15 bands of 16 + 10 rows)."""
    scene = os.path.join(GOLDEN, "scenes", "scene9.dat")
    imgs = []
    for g in (1, 3, 8):
        out = tmp_path / f"s9_{g}.ppm"
        r = subprocess.run([host.CLI_PATH, scene, "-x", "333", "-y", "250", "--depth", "4", "--gpus", str(g), "--out", str(out)],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        imgs.append(out.read_bytes())
    assert imgs[0].startswith(b"P6\n333 250\n255\n") and len(imgs[0]) > 333 * 250 * 3
    assert imgs[1] == imgs[0] and imgs[2] == imgs[0]
    r = subprocess.run([host.CLI_PATH, scene, "--gpus", "2", "--orbit", "3"], capture_output=True, text=True)
    assert r.returncode != 0


def _write_obj(path, nq=12, size=60.0, y0=5.0):
    """A bumpy nq x nq quad grid (two conventions of face records, one 4-gon per quad)."""
    xs = np.linspace(-size, size, nq + 1)
    lines = ["# test mesh"]
    for j in range(nq + 1):
        for i in range(nq + 1):
            lines.append(f"v {xs[i]:.4f} {y0 + 6.0 * np.sin(0.7 * i) * np.cos(0.9 * j):.4f} {xs[j]:.4f}")
    for j in range(nq):
        for i in range(nq):
            a = j * (nq + 1) + i + 1
            b, c, d = a + 1, a + nq + 2, a + nq + 1
            if (i + j) % 2:
                lines.append(f"f {a}/1/1 {d}/2/1 {c}/3/1 {b}/4/1")        # quad, v/vt/vn form
            else:
                lines.append(f"f {a - (nq + 1) ** 2 - 1} {d}//7 {c} {b}")   # relative index, v//vn form
    path.write_text("\n".join(lines) + "\n")
    return 2 * nq * nq


def test_obj_loader_appends_triangles_in_file_order(rt, host, tmp_path):
    """CScene::ChargerMaillageOBJ: fan-triangulated faces appended through AjouterSurface after the
    scene file's own surfaces, transformed by the prototype's matrix in Pretraitement."""
    obj = tmp_path / "grid.obj"
    n_tri = _write_obj(obj)
    scene = os.path.join(GOLDEN, "scenes", "scene1.dat")
    base, _ = host.load_and_pack(scene, 64, 48)
    host.lib().rtscene_load(os.fsencode(scene), 2)
    assert host.add_obj(str(obj), rgb=(0.9, 0.2, 0.1), reflect=0.3, scale=2.0, translate=(10.0, 0.0, -5.0)) == n_tri
    import ctypes as C
    s, f = rt.abi.Scene(), rt.abi.Frame()
    host.lib().rtscene_pack(64, 48, C.byref(s), C.byref(f))
    T0 = len(base["tri_order"])
    assert s.n_triangles == T0 + n_tri and s.n_quadrics == len(base["quad_order"])
    pts = np.ctypeslib.as_array(s.tri_points, shape=(s.n_triangles * 9,)).reshape(-1, 3, 3)
    order = np.ctypeslib.as_array(s.tri_order, shape=(s.n_triangles,))
    n_base = T0 + len(base["quad_order"]) + len(base["plane_order"])
    assert list(order[T0:]) == list(range(n_base, n_base + n_tri))      # appended after every scene-file surface
    # first face: quad (v1, v14, v15, v2) fanned -> (v1, v14, v15), scaled by 2 then translated
    xs = np.linspace(-60.0, 60.0, 13)
    v1 = np.array([xs[0], 5.0, xs[0]], np.float32) * np.float32(2.0) + np.array([10.0, 0.0, -5.0], np.float32)
    assert np.allclose(pts[T0, 0], v1, atol=1e-3)
    with pytest.raises(FileNotFoundError):
        host.add_obj(str(tmp_path / "missing.obj"))


@pytest.mark.gpu
def test_obj_mesh_renders_through_lbvh_like_the_oracle(rt, host, port, tmp_path):
    """scene file + OBJ mesh -> CScene::LancerRayons with RTB200_F_USE_BVH == the oracle's frame."""
    obj = tmp_path / "grid.obj"
    _write_obj(obj, nq=24)
    scene = os.path.join(GOLDEN, "scenes", "scene7.dat")
    host.lib().rtscene_load(os.fsencode(scene), 3)
    host.add_obj(str(obj), rgb=(0.2, 0.7, 0.9), reflect=0.4, scale=1.5, translate=(0.0, -20.0, -60.0))
    f, q, st = host.render_loaded(200, 150, flags=rt.F_USE_BVH)
    assert st.path == 2 and st.bvh_nodes_visited > 0
    # the same scene through the packer into the CPU oracle
    import ctypes as C
    s, fr = rt.abi.Scene(), rt.abi.Frame()
    host.lib().rtscene_pack(200, 150, C.byref(s), C.byref(fr))
    want = np.empty((150, 200, 3), np.float32)
    port.lib.rto_render(C.byref(s), C.byref(fr), want, port.max_threads)
    assert np.allclose(f, want, rtol=0, atol=4e-6)
    assert (np.abs(f - want) > 0).mean() < 0.01


def _vec(a):
    return np.array(a[:], np.float32)


def test_orbit_camera_blocks(rt, host):
    """CScene::OrbiteCamera / CadreDepuis: frame 0 is the scene's own camera bit for bit, every other
    camera keeps height, distance to the point looked at and field of view, and its basis is the one
    InitialiserCamera derives (orthonormal, N pointing from the look-at point to the camera)."""
    path = os.path.join(GOLDEN, "scenes", "scene7.dat")
    ps, own = host.load_and_pack(path, 96, 64)
    frames = host.orbit_frames(96, 64, 8)
    assert bytes(frames[0]) == bytes(own)
    p0 = _vec(own.cam_pos).astype(np.float64)
    # the point looked at lies on the line p0 - t * N
    n0 = _vec(own.cam_n).astype(np.float64)
    for k, f in enumerate(frames):
        u, v, n = (_vec(x).astype(np.float64) for x in (f.cam_u, f.cam_v, f.cam_n))
        for a in (u, v, n):
            assert abs(np.dot(a, a) - 1.0) < 1e-5
        assert abs(np.dot(u, v)) < 1e-5 and abs(np.dot(u, n)) < 1e-5 and abs(np.dot(v, n)) < 1e-5
        assert np.allclose(np.cross(v, n), u, atol=1e-5)
        assert f.half_h == pytest.approx(own.half_h, rel=1e-5) and f.half_w == pytest.approx(own.half_w, rel=1e-5)
        assert (f.width, f.height, f.max_bounces) == (96, 64, own.max_bounces)
        assert _vec(f.cam_pos)[1] == pytest.approx(p0[1], abs=1e-3)       # the circle is horizontal
        assert abs(n[1] - n0[1]) < 1e-5                                    # same elevation angle
    assert len({tuple(_vec(f.cam_pos)) for f in frames}) == 8


@pytest.mark.gpu
@pytest.mark.parametrize("n,flags_name", [(1, None), (4, None), (7, None), (4, "F_USE_BVH")])
def test_camera_path_equals_single_frames(rt, host, renderer, n, flags_name):
    """CScene::LancerTrajectoire (one pipelined rtb200_render_sequence: depth 0 keeps the fast path;
    depth 4 / LBVH take the frame-by-frame form) == the same cameras rendered one by one."""
    flags = getattr(rt, flags_name) if flags_name else 0
    path = os.path.join(GOLDEN, "scenes", f"scene{n}.dat")
    for depth in (0, 4):
        ps, _ = host.load_and_pack(path, 160, 120, depth)
        q, frames, stats = host.orbit_loaded(160, 120, 6, flags)
        renderer.set_scene(ps)
        if flags:
            renderer.build_bvh()
        for i, fr in enumerate(frames):
            assert fr.flags == flags and fr.max_bounces == depth
            _, want8, st = renderer.render(fr)
            assert np.array_equal(q[i], want8), (depth, i)
            assert stats[i].rays_primary == 160 * 120 and stats[i].rays_shadow == st.rays_shadow
        assert not np.array_equal(q[0], q[3])                             # the camera did move


@pytest.mark.gpu
def test_cli_orbit_writes_numbered_ppm(rt, host, renderer, tmp_path):
    out = tmp_path / "fly.ppm"
    path = os.path.join(GOLDEN, "scenes", "scene1.dat")
    r = subprocess.run([host.CLI_PATH, path, "-x", "96", "-y", "64", "--depth", "0", "--orbit", "3", "--out", str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    ps, _ = host.load_and_pack(path, 96, 64, 0)
    frames = host.orbit_frames(96, 64, 3)
    renderer.set_scene(ps)
    head = b"P6\n96 64\n255\n"
    for i, fr in enumerate(frames):
        data = (tmp_path / f"fly_{i:03d}.ppm").read_bytes()
        assert data.startswith(head)
        img = np.frombuffer(data[len(head):], np.uint8).reshape(64, 96, 3)[::-1]
        _, want8, _ = renderer.render(fr)
        assert np.array_equal(img, want8[..., :3]), i
    assert not (tmp_path / "fly_003.ppm").exists() and not out.exists()
