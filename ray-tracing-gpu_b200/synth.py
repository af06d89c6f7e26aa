"""Seeded synthetic scenes of the shapes BASELINE.json names.  Geometry lives at
bundled-scene scale (edges >> sqrt(EPSILON), extents 10^2..10^3): the reference's absolute
EPSILON = 0.01 makes triangles with |e1||e2| <~ 0.01 invisible to every ray.

The generators only produce INPUTS (numpy float32 arrays in the layout of include/rtb200.h);
face normals follow CTriangle::CalculerNormale (Triangle.cpp:199-204) in binary32."""
import numpy as np

from . import abi, hostbind

F = np.float32


def face_normals(tp):
    """normalise((p1-p0) x (p2-p0)) in binary32, one rounding per operation; ZERO if |n| <= 0.01."""
    p0, p1, p2 = tp[:, 0:3], tp[:, 3:6], tp[:, 6:9]
    e1, e2 = (p1 - p0).astype(F), (p2 - p0).astype(F)
    n = np.stack([e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1],
                  e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2],
                  e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]], axis=1).astype(F)
    ln = np.sqrt((n[:, 0] * n[:, 0] + n[:, 1] * n[:, 1]).astype(F) + n[:, 2] * n[:, 2]).astype(F)
    ok = ln > F(0.01)
    inv = np.where(ok, F(1.0) / np.where(ok, ln, F(1.0)), F(0.0)).astype(F)
    return (n * inv[:, None]).astype(F)


def _materials(rng, n, p_reflect, p_refract, specular=(0.4, 20.0), reflect=0.5, refract=(0.6, 1.3)):
    """A small palette of materials and a per-surface index (colour r g b | ka kd ks shin kr kt ior)."""
    n_pal = 64
    pal = np.zeros((n_pal, 10), F)
    pal[:, 0:3] = rng.uniform(0.15, 1.0, (n_pal, 3))
    pal[:, 3], pal[:, 4] = 0.2, 0.8
    pal[:, 5], pal[:, 6] = specular
    kind = rng.uniform(0, 1, n_pal)
    pal[kind < p_reflect, 7] = reflect
    sel = (kind >= p_reflect) & (kind < p_reflect + p_refract)
    pal[sel, 8], pal[sel, 9] = refract
    return pal, rng.integers(0, n_pal, n).astype(np.int32)


def _ellipsoid(c, a):
    """(x-cx)^2/ax^2 + ... = 1 as the 10 expanded coefficients A B C | yz xz xy | lx ly lz | K."""
    q = (1.0 / (a * a))
    lin = -2.0 * c * q
    K = float(np.sum(c * c * q) - 1.0)
    return np.array([q[0], q[1], q[2], 0, 0, 0, lin[0], lin[1], lin[2], K], F)


def _lights(rng, n, height=(400.0, 800.0), spread=500.0, intensity=0.35):
    L = np.zeros((n, 7), F)
    L[:, 0] = rng.uniform(-spread, spread, n)
    L[:, 1] = rng.uniform(height[0], height[1], n)
    L[:, 2] = rng.uniform(-spread, spread, n)
    L[:, 3:6] = 1.0
    L[:, 6] = intensity
    return L


def _finish(tri, tri_mat, quads, quad_mat, planes, plane_mat, mats, lights, camera, background=(0.0, 0.0, 0.6)):
    T, Q, P = len(tri), len(quads), len(planes)
    arrays = {
        "tri_points": tri, "tri_normal": face_normals(tri) if T else np.zeros((0, 3), F),
        "tri_material": tri_mat, "tri_order": np.arange(T, dtype=np.int32),
        "quad_coef": quads, "quad_material": quad_mat, "quad_order": np.arange(T, T + Q, dtype=np.int32),
        "plane_coef": planes, "plane_material": plane_mat, "plane_order": np.arange(T + Q, T + Q + P, dtype=np.int32),
        "materials": mats, "lights": lights,
    }
    frame = hostbind.camera_frame(*camera)
    frame["background"] = np.asarray(background, F)
    return abi.PackedScene(arrays, frame)


def soup(n_tris=10000, n_quads=64, n_lights=4, seed=1234, radius=12.0):
    """BASELINE config 3: triangle soup + ellipsoids + ground plane y = -45, camera (0,150,900)."""
    rng = np.random.default_rng(seed)
    lo, hi = np.array([-400.0, -40.0, -400.0]), np.array([400.0, 360.0, 400.0])
    c = rng.uniform(lo, hi, (n_tris, 3))
    tri = (c[:, None, :] + radius * rng.uniform(-1, 1, (n_tris, 3, 3))).reshape(n_tris, 9).astype(F)
    qc = rng.uniform(lo, hi, (n_quads, 3))
    qa = rng.uniform(10.0, 40.0, (n_quads, 3))
    quads = np.stack([_ellipsoid(qc[i], qa[i]) for i in range(n_quads)]).astype(F) if n_quads else np.zeros((0, 10), F)
    mats, idx = _materials(rng, n_tris + n_quads, 0.2, 0.1)
    ground = np.array([[0.04, 1.0, 0.043, 0.3, 0.7, 0, 0, 0, 0, 0]], F)
    mats = np.concatenate([mats, ground]).astype(F)
    planes = np.array([[0.0, 1.0, 0.0, 45.0]], F)
    return _finish(tri, idx[:n_tris], quads, idx[n_tris:], planes, np.array([len(mats) - 1], np.int32), mats,
                   _lights(rng, n_lights), ((0.0, 150.0, 900.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0)))


def heightfield(n_tris=1000000, extent=1000.0, n_lights=8, seed=4321, amplitude=60.0, p_mirror=0.1, camera=None):
    """BASELINE configs 4/5: a square height-field grid (2 triangles per quad, trimmed to exactly
    n_tris) over [-extent, extent]^2, height = sum of 4 seeded sinusoids, plus the ground plane."""
    rng = np.random.default_rng(seed)
    nq = int(np.ceil(np.sqrt(n_tris / 2.0)))
    xs = np.linspace(-extent, extent, nq + 1)
    X, Z = np.meshgrid(xs, xs, indexing="xy")
    Y = np.zeros_like(X)
    for _ in range(4):
        kx, kz = rng.uniform(0.5, 3.0, 2) * np.pi / extent
        ph = rng.uniform(0, 2 * np.pi)
        Y += (amplitude / 4.0) * np.sin(kx * X + kz * Z + ph)
    V = np.stack([X, Y, Z], axis=-1)
    a, b, c, d = V[:-1, :-1], V[:-1, 1:], V[1:, :-1], V[1:, 1:]
    t1 = np.concatenate([a, c, b], axis=-1).reshape(-1, 9)          # counter-clockwise seen from +y
    t2 = np.concatenate([b, c, d], axis=-1).reshape(-1, 9)
    tri = np.empty((2 * len(t1), 9))
    tri[0::2], tri[1::2] = t1, t2
    tri = tri[:n_tris].astype(F)
    mats, idx = _materials(rng, n_tris, p_mirror, 0.0, specular=(0.3, 30.0), reflect=0.6)
    ground = np.array([[0.04, 1.0, 0.043, 0.3, 0.7, 0, 0, 0, 0, 0]], F)
    mats = np.concatenate([mats, ground]).astype(F)
    planes = np.array([[0.0, 1.0, 0.0, amplitude + 45.0]], F)
    cam = camera or ((0.0, 0.9 * extent, 1.6 * extent), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0))
    return _finish(tri, idx, np.zeros((0, 10), F), np.zeros(0, np.int32), planes, np.array([len(mats) - 1], np.int32),
                   mats, _lights(rng, n_lights, height=(0.8 * extent, 1.6 * extent), spread=extent, intensity=2.0 / max(n_lights, 1)), cam)
