"""ctypes binding of oracle/_ref/libref_*.so -- TEST INFRASTRUCTURE ONLY.

The library is the UNMODIFIED reference renderer (Rodyll/Ray-Tracing-GPU,
Projet-INF8702/*.cpp) compiled in place by oracle/Makefile plus the headless
driver oracle/ref_driver.cpp.  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may import this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
FLAVOURS = {"shipped": "libref_shipped.so", "rec": "libref_rec.so", "rec_ndebug": "libref_rec_ndebug.so",
            "gpu": "libref_gpu.so"}      # gpu: the reference with the INTEGRATION.md binding behind g_ComputerShadersON

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def available(flavour="rec"):
    return os.path.exists(os.path.join(_HERE, "_ref", FLAVOURS[flavour]))


class RefRenderer:
    """One loaded copy of the reference renderer (the reference's CScene is a process-wide
    singleton, so there is one live scene per loaded flavour)."""

    def __init__(self, flavour="rec"):
        path = os.path.join(_HERE, "_ref", FLAVOURS[flavour])
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path}: build it with `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(path)
        self.lib = L
        L.ref_flavour.restype = C.c_char_p
        L.ref_max_threads.restype = C.c_int
        L.ref_load_scene.argtypes = [C.c_char_p]
        L.ref_load_scene.restype = C.c_int
        L.ref_set_camera.argtypes = [_f32p, _f32p, _f32p]
        L.ref_set_background.argtypes = [_f32p]
        L.ref_set_params.argtypes = [C.c_int, C.c_float, C.c_float]
        L.ref_add_triangle.argtypes = [_f32p, _f32p]
        L.ref_add_quadric.argtypes = [_f32p, _f32p]
        L.ref_add_plane.argtypes = [_f32p, _f32p]
        L.ref_add_triangles.argtypes = [C.c_int, _f32p, _f32p, _i32p]
        L.ref_add_quadrics.argtypes = [C.c_int, _f32p, _f32p, _i32p]
        L.ref_add_light.argtypes = [_f32p, _f32p, C.c_float]
        L.ref_render_full.argtypes = [C.c_int, C.c_int, _f32p]
        L.ref_render_full.restype = C.c_double
        L.ref_prepare.argtypes = [C.c_int, C.c_int]
        L.ref_prepare.restype = C.c_int
        L.ref_dump_surface.argtypes = [C.c_int, C.POINTER(C.c_int), _f32p, _f32p]
        L.ref_dump_light.argtypes = [C.c_int, _f32p]
        L.ref_dump_camera.argtypes = [_f32p]
        L.ref_render_region.argtypes = [C.c_int] * 7 + [_f32p, C.c_int]
        L.ref_render_region.restype = C.c_double
        L.ref_use_gpu_branch.argtypes = [C.c_int]
        L.ref_use_gpu_branch.restype = C.c_int
        L.ref_trace_rays.argtypes = [C.c_int, _f32p, _f32p, _f32p, C.c_int]
        L.ref_nearest_hits.argtypes = [C.c_int, _f32p, _f32p, _i32p, _f32p, _f32p, C.c_int]
        L.ref_shadow_filters.argtypes = [C.c_int, _f32p, _f32p, _f32p, C.c_int]
        self.flavour = L.ref_flavour().decode()
        self.max_threads = L.ref_max_threads()

    # -- scene construction ------------------------------------------------------------
    def load(self, path, max_bounces=-1, min_energy=-1.0, scene_ior=-1.0):
        n = self.lib.ref_load_scene(os.fsencode(path))
        if n < 0:
            raise FileNotFoundError(path)
        self.lib.ref_set_params(max_bounces, min_energy, scene_ior)
        return n

    def new_scene(self, camera=None, background=None, max_bounces=-1, min_energy=-1.0, scene_ior=-1.0):
        self.lib.ref_new_scene()
        if camera is not None:
            pos, look, up = (np.ascontiguousarray(v, np.float32) for v in camera)
            self.lib.ref_set_camera(pos, look, up)
        if background is not None:
            self.lib.ref_set_background(np.ascontiguousarray(background, np.float32))
        self.lib.ref_set_params(max_bounces, min_energy, scene_ior)

    def add_packed(self, ps):
        """Feeds a packed scene (dict of numpy arrays, layout of include/rtb200.h, geometry
        ALREADY final -- transforms are identity) through the reference's public API, in
        insertion order."""
        T, Q, P = len(ps["tri_order"]), len(ps["quad_order"]), len(ps["plane_order"])
        if T > 1000 and np.array_equal(ps["tri_order"], np.arange(T)):
            # large synthetic meshes: the triangles come first in insertion order -> one bulk call
            self.lib.ref_add_triangles(T, np.ascontiguousarray(ps["tri_points"], np.float32).reshape(-1),
                                       np.ascontiguousarray(ps["materials"], np.float32).reshape(-1),
                                       np.ascontiguousarray(ps["tri_material"], np.int32))
            rest = {k: ps[k] for k in ("quad_order", "quad_coef", "quad_material", "plane_order", "plane_coef", "plane_material", "materials", "lights")}
            rest.update({"tri_order": np.zeros(0, np.int32), "tri_points": np.zeros((0, 9), np.float32), "tri_material": np.zeros(0, np.int32)})
            self._add_items(rest)
            assert T + Q + P == self.lib.ref_num_surfaces()
            return
        self._add_items(ps)
        assert T + Q + P == self.lib.ref_num_surfaces()

    def _add_items(self, ps):
        items = [(int(o), 0, i) for i, o in enumerate(ps["tri_order"])]
        items += [(int(o), 1, i) for i, o in enumerate(ps["quad_order"])]
        items += [(int(o), 2, i) for i, o in enumerate(ps["plane_order"])]
        items.sort()
        mats = np.ascontiguousarray(ps["materials"], np.float32).reshape(-1, 10)
        tp = np.ascontiguousarray(ps["tri_points"], np.float32).reshape(-1, 9)
        qc = np.ascontiguousarray(ps["quad_coef"], np.float32).reshape(-1, 10)
        pc = np.ascontiguousarray(ps["plane_coef"], np.float32).reshape(-1, 4)
        for _, kind, i in items:
            if kind == 0:
                self.lib.ref_add_triangle(tp[i], mats[int(ps["tri_material"][i])])
            elif kind == 1:
                self.lib.ref_add_quadric(qc[i], mats[int(ps["quad_material"][i])])
            else:
                self.lib.ref_add_plane(pc[i], mats[int(ps["plane_material"][i])])
        lights = np.ascontiguousarray(ps["lights"], np.float32).reshape(-1, 7)
        for l in lights:
            self.lib.ref_add_light(np.ascontiguousarray(l[0:3]), np.ascontiguousarray(l[3:6]), float(l[6]))

    def use_gpu_branch(self, on):
        """CVar::g_ComputerShadersON for the next render_full (Scene.cpp:692); only the "gpu" flavour accepts it."""
        if self.lib.ref_use_gpu_branch(1 if on else 0) != 0:
            raise RuntimeError("this flavour has no GPU branch that can run here")

    # -- rendering ---------------------------------------------------------------------
    def render_full(self, W, H):
        out = np.empty((H, W, 3), np.float32)
        secs = self.lib.ref_render_full(W, H, out)
        if secs < 0:
            raise RuntimeError(f"ref_render_full failed ({secs})")
        return out, secs

    def prepare(self, W, H):
        if self.lib.ref_prepare(W, H) != 0:
            raise RuntimeError("scene already prepared (Pretraitement is not idempotent)")

    def render_region(self, W, H, k, x0, y0, w, h, threads=1):
        out = np.empty((h, w, 3), np.float32)
        secs = self.lib.ref_render_region(W, H, k, x0, y0, w, h, out, threads)
        return out, secs

    def trace_rays(self, org, dirs, threads=1):
        org = np.ascontiguousarray(org, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.empty_like(org)
        self.lib.ref_trace_rays(len(org), org, dirs, out, threads)
        return out

    def nearest_hits(self, org, dirs, threads=1):
        org = np.ascontiguousarray(org, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = len(org)
        idx = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        nrm = np.empty((n, 3), np.float32)
        self.lib.ref_nearest_hits(n, org, dirs, idx, t, nrm, threads)
        return idx, t, nrm

    def shadow_filters(self, pts, lights, threads=1):
        pts = np.ascontiguousarray(pts, np.float32)
        lights = np.ascontiguousarray(lights, np.float32)
        out = np.empty_like(pts)
        self.lib.ref_shadow_filters(len(pts), pts, lights, out, threads)
        return out

    # -- dumps -------------------------------------------------------------------------
    def dump(self):
        """Post-Pretraitement scene as seen through the reference's getters."""
        n = self.lib.ref_num_surfaces()
        types = np.empty(n, np.int32)
        geom = np.zeros((n, 16), np.float32)
        mats = np.zeros((n, 10), np.float32)
        t = C.c_int()
        for i in range(n):
            self.lib.ref_dump_surface(i, C.byref(t), geom[i], mats[i])
            types[i] = t.value
        nl = self.lib.ref_num_lights()
        lights = np.zeros((nl, 7), np.float32)
        for i in range(nl):
            self.lib.ref_dump_light(i, lights[i])
        cam = np.zeros(21, np.float32)
        self.lib.ref_dump_camera(cam)
        return {"types": types, "geom": geom, "materials": mats, "lights": lights, "camera": cam}

    def free(self):
        self.lib.ref_free_scene()


def quantise_rgba8(rgb):
    """The 8-bit quantiser SURVEY.md 8a/a10 fixes for both sides (GL_RGBA8 upload semantics):
    q = floor(clamp(c,0,1)*255 + 0.5), NaN -> 0, alpha = 255."""
    c = np.nan_to_num(np.asarray(rgb, np.float32), nan=0.0)
    c = np.minimum(np.maximum(c, np.float32(0)), np.float32(1))
    q = np.floor(c * np.float32(255) + np.float32(0.5)).astype(np.uint8)
    a = np.full(q.shape[:-1] + (1,), 255, np.uint8)
    return np.concatenate([q, a], axis=-1)
