"""ctypes mirror of include/rtb200.h (struct layouts and a container for packed scenes).

Python is only the test / benchmark harness of this repository: the product is
librtb200.so (CUDA, C ABI) and the C++ host scene API under host/.  Nothing in this
module computes anything -- it moves numpy arrays across the C ABI.
"""
import ctypes as C

import numpy as np

UNIQUE_ID_BYTES = 128

F_USE_BVH = 1 << 0
F_NO_SECONDARY = 1 << 1
F_EXACT_FP = 1 << 2
F_FORCE_STREAM = 1 << 3
F_FORCE_RESIDENT = 1 << 4
F_FUSED_FP = 1 << 5
F_GATHER_F32 = 1 << 6
F_HOST_BANDS = 1 << 7

_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int32)


class Material(C.Structure):
    _fields_ = [("color", C.c_float * 3), ("ambient", C.c_float), ("diffuse", C.c_float),
                ("specular", C.c_float), ("shininess", C.c_float), ("reflect", C.c_float),
                ("refract", C.c_float), ("ior", C.c_float)]


class Light(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("color", C.c_float * 3), ("intensity", C.c_float)]


class Scene(C.Structure):
    _fields_ = [("n_triangles", C.c_int32), ("n_quadrics", C.c_int32), ("n_planes", C.c_int32),
                ("n_lights", C.c_int32), ("n_materials", C.c_int32),
                ("tri_points", _fp), ("tri_normal", _fp), ("tri_material", _ip), ("tri_order", _ip),
                ("quad_coef", _fp), ("quad_material", _ip), ("quad_order", _ip),
                ("plane_coef", _fp), ("plane_material", _ip), ("plane_order", _ip),
                ("materials", C.POINTER(Material)), ("lights", C.POINTER(Light))]


class Frame(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp_grid", C.c_int32),
                ("max_bounces", C.c_int32), ("min_energy", C.c_float), ("scene_ior", C.c_float),
                ("background", C.c_float * 3), ("cam_pos", C.c_float * 3),
                ("cam_u", C.c_float * 3), ("cam_v", C.c_float * 3), ("cam_n", C.c_float * 3),
                ("half_w", C.c_float), ("half_h", C.c_float), ("flags", C.c_uint32),
                ("band_rows", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("ms_render", C.c_double), ("ms_gather", C.c_double), ("ms_h2d", C.c_double),
                ("ms_d2h", C.c_double),
                ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("rays_reflect", C.c_uint64), ("rays_refract", C.c_uint64),
                ("tests_triangle", C.c_uint64), ("tests_quadric", C.c_uint64),
                ("tests_plane", C.c_uint64), ("bvh_nodes_visited", C.c_uint64),
                ("nan_pixels", C.c_uint64),
                ("kernel_launches", C.c_uint32), ("depth_levels", C.c_uint32),
                ("path", C.c_uint32), ("chunks", C.c_uint32),
                ("ms_trace_nearest", C.c_double), ("ms_trace_shadow", C.c_double),
                ("ms_shade", C.c_double), ("ms_primary_kernel", C.c_double),
                ("shadow_tests_triangle", C.c_uint64), ("shadow_bvh_nodes", C.c_uint64),
                ("level_loop_redone", C.c_uint64), ("tests_triangle_executed", C.c_uint64),
                ("shadow_fixups", C.c_uint64), ("bvh_out_of_radius", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}

    @property
    def rays(self):
        return self.rays_primary + self.rays_shadow + self.rays_reflect + self.rays_refract


_KEYS_F = {"tri_points": 9, "tri_normal": 3, "quad_coef": 10, "plane_coef": 4, "materials": 10, "lights": 7}
_KEYS_I = ("tri_material", "tri_order", "quad_material", "quad_order", "plane_material", "plane_order")


class PackedScene:
    """Packed scene = the arrays of `rtb200_scene` as numpy arrays (kept alive here) plus the
    camera/frame defaults that came with the scene (dict `frame`)."""

    def __init__(self, arrays, frame=None):
        self.a = {}
        for k, w in _KEYS_F.items():
            self.a[k] = np.ascontiguousarray(np.asarray(arrays.get(k, np.zeros((0, w))), np.float32).reshape(-1, w))
        for k in _KEYS_I:
            self.a[k] = np.ascontiguousarray(np.asarray(arrays.get(k, np.zeros(0)), np.int32).reshape(-1))
        self.frame = dict(frame or {})

    def __getitem__(self, k):
        return self.a[k]

    @property
    def counts(self):
        return (len(self.a["tri_order"]), len(self.a["quad_order"]), len(self.a["plane_order"]), len(self.a["lights"]))

    def c_struct(self):
        """The `rtb200_scene` view of the arrays (pointers into them; built once, the arrays are never rebound)."""
        if getattr(self, "_c_struct", None) is not None:
            return self._c_struct
        a = self.a
        s = Scene()
        s.n_triangles, s.n_quadrics, s.n_planes, s.n_lights = self.counts
        s.n_materials = len(a["materials"])
        for k in _KEYS_F:
            if k in ("materials", "lights"):
                continue
            setattr(s, k, a[k].ctypes.data_as(_fp))
        for k in _KEYS_I:
            setattr(s, k, a[k].ctypes.data_as(_ip))
        s.materials = a["materials"].ctypes.data_as(C.POINTER(Material))
        s.lights = a["lights"].ctypes.data_as(C.POINTER(Light))
        self._c_struct = s
        return s

    def nbytes(self):
        return int(sum(v.nbytes for v in self.a.values()))

    def save(self, path):
        np.savez_compressed(path, **self.a, **{"frame_" + k: np.asarray(v) for k, v in self.frame.items()})

    @staticmethod
    def load(path):
        z = np.load(path)
        arrays = {k: z[k] for k in z.files if not k.startswith("frame_")}
        frame = {k[6:]: z[k] for k in z.files if k.startswith("frame_")}
        frame = {k: (v.item() if v.ndim == 0 else v) for k, v in frame.items()}
        return PackedScene(arrays, frame)


def make_frame(ps_frame, width, height, max_bounces=None, spp_grid=1, flags=0, band_rows=0, min_energy=None):
    """Builds an rtb200_frame from the camera block a packed scene carries.  half_w follows
    Scene.cpp:677: (float(W) / H) * half_h in binary32."""
    f = Frame()
    f.width, f.height, f.spp_grid = int(width), int(height), int(spp_grid)
    f.max_bounces = int(ps_frame.get("max_bounces", 20) if max_bounces is None else max_bounces)
    f.min_energy = float(ps_frame.get("min_energy", 0.01) if min_energy is None else min_energy)
    f.scene_ior = float(ps_frame.get("scene_ior", 1.0))
    for name in ("background", "cam_pos", "cam_u", "cam_v", "cam_n"):
        v = np.asarray(ps_frame[name], np.float32)
        for i in range(3):
            getattr(f, name)[i] = float(v[i])
    hh = np.float32(ps_frame["half_h"])
    f.half_h = float(hh)
    f.half_w = float((np.float32(width) / np.float32(height)) * hh)
    f.flags = int(flags)
    f.band_rows = int(band_rows)
    return f
