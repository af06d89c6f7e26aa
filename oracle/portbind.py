"""ctypes binding of oracle/librt_oracle.so (the plain-C restatement, rt_oracle.c) --
TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, never from the product."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.abspath(os.environ["RTB200_ORACLE_LIB"]) if os.environ.get("RTB200_ORACLE_LIB") else os.path.join(_HERE, "librt_oracle.so")

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, _LIB])


class PortOracle:
    def __init__(self, abi):
        """`abi` = the product package's abi module (struct layouts of include/rtb200.h only)."""
        if not os.path.exists(_LIB):
            build()
        self.abi = abi
        L = C.CDLL(_LIB)
        self.lib = L
        sp, fp = C.POINTER(abi.Scene), C.POINTER(abi.Frame)
        L.rto_render.argtypes = [sp, fp, _f32p, C.c_int]
        L.rto_render.restype = C.c_double
        L.rto_render_region.argtypes = [sp, fp, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, C.c_int]
        L.rto_render_region.restype = C.c_double
        L.rto_signatures.argtypes = [sp, fp, _u64p, _i32p, C.c_int]
        L.rto_nearest.argtypes = [sp, C.c_int, _f32p, _f32p, _i32p, _f32p, _f32p, C.c_int]
        L.rto_shadow.argtypes = [sp, C.c_int, _f32p, _f32p, _f32p, C.c_int]
        L.rto_trace.argtypes = [sp, fp, C.c_int, _f32p, _f32p, _f32p, C.c_int]
        L.rto_quantise.argtypes = [_f32p, C.c_size_t, _u8p]
        L.rto_quantise.restype = C.c_uint64
        L.rto_last_ray_counts.argtypes = [_u64p]
        L.rto_max_threads.restype = C.c_int
        self.max_threads = L.rto_max_threads()

    def render(self, ps, frame, threads=1):
        s = ps.c_struct()
        out = np.empty((frame.height, frame.width, 3), np.float32)
        secs = self.lib.rto_render(C.byref(s), C.byref(frame), out, threads)
        return out, secs

    def render_region(self, ps, frame, x0, y0, w, h, threads=1):
        s = ps.c_struct()
        out = np.empty((h, w, 3), np.float32)
        secs = self.lib.rto_render_region(C.byref(s), C.byref(frame), x0, y0, w, h, out, threads)
        return out, secs

    def ray_counts(self):
        c = np.zeros(4, np.uint64)
        self.lib.rto_last_ray_counts(c)
        return {"primary": int(c[0]), "shadow": int(c[1]), "reflect": int(c[2]), "refract": int(c[3])}

    def signatures(self, ps, frame, threads=1):
        s = ps.c_struct()
        sig = np.empty((frame.height, frame.width), np.uint64)
        hit = np.empty((frame.height, frame.width), np.int32)
        self.lib.rto_signatures(C.byref(s), C.byref(frame), sig, hit, threads)
        return sig, hit

    def nearest(self, ps, org, dirs, threads=1):
        s = ps.c_struct()
        org = np.ascontiguousarray(org, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = len(org)
        idx = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        nrm = np.empty((n, 3), np.float32)
        self.lib.rto_nearest(C.byref(s), n, org, dirs, idx, t, nrm, threads)
        return idx, t, nrm

    def shadow(self, ps, pts, lights, threads=1):
        s = ps.c_struct()
        pts = np.ascontiguousarray(pts, np.float32)
        lights = np.ascontiguousarray(lights, np.float32)
        out = np.empty_like(pts)
        self.lib.rto_shadow(C.byref(s), len(pts), pts, lights, out, threads)
        return out

    def trace(self, ps, frame, org, dirs, threads=1):
        s = ps.c_struct()
        org = np.ascontiguousarray(org, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.empty_like(org)
        self.lib.rto_trace(C.byref(s), C.byref(frame), len(org), org, dirs, out, threads)
        return out

    def quantise(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.float32)
        out = np.empty(rgb.shape[:-1] + (4,), np.uint8)
        nans = self.lib.rto_quantise(rgb.reshape(-1), rgb.size // 3, out.reshape(-1))
        return out, int(nans)


def packed_from_ref_dump(abi, dump):
    """Packed scene from the reference's own post-Pretraitement getters (RefRenderer.dump())."""
    types, geom, mats = dump["types"], dump["geom"], dump["materials"]
    cam = dump["camera"]
    arrays = {
        "tri_points": geom[types == 0][:, 0:9], "tri_normal": geom[types == 0][:, 9:12],
        "tri_material": np.nonzero(types == 0)[0], "tri_order": np.nonzero(types == 0)[0],
        "quad_coef": geom[types == 1][:, 0:10],
        "quad_material": np.nonzero(types == 1)[0], "quad_order": np.nonzero(types == 1)[0],
        "plane_coef": geom[types == 2][:, 0:4],
        "plane_material": np.nonzero(types == 2)[0], "plane_order": np.nonzero(types == 2)[0],
        "materials": mats, "lights": dump["lights"],
    }
    frame = {"cam_pos": cam[0:3], "cam_u": cam[3:6], "cam_v": cam[6:9], "cam_n": cam[9:12],
             "angle": float(cam[12]), "half_h": float(cam[14]), "background": cam[15:18],
             "max_bounces": int(cam[18]), "min_energy": float(cam[19]), "scene_ior": float(cam[20])}
    return abi.PackedScene(arrays, frame)
