"""ctypes access to librtb200_scene.so: the fresh C++ host scene API (CScene, ISurface
hierarchy, parser, SoA packer) through its small extern "C" test surface."""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# RTB200_SCENE_LIB: another build of the same library (e.g. the ASan / UBSan build, `make asan`)
LIB_PATH = os.path.abspath(os.environ["RTB200_SCENE_LIB"]) if os.environ.get("RTB200_SCENE_LIB") else os.path.join(_HERE, "lib", "librtb200_scene.so")
CLI_PATH = os.path.join(_HERE, "lib", "rt_cli")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} missing: run `make host`")
        L = C.CDLL(LIB_PATH)
        L.rtscene_load.argtypes = [C.c_char_p, C.c_int]
        L.rtscene_pack.argtypes = [C.c_int, C.c_int, C.POINTER(abi.Scene), C.POINTER(abi.Frame)]
        L.rtscene_render.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint, C.c_void_p, C.c_void_p, C.POINTER(abi.Stats)]
        L.rtscene_add_obj.argtypes = [C.c_char_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p]
        L.rtscene_orbit.argtypes = [C.c_int, C.c_int, C.c_uint, C.c_int, C.c_void_p, C.POINTER(abi.Frame), C.POINTER(abi.Stats)]
        L.rtscene_orbit_frames.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(abi.Frame)]
        L.rtscene_free.restype = None
        L.rtscene_camera.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(abi.Frame)]
        _lib = L
    return _lib


def _arr(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def load_and_pack(path, W, H, max_bounces=-1):
    """Parses a scene file with the host CScene, runs Pretraitement + the SoA packer and
    returns (PackedScene, Frame) as copies."""
    L = lib()
    n = L.rtscene_load(os.fsencode(path), max_bounces)
    if n < 0:
        raise FileNotFoundError(path)
    s, f = abi.Scene(), abi.Frame()
    L.rtscene_pack(W, H, C.byref(s), C.byref(f))
    T, Q, P, Ln, M = s.n_triangles, s.n_quadrics, s.n_planes, s.n_lights, s.n_materials
    arrays = {
        "tri_points": _arr(s.tri_points, 9 * T, np.float32), "tri_normal": _arr(s.tri_normal, 3 * T, np.float32),
        "tri_material": _arr(s.tri_material, T, np.int32), "tri_order": _arr(s.tri_order, T, np.int32),
        "quad_coef": _arr(s.quad_coef, 10 * Q, np.float32), "quad_material": _arr(s.quad_material, Q, np.int32),
        "quad_order": _arr(s.quad_order, Q, np.int32), "plane_coef": _arr(s.plane_coef, 4 * P, np.float32),
        "plane_material": _arr(s.plane_material, P, np.int32), "plane_order": _arr(s.plane_order, P, np.int32),
        "materials": _arr(C.cast(s.materials, C.POINTER(C.c_float)), 10 * M, np.float32),
        "lights": _arr(C.cast(s.lights, C.POINTER(C.c_float)), 7 * Ln, np.float32),
    }
    frame = {"cam_pos": np.array(f.cam_pos[:], np.float32), "cam_u": np.array(f.cam_u[:], np.float32),
             "cam_v": np.array(f.cam_v[:], np.float32), "cam_n": np.array(f.cam_n[:], np.float32),
             "half_h": np.float32(f.half_h).item(), "background": np.array(f.background[:], np.float32),
             "max_bounces": f.max_bounces, "min_energy": np.float32(f.min_energy).item(), "scene_ior": np.float32(f.scene_ior).item()}
    fcopy = abi.Frame()
    C.memmove(C.byref(fcopy), C.byref(f), C.sizeof(f))
    return abi.PackedScene(arrays, frame), fcopy


def render_loaded(W, H, spp_grid=1, flags=0):
    """CScene::LancerRayons() of the scene loaded by load_and_pack/rtscene_load (drop-in path)."""
    L = lib()
    f = np.empty((H, W, 3), np.float32)
    q = np.empty((H, W, 4), np.uint8)
    st = abi.Stats()
    L.rtscene_render(W, H, spp_grid, flags, f.ctypes.data, q.ctypes.data, C.byref(st))
    return f, q, st


def orbit_frames(W, H, n):
    """Camera blocks of CScene::OrbiteCamera(n) for the loaded scene (host only, no device)."""
    frames = (abi.Frame * n)()
    lib().rtscene_orbit_frames(W, H, n, frames)
    return list(frames)


def orbit_loaded(W, H, n, flags=0):
    """CScene::LancerTrajectoire over CScene::OrbiteCamera(n): (n, H, W, 4) RGBA8 frames, the camera
    blocks they were drawn with and the per-frame statistics."""
    q = np.empty((n, H, W, 4), np.uint8)
    frames = (abi.Frame * n)()
    st = (abi.Stats * n)()
    lib().rtscene_orbit(W, H, flags, n, q.ctypes.data, frames, st)
    return q, list(frames), list(st)


def add_obj(path, rgb=(0.8, 0.8, 0.8), reflect=0.0, scale=1.0, translate=(0.0, 0.0, 0.0)):
    """CScene::ChargerMaillageOBJ on the currently loaded scene; returns the triangles added."""
    rgb = np.ascontiguousarray(rgb, np.float32)
    tr = np.ascontiguousarray(translate, np.float32)
    n = lib().rtscene_add_obj(os.fsencode(path), rgb.ctypes.data, float(reflect), float(scale), tr.ctypes.data)
    if n < 0:
        raise FileNotFoundError(path)
    return n


def free():
    lib().rtscene_free()


def camera_frame(pos, look, up):
    """Frame dict (basis U,V,N, half_h) from CScene::InitialiserCamera for the given camera."""
    pos, look, up = (np.ascontiguousarray(v, np.float32) for v in (pos, look, up))
    f = abi.Frame()
    lib().rtscene_camera(pos.ctypes.data, look.ctypes.data, up.ctypes.data, 16, 16, C.byref(f))
    return {"cam_pos": np.array(f.cam_pos[:], np.float32), "cam_u": np.array(f.cam_u[:], np.float32),
            "cam_v": np.array(f.cam_v[:], np.float32), "cam_n": np.array(f.cam_n[:], np.float32),
            "half_h": np.float32(f.half_h).item(), "background": np.zeros(3, np.float32),
            "max_bounces": 20, "min_energy": np.float32(0.01).item(), "scene_ior": 1.0,
            "camera_pos": pos, "camera_look": look, "camera_up": up}
