"""rtb200 -- Python harness over librtb200.so (the hand-written sm_100a CUDA render path).

The directory name carries a hyphen, so the package is loaded under the name ``rtb200``
by ``__graft_entry__.load_package()``.  Everything here is a thin ctypes shim: the product
is the shared library (C ABI in include/rtb200.h) and the C++ host API under host/.
There is no CPU fallback -- if the library is missing or no CUDA device is usable the
calls raise.
"""
import ctypes as C
import os

import numpy as np

from . import abi
from .abi import (F_EXACT_FP, F_FORCE_RESIDENT, F_FORCE_STREAM, F_FUSED_FP, F_GATHER_F32, F_HOST_BANDS, F_NO_SECONDARY, F_USE_BVH,  # noqa: F401
                  Frame, PackedScene, Stats, make_frame)

_HERE = os.path.dirname(os.path.abspath(__file__))
# RTB200_LIB: another build of the SAME library (e.g. ab/checked/librtb200.so, `make checked`) -- never a fallback
LIB_PATH = os.path.abspath(os.environ["RTB200_LIB"]) if os.environ.get("RTB200_LIB") else os.path.join(_HERE, "lib", "librtb200.so")

EXPORTS = [
    "rtb200_abi_version", "rtb200_create", "rtb200_destroy", "rtb200_error_string", "rtb200_last_error",
    "rtb200_set_scene", "rtb200_validate_scene", "rtb200_build_bvh", "rtb200_bvh_info", "rtb200_render", "rtb200_render_sequence", "rtb200_read_frame",
    "rtb200_trace_nearest", "rtb200_trace_nearest_static", "rtb200_trace_shadow", "rtb200_comm_unique_id", "rtb200_comm_init", "rtb200_band_map",
    "rtb200_host_alloc", "rtb200_host_free", "rtb200_host_register", "rtb200_host_unregister", "rtb200_set_sharding", "rtb200_measure_fp32_peak", "rtb200_fp32_peak_detail", "rtb200_device_info", "rtb200_device_count",
    "rtb200_timer_mark", "rtb200_timer_elapsed",
]

_lib = None


class RtError(RuntimeError):
    def __init__(self, status, text):
        super().__init__(f"rtb200 status {status}: {text}")
        self.status = status


def lib():
    """Loads librtb200.so (built in-tree by `make lib`); raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} missing: run `make lib` (no CPU fallback exists)")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.rtb200_abi_version.restype = C.c_int
        L.rtb200_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.rtb200_destroy.argtypes = [vp]
        L.rtb200_destroy.restype = None
        L.rtb200_error_string.argtypes = [C.c_int]
        L.rtb200_error_string.restype = C.c_char_p
        L.rtb200_last_error.argtypes = [vp]
        L.rtb200_last_error.restype = C.c_char_p
        L.rtb200_set_scene.argtypes = [vp, C.POINTER(abi.Scene)]
        L.rtb200_build_bvh.argtypes = [vp]
        L.rtb200_bvh_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.rtb200_render.argtypes = [vp, C.POINTER(abi.Frame), vp, vp, C.POINTER(abi.Stats)]
        L.rtb200_read_frame.argtypes = [vp, vp, vp]
        L.rtb200_render_sequence.argtypes = [vp, vp, C.c_int, vp, vp]
        L.rtb200_trace_nearest.argtypes = [vp, C.c_int, vp, vp, C.c_uint32, vp, vp, vp]
        L.rtb200_trace_nearest_static.argtypes = [vp, C.c_int, vp, vp, C.c_uint32, vp, vp, vp]
        L.rtb200_trace_shadow.argtypes = [vp, C.c_int, vp, vp, C.c_uint32, vp]
        L.rtb200_comm_unique_id.argtypes = [vp]
        L.rtb200_comm_init.argtypes = [vp, C.c_int, C.c_int, vp]
        L.rtb200_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
        L.rtb200_host_free.argtypes = [vp]
        L.rtb200_host_free.restype = None
        L.rtb200_host_register.argtypes = [vp, C.c_size_t]
        L.rtb200_host_unregister.argtypes = [vp]
        L.rtb200_set_sharding.argtypes = [vp, C.c_int, C.c_int]
        L.rtb200_measure_fp32_peak.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.rtb200_fp32_peak_detail.argtypes = [vp, C.POINTER(C.c_double), C.c_int]
        L.rtb200_device_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_char_p, C.c_int]
        L.rtb200_band_map.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.POINTER(C.c_int)]
        L.rtb200_validate_scene.argtypes = [vp, C.POINTER(C.c_char_p)]
        L.rtb200_timer_mark.argtypes = [vp, C.c_int]
        L.rtb200_timer_elapsed.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_double)]
        _lib = L
    return _lib


def validate_scene(ps):
    """rtb200_validate_scene on a PackedScene: (ok, reason).  Needs no GPU."""
    why = C.c_char_p()
    s = ps.c_struct()
    rc = lib().rtb200_validate_scene(C.byref(s), C.byref(why))
    return rc == 0, (why.value or b"").decode()


def band_map(height, band_rows, n_ranks, rank):
    """Image rows rendered by `rank` (global row of each of its local rows) -- rtb200_band_map."""
    L = lib()
    n = C.c_int()
    rc = L.rtb200_band_map(height, band_rows, n_ranks, rank, None, 0, C.byref(n))
    if rc != 0:
        raise RtError(rc, "rtb200_band_map")
    rows = np.empty(n.value, np.int32)
    rc = L.rtb200_band_map(height, band_rows, n_ranks, rank, rows.ctypes.data, len(rows), C.byref(n))
    if rc != 0:
        raise RtError(rc, "rtb200_band_map")
    return rows


def pinned_empty(shape, dtype):
    """numpy array over page-locked memory from rtb200_host_alloc (freed with the array)."""
    L = lib()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    rc = L.rtb200_host_alloc(max(n, 1), C.byref(p))
    if rc != 0:
        raise RtError(rc, "rtb200_host_alloc")
    buf = (C.c_ubyte * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    class _Owner:
        def __init__(self, ptr):
            self.ptr = ptr

        def __del__(self):
            try:
                L.rtb200_host_free(self.ptr)
            except Exception:
                pass
    _keep[id(buf)] = _Owner(p)
    return arr


_keep = {}


class Renderer:
    """One context on one GPU (the replacement of CNuanceurCalculProg + the GL dispatch)."""

    def __init__(self, device=0):
        self.L = lib()
        h = C.c_void_p()
        rc = self.L.rtb200_create(int(device), C.byref(h))
        if rc != 0:
            raise RtError(rc, self.L.rtb200_error_string(rc).decode() + " in rtb200_create: " + self.L.rtb200_last_error(None).decode())
        self.h = h
        self.device = device
        self._scene = None

    def _check(self, rc):
        if rc != 0:
            raise RtError(rc, self.L.rtb200_error_string(rc).decode() + ": " + self.L.rtb200_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.rtb200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def device_info(self):
        sm, l2 = C.c_int(), C.c_int()
        name = C.create_string_buffer(128)
        self._check(self.L.rtb200_device_info(self.h, C.byref(sm), C.byref(l2), name, 128))
        return {"sm_count": sm.value, "l2_bytes": l2.value, "name": name.value.decode()}

    def fp32_peak(self):
        tf, mhz = C.c_double(), C.c_double()
        self._check(self.L.rtb200_measure_fp32_peak(self.h, C.byref(tf), C.byref(mhz)))
        return tf.value, mhz.value

    def fp32_peak_detail(self):
        """The FFMA-chain measurement with its parts (rtb200_fp32_peak_detail)."""
        d = (C.c_double * 10)()
        self._check(self.L.rtb200_fp32_peak_detail(self.h, d, 10))
        keys = ("tflops", "ffma_per_thread", "threads_per_block", "blocks", "elapsed_ms", "block_cycles_median",
                "block_ns_median", "sm_mhz_from_clock64_over_globaltimer", "ffma_per_clock_per_sm", "sm_count")
        return dict(zip(keys, [float(v) for v in d]))

    def timer_mark(self, slot):
        self._check(self.L.rtb200_timer_mark(self.h, slot))

    def timer_elapsed(self, a, b):
        ms = C.c_double()
        self._check(self.L.rtb200_timer_elapsed(self.h, a, b, C.byref(ms)))
        return ms.value

    def set_scene(self, ps):
        self._scene = ps                       # keep the arrays alive during the call
        s = ps.c_struct()
        self._check(self.L.rtb200_set_scene(self.h, C.byref(s)))

    def build_bvh(self):
        self._check(self.L.rtb200_build_bvh(self.h))
        return self.bvh_info()

    def bvh_info(self):
        nn, nt = C.c_int(), C.c_int()
        ms, rad, nb = C.c_double(), C.c_double(), C.c_double()
        self._check(self.L.rtb200_bvh_info(self.h, C.byref(nn), C.byref(nt), C.byref(ms), C.byref(rad), C.byref(nb)))
        return {"nodes": nn.value, "triangles": nt.value, "build_ms": ms.value, "pad_radius": rad.value, "node_bytes": nb.value}

    def render(self, frame, want_f32=True, want_rgba8=True, out_f32=None, out_rgba8=None):
        """Returns (float RGB [H,W,3] or None, RGBA8 [H,W,4] or None, Stats)."""
        H, W = frame.height, frame.width
        if want_f32 and out_f32 is None:
            out_f32 = np.empty((H, W, 3), np.float32)
        if want_rgba8 and out_rgba8 is None:
            out_rgba8 = np.empty((H, W, 4), np.uint8)
        st = abi.Stats()
        pf = out_f32.ctypes.data if (want_f32 and out_f32 is not None) else None
        p8 = out_rgba8.ctypes.data if (want_rgba8 and out_rgba8 is not None) else None
        self._check(self.L.rtb200_render(self.h, C.byref(frame), pf, p8, C.byref(st)))
        return (out_f32 if want_f32 else None), (out_rgba8 if want_rgba8 else None), st

    def render_sequence(self, frames, out_rgba8=None, want_stats=True):
        """frames: list of abi.Frame (same size / flags).  out_rgba8: None or a uint8 array
        [n, H, W, 4] (pinned for true overlap).  Returns (out_rgba8, list of Stats)."""
        n = len(frames)
        arr = (abi.Frame * n)(*frames)
        sts = (abi.Stats * n)() if want_stats else None
        p8 = out_rgba8.ctypes.data if out_rgba8 is not None else None
        self._check(self.L.rtb200_render_sequence(self.h, C.byref(arr), n, p8, C.byref(sts) if want_stats else None))
        return out_rgba8, (list(sts) if want_stats else None)

    def read_frame(self, H, W, want_f32=True):
        f = np.empty((H, W, 3), np.float32) if want_f32 else None
        q = np.empty((H, W, 4), np.uint8)
        self._check(self.L.rtb200_read_frame(self.h, f.ctypes.data if want_f32 else None, q.ctypes.data))
        return f, q

    def trace_nearest(self, org, dirs, flags=0):
        org = np.ascontiguousarray(org, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = len(org)
        idx = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        nrm = np.empty((n, 3), np.float32)
        self._check(self.L.rtb200_trace_nearest(self.h, n, org.ctypes.data, dirs.ctypes.data, flags,
                                                idx.ctypes.data, t.ctypes.data, nrm.ctypes.data))
        return idx, t, nrm

    def trace_nearest_static(self, org, dirs, flags=0):
        """rtb200_trace_nearest_static: the same query with its scratch in static device memory (no allocation)."""
        org = np.ascontiguousarray(org, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = len(org)
        idx = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        nrm = np.empty((n, 3), np.float32)
        self._check(self.L.rtb200_trace_nearest_static(self.h, n, org.ctypes.data, dirs.ctypes.data, flags,
                                                       idx.ctypes.data, t.ctypes.data, nrm.ctypes.data))
        return idx, t, nrm

    def trace_shadow(self, pts, lights, flags=0):
        pts = np.ascontiguousarray(pts, np.float32)
        lights = np.ascontiguousarray(lights, np.float32)
        out = np.empty_like(pts)
        self._check(self.L.rtb200_trace_shadow(self.h, len(pts), pts.ctypes.data, lights.ctypes.data, flags, out.ctypes.data))
        return out

    # multi-GPU ----------------------------------------------------------------------------
    def set_sharding(self, rank, n_ranks):
        """Band sharding without a communicator (rtb200_set_sharding)."""
        self._check(self.L.rtb200_set_sharding(self.h, rank, n_ranks))

    def host_register(self, arr):
        self._check(self.L.rtb200_host_register(arr.ctypes.data, arr.nbytes))

    def host_unregister(self, arr):
        self._check(self.L.rtb200_host_unregister(arr.ctypes.data))

    def comm_unique_id(self):
        buf = (C.c_ubyte * abi.UNIQUE_ID_BYTES)()
        rc = self.L.rtb200_comm_unique_id(buf)
        if rc != 0:
            raise RtError(rc, "rtb200_comm_unique_id")
        return bytes(buf)

    def comm_init(self, rank, n_ranks, uid):
        buf = (C.c_ubyte * abi.UNIQUE_ID_BYTES).from_buffer_copy(uid)
        self._check(self.L.rtb200_comm_init(self.h, rank, n_ranks, buf))
