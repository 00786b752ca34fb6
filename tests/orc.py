"""ctypes binding of the ORACLE (oracle/liborc.so) — test infrastructure only.

Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
Never imported by the product package.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import rtc_b200 as rtc  # noqa: E402
from rtc_b200 import ffi  # noqa: E402

ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "liborc.so")
KAT_PATH = os.path.join(ORACLE_DIR, "orc_kat")


class OrcCounters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("shaded_hits", C.c_uint64),
                ("sphere_tests", C.c_uint64), ("plane_tests", C.c_uint64), ("cube_tests", C.c_uint64),
                ("seconds", C.c_double)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None


def build():
    subprocess.run(["make", "-j2"], cwd=ORACLE_DIR, check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_last_error.restype = C.c_char_p
        L.orc_world_create.restype = C.c_void_p
        L.orc_world_create.argtypes = [C.POINTER(ffi.RtcLight), C.POINTER(ffi.RtcShape), C.c_size_t]
        L.orc_world_destroy.argtypes = [C.c_void_p]
        L.orc_world_add_light.argtypes = [C.c_void_p, C.POINTER(ffi.RtcLight)]
        L.orc_render.restype = C.c_int
        L.orc_render.argtypes = [C.c_void_p, C.POINTER(ffi.RtcCamera), C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(OrcCounters)]
        L.orc_matrix_chain.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double)]
        L.orc_matrix_inverse.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.orc_matrix_transpose.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.orc_view_transform.argtypes = [C.POINTER(C.c_double)] * 4
        L.orc_camera_new.argtypes = [C.c_uint32, C.c_uint32, C.c_double, C.POINTER(C.c_double), C.POINTER(ffi.RtcCamera)]
        L.orc_quantize_u8.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_sample.restype = C.c_int
        L.orc_sample.argtypes = [C.c_void_p, C.POINTER(ffi.RtcCamera), C.c_uint32, C.c_double, C.c_uint32, C.c_double, C.c_int,
                                 C.POINTER(C.c_double)]
        L.orc_resample_offset.restype = C.c_double
        L.orc_resample_offset.argtypes = [C.c_uint32] * 6
        _lib = L
    return _lib


class OracleWorld:
    def __init__(self, world=None, flat=None):
        light, shapes, n = flat if flat is not None else world.flatten()
        self._h = lib().orc_world_create(C.byref(light), shapes, n)
        if not self._h:
            raise RuntimeError("oracle: " + lib().orc_last_error().decode())
        if flat is None and getattr(world, "extra_lights", None):
            for l in world.flatten_lights()[1:]:
                lib().orc_world_add_light(self._h, C.byref(l))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_world_destroy(self._h)
            self._h = None

    def render(self, cam, rect=ffi.RECT_FULL, max_depth=5, threads=1, pixels=None):
        """-> (rgb f64, hit u8, counters).  Full frame: shapes (h,w,3)/(h,w); with `pixels`
        (flat y*w+x indices): (n,3)/(n,)."""
        desc = cam.flatten() if hasattr(cam, "flatten") else cam
        h, w = desc.vsize, desc.hsize
        cnt = OrcCounters()
        if pixels is None:
            rgb = np.zeros((h, w, 3), dtype=np.float64)
            hit = np.zeros((h, w), dtype=np.uint8)
            idx_p, n = None, 0
        else:
            idx = np.ascontiguousarray(pixels, dtype=np.uint32)
            rgb = np.zeros((idx.size, 3), dtype=np.float64)
            hit = np.zeros((idx.size,), dtype=np.uint8)
            idx_p, n = idx.ctypes.data_as(C.c_void_p), idx.size
        rc = lib().orc_render(self._h, C.byref(desc), rect, max_depth, threads, idx_p, n,
                              rgb.ctypes.data_as(C.c_void_p), hit.ctypes.data_as(C.c_void_p), C.byref(cnt))
        if rc != 0:
            raise RuntimeError("oracle: " + lib().orc_last_error().decode())
        return rgb, hit, cnt


def sample(oracle_world, cam, x, xo, y, yo, max_depth=5):
    """world.color_at(camera.ray_for_pixel_offset(x, xo, y, yo), max_depth) -> (r, g, b) as Python floats."""
    desc = cam.flatten() if hasattr(cam, "flatten") else cam
    out = (C.c_double * 3)()
    if lib().orc_sample(oracle_world._h, C.byref(desc), x, xo, y, yo, max_depth, out) != 0:
        raise RuntimeError("oracle: " + lib().orc_last_error().decode())
    return (out[0], out[1], out[2])


def resample_offset(aa_seed, hsize, x, y, k, axis):
    return lib().orc_resample_offset(aa_seed, hsize, x, y, k, axis)


def render(world, cam, rect=ffi.RECT_FULL, max_depth=5, threads=1, pixels=None):
    return OracleWorld(world).render(cam, rect, max_depth, threads, pixels)


def quantize_u8(rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.float64)
    out = np.empty(rgb.shape, dtype=np.uint8)
    lib().orc_quantize_u8(rgb.ctypes.data_as(C.c_void_p), rgb.size, out.ctypes.data_as(C.c_void_p))
    return out


def lattice(hsize, vsize, step=8, n_random=4096, seed=1234):
    """Pixel subset used for parity at sizes the oracle cannot finish in seconds: every
    `step`-th pixel in x and y plus `n_random` random pixels."""
    ys, xs = np.meshgrid(np.arange(0, vsize, step), np.arange(0, hsize, step), indexing="ij")
    idx = (ys * hsize + xs).reshape(-1)
    rng = np.random.default_rng(seed)
    extra = rng.integers(0, hsize * vsize, size=n_random)
    return np.unique(np.concatenate([idx, extra])).astype(np.uint32)
