"""ctypes binding of the C ABI in ``include/rtc.h`` (the same declarations a Rust
``extern "C"`` block would carry).

There is no fallback of any kind: if ``librtc_b200.so`` has not been built, or the machine has
no CUDA device, every product entry point raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTC_B200_LIB") or os.path.join(_HERE, "csrc", "librtc_b200.so")   # override: kernel-variant experiments

RTC_OK = 0
RTC_ERR_INVALID_ARG, RTC_ERR_UNSUPPORTED, RTC_ERR_NO_DEVICE, RTC_ERR_CUDA, RTC_ERR_NO_SCENE, RTC_ERR_NCCL, RTC_ERR_TIMEOUT = 1, 2, 3, 4, 5, 6, 7
SHAPE_SPHERE, SHAPE_PLANE, SHAPE_CUBE = 0, 1, 2
PATTERN_NONE, PATTERN_TEST, PATTERN_STRIPE, PATTERN_GRADIENT, PATTERN_RING, PATTERN_CHECKER, PATTERN_GRID = range(7)
PRECISION_F64_PARITY, PRECISION_F32_FAST = 0, 1
RECT_REFERENCE_RENDER, RECT_FULL = 0, 1
OUT_F64_RGB, OUT_F32_RGB, OUT_U8_RGB = 0, 1, 2
RTC_MAX_DEPTH = 8
MAX_LIGHTS = 8   # RTC_MAX_LIGHTS
OPT_DST_FULL_FRAME = 1
OPT_F32_GENERAL_KERNEL = 2
OPT_LINEAR_TILE_ORDER = 4
OPT_NO_BOUND_FILTER = 8
OPT_DELIVER_DIRECT = 16
OPT_DELIVER_STAGED = 32
NOALLOC_MAX_TILES = 1 << 20   # RTC_NOALLOC_MAX_TILES
IPC_HANDLE_BYTES = 64


class RtcShape(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("world_id", C.c_int32),
        ("inv", C.c_double * 16), ("inv_t", C.c_double * 16),
        ("has_color", C.c_int32), ("pattern_kind", C.c_int32),
        ("color", C.c_double * 3),
        ("ambient", C.c_double), ("diffuse", C.c_double), ("specular", C.c_double),
        ("shininess", C.c_double), ("reflective", C.c_double), ("transparency", C.c_double),
        ("refractive_index", C.c_double),
        ("pat_inv", C.c_double * 16), ("pat_a", C.c_double * 3), ("pat_b", C.c_double * 3),
    ]


class RtcLight(C.Structure):
    _fields_ = [("intensity", C.c_double * 3), ("position", C.c_double * 3)]


class RtcCamera(C.Structure):
    _fields_ = [
        ("hsize", C.c_uint32), ("vsize", C.c_uint32),
        ("half_width", C.c_double), ("half_height", C.c_double), ("pixel_size", C.c_double),
        ("view_inv", C.c_double * 16),
        ("samples", C.c_uint32), ("aa_seed", C.c_uint32),
    ]


class RtcRenderOpts(C.Structure):
    _fields_ = [
        ("precision", C.c_int32), ("rect", C.c_int32), ("max_depth", C.c_int32), ("out_format", C.c_int32),
        ("shard_index", C.c_int32), ("shard_count", C.c_int32), ("tile_rows", C.c_int32), ("flags", C.c_int32),
    ]


class RtcStats(C.Structure):
    _fields_ = [
        ("rays", C.c_uint64), ("primary_hits", C.c_uint64), ("shaded_hits", C.c_uint64),
        ("shadow_rays", C.c_uint64), ("kernel_launches", C.c_uint64),
        ("h2d_ms", C.c_double), ("kernel_ms", C.c_double), ("gather_ms", C.c_double),
        ("d2h_ms", C.c_double), ("total_ms", C.c_double),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


# every symbol include/rtc.h declares: (name, restype, argtypes)
EXPORTS = [
    ("rtc_abi_version", C.c_int, []),
    ("rtc_last_error", C.c_char_p, []),
    ("rtc_device_count", C.c_int, []),
    ("rtc_context_create", C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]),
    ("rtc_context_destroy", None, [C.c_void_p]),
    ("rtc_scene_upload", C.c_int, [C.c_void_p, C.POINTER(RtcLight), C.POINTER(RtcShape), C.c_size_t]),
    ("rtc_scene_upload_lights", C.c_int, [C.c_void_p, C.POINTER(RtcLight), C.c_size_t, C.POINTER(RtcShape), C.c_size_t]),
    ("rtc_render", C.c_int, [C.c_void_p, C.POINTER(RtcCamera), C.POINTER(RtcRenderOpts), C.c_void_p,
                             C.c_void_p, C.POINTER(RtcStats)]),
    ("rtc_shard_rows", C.c_int, [C.POINTER(RtcCamera), C.POINTER(RtcRenderOpts)]),
    ("rtc_render_device", C.c_int, [C.c_void_p, C.POINTER(RtcCamera), C.POINTER(RtcRenderOpts), C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.POINTER(RtcStats)]),
    ("rtc_assemble_device", C.c_int, [C.c_void_p, C.POINTER(RtcCamera), C.POINTER(RtcRenderOpts), C.c_void_p,
                                      C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    ("rtc_ipc_export", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]),
    ("rtc_ipc_open", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p)]),
    ("rtc_ipc_close", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    ("rtc_signal_device", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    ("rtc_wait_device", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_void_p]),
    ("rtc_wait_signal_device", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p]),
    ("rtc_set_wait_timeout", C.c_int, [C.c_void_p, C.c_double]),
    ("rtc_wait_result", C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]),
    ("rtc_render_device_synced", C.c_int, [C.c_void_p, C.POINTER(RtcCamera), C.POINTER(RtcRenderOpts), C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]),
    ("rtc_render_device_noalloc", C.c_int, [C.c_void_p, C.POINTER(RtcCamera), C.POINTER(RtcRenderOpts), C.c_void_p, C.c_void_p, C.c_void_p]),
    ("rtc_register_host_buffer", C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    ("rtc_unregister_host_buffer", C.c_int, [C.c_void_p, C.c_void_p]),
    ("rtc_delivery_info", C.c_int, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("rtc_pack_rgba8_device", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_float, C.c_void_p, C.c_void_p]),
    ("rtc_download_frame", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]),
    ("rtc_measure_peak_tflops", C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_double)]),
    ("rtc_scene_selfcheck", C.c_int, [C.POINTER(RtcLight), C.c_size_t, C.POINTER(RtcShape), C.c_size_t, C.c_void_p]),
]

_lib = None


class RtcSelfcheck(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_planes", "n_uniform_spheres", "n_general_spheres", "n_cubes", "streams_are_permutations",
                                          "morton_sorted", "n_groups_f32", "n_blocks_f32", "n_groups_f64", "n_blocks_f64",
                                          "small_scene_form", "reserved")] + [("worst_enclosure_f32", C.c_double), ("worst_enclosure_f64", C.c_double)]


class RtcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"rtc error {code}: {msg}")
        self.code = code


def lib():
    """Load librtc_b200.so (built in-tree by __graft_entry__.build()).  Raises if missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback for the render path)")
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, res, args in EXPORTS:
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code):
    if code != RTC_OK:
        msg = lib().rtc_last_error()
        raise RtcError(code, msg.decode() if msg else "")
