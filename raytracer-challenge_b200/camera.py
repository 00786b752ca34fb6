"""Host-side mirror of ``Camera`` (``src/camera.rs:16-162``) and ``Canvas`` (``src/canvas.rs``),
plus ``Renderer``: the thin owner of an ``rtc_context`` through which ``camera.render(world)``
reaches the CUDA kernels.  ``Camera.render`` / ``render_async`` are the drop-in entry points;
everything they compute happens on the GPU behind the C ABI (no CPU fallback).
"""
import ctypes as C
import math

import numpy as np

from . import ffi
from .transform import Matrix


class Canvas:
    """src/canvas.rs:16-51: ``pixels`` is row-major ``y*width+x``, 3 x f64 per pixel."""

    def __init__(self, w, h, pixels=None, hit_mask=None, stats=None):
        self.width, self.height = w, h
        self.pixels = pixels if pixels is not None else np.zeros((h, w, 3), dtype=np.float64)
        self.hit_mask = hit_mask
        self.stats = stats
        self.gamma = 1.0

    def write_pixel(self, x, y, c):
        self.pixels[y, x] = tuple(c)

    def get_pixel(self, x, y):
        from .scene import Color
        p = self.pixels[y, x]
        return Color(p[0], p[1], p[2])

    def to_u8(self):
        """Color::scale's 8-bit rule: ``clamp((c*255.0) as i32, 0, 255)`` — truncation toward
        zero, saturating, NaN -> 0 (src/color.rs:100-114)."""
        v = self.pixels.astype(np.float64) * 255.0
        v = np.nan_to_num(v, nan=0.0, posinf=3e9, neginf=-3e9)
        return np.clip(np.trunc(v), 0, 255).astype(np.uint8)

    def write_to_file_simple(self, file_name):
        """PPM P3, one image row per line (src/canvas.rs:86-109)."""
        u8 = self.to_u8()
        with open(file_name, "w") as f:
            f.write(f"P3\n{self.width} {self.height}\n255\n")
            for row in u8:
                f.write(" ".join(str(int(v)) for v in row.reshape(-1)))
                f.write("\n")


class Renderer:
    """Owns one ``rtc_context`` (streams, device scene, frame buffers).  Not thread-safe per
    instance; create one per thread for concurrent renders (include/rtc.h)."""

    def __init__(self, devices=None):
        self._lib = ffi.lib()
        self._ctx = C.c_void_p()
        if devices is None:
            devices = [0]
        ids = (C.c_int * len(devices))(*devices)
        ffi.check(self._lib.rtc_context_create(ids, len(devices), C.byref(self._ctx)))
        self.devices = list(devices)
        self._world_key = None

    def close(self):
        if self._ctx:
            self._lib.rtc_context_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._ctx

    def upload(self, world):
        light, shapes, n = world.flatten()
        if getattr(world, "extra_lights", None):
            lights = world.flatten_lights()
            ffi.check(self._lib.rtc_scene_upload_lights(self._ctx, lights, len(lights), shapes, n))
        else:
            ffi.check(self._lib.rtc_scene_upload(self._ctx, C.byref(light), shapes, n))

    def upload_flat(self, light, shapes, n):
        ffi.check(self._lib.rtc_scene_upload(self._ctx, C.byref(light), shapes, n))

    @staticmethod
    def make_opts(precision=ffi.PRECISION_F64_PARITY, rect=ffi.RECT_FULL, max_depth=5,
                  out_format=ffi.OUT_F64_RGB, shard_index=0, shard_count=1, tile_rows=16, flags=0):
        o = ffi.RtcRenderOpts()
        o.precision, o.rect, o.max_depth, o.out_format = precision, rect, max_depth, out_format
        o.shard_index, o.shard_count, o.tile_rows, o.flags = shard_index, shard_count, tile_rows, flags
        return o

    def render(self, cam_desc, opts, want_mask=True, out=None, want_stats=True):
        """rtc_render into host numpy buffers.  Returns (rgb, hit_mask, stats); stats is None with
        want_stats=False (the library then launches the kernels without their work counters, as the
        Rust shim of INTEGRATION.md does)."""
        h, w = cam_desc.vsize, cam_desc.hsize
        dt = {ffi.OUT_F64_RGB: np.float64, ffi.OUT_F32_RGB: np.float32, ffi.OUT_U8_RGB: np.uint8}[opts.out_format]
        rgb = out if out is not None else np.empty((h, w, 3), dtype=dt)
        assert rgb.dtype == dt and rgb.flags["C_CONTIGUOUS"] and rgb.size == h * w * 3
        mask = np.empty((h, w), dtype=np.uint8) if want_mask else None
        stats = ffi.RtcStats() if want_stats else None
        ffi.check(self._lib.rtc_render(
            self._ctx, C.byref(cam_desc), C.byref(opts), rgb.ctypes.data_as(C.c_void_p),
            mask.ctypes.data_as(C.c_void_p) if mask is not None else None, C.byref(stats) if want_stats else None))
        return rgb, mask, stats

    def render_device(self, cam_desc, opts, d_rgb, d_mask=None, stream=None, stats=None):
        """rtc_render_device: d_rgb / d_mask are raw device pointers (ints)."""
        ffi.check(self._lib.rtc_render_device(
            self._ctx, C.byref(cam_desc), C.byref(opts), C.c_void_p(d_rgb),
            C.c_void_p(d_mask) if d_mask else None, C.c_void_p(stream) if stream else None,
            C.byref(stats) if stats is not None else None))

    def render_device_synced(self, cam_desc, opts, d_rgb, stream, wait_flag=None, wait_value=0, done_flag=None, done_value=0, d_mask=None):
        """rtc_render_device_synced: render ordered after *wait_flag >= wait_value, done_value release-stored to *done_flag
        when the pixels are out (raw device pointers; peer-mapped words allowed)."""
        ffi.check(self._lib.rtc_render_device_synced(
            self._ctx, C.byref(cam_desc), C.byref(opts), C.c_void_p(d_rgb), C.c_void_p(d_mask) if d_mask else None, C.c_void_p(stream),
            C.c_void_p(wait_flag) if wait_flag else None, wait_value, C.c_void_p(done_flag) if done_flag else None, done_value))

    def render_device_noalloc(self, cam_desc, opts, d_rgb, stream, d_mask=None):
        """rtc_render_device_noalloc: the small-scene fast kernel with its scratch in a static device array — no cudaMalloc."""
        ffi.check(self._lib.rtc_render_device_noalloc(self._ctx, C.byref(cam_desc), C.byref(opts), C.c_void_p(d_rgb),
                                                      C.c_void_p(d_mask) if d_mask else None, C.c_void_p(stream)))

    def assemble_device(self, cam_desc, opts, d_gathered, shard_stride_rows, elem_kind, d_frame, stream=None):
        ffi.check(self._lib.rtc_assemble_device(
            self._ctx, C.byref(cam_desc), C.byref(opts), C.c_void_p(d_gathered), shard_stride_rows,
            elem_kind, C.c_void_p(d_frame), C.c_void_p(stream) if stream else None))

    def pack_rgba8_device(self, d_frame, frame_format, n_pixels, gamma, d_out_rgba, stream=None):
        """rtc_pack_rgba8_device: Canvas::to_imgbuf on the device (gamma, truncate-then-clamp, alpha 255)."""
        ffi.check(self._lib.rtc_pack_rgba8_device(self._ctx, C.c_void_p(d_frame), frame_format, n_pixels, C.c_float(gamma),
                                                  C.c_void_p(d_out_rgba), C.c_void_p(stream) if stream else None))

    def download_frame(self, d_frame, dev_format, hsize, vsize, out, stream=None):
        """rtc_download_frame: a finished device frame into the host numpy array `out` (same format, or
        f32 -> f64 widening), after the work already enqueued on `stream`."""
        fmt = {np.dtype(np.float64): ffi.OUT_F64_RGB, np.dtype(np.float32): ffi.OUT_F32_RGB, np.dtype(np.uint8): ffi.OUT_U8_RGB}[out.dtype]
        assert out.flags["C_CONTIGUOUS"] and out.size == hsize * vsize * 3
        ffi.check(self._lib.rtc_download_frame(self._ctx, C.c_void_p(d_frame), dev_format, hsize, vsize, fmt,
                                               out.ctypes.data_as(C.c_void_p), C.c_void_p(stream) if stream else None))
        return out

    def ipc_export(self, d_ptr):
        """-> (handle bytes, offset) for the allocation containing device pointer d_ptr."""
        h = (C.c_ubyte * ffi.IPC_HANDLE_BYTES)()
        off = C.c_uint64()
        ffi.check(self._lib.rtc_ipc_export(self._ctx, C.c_void_p(d_ptr), h, C.byref(off)))
        return bytes(h), off.value

    def ipc_open(self, handle, offset):
        """Map a peer process's buffer; -> device pointer (int) valid in this process."""
        h = (C.c_ubyte * ffi.IPC_HANDLE_BYTES).from_buffer_copy(handle)
        out = C.c_void_p()
        ffi.check(self._lib.rtc_ipc_open(self._ctx, h, offset, C.byref(out)))
        return out.value

    def ipc_close(self, d_ptr, offset):
        ffi.check(self._lib.rtc_ipc_close(self._ctx, C.c_void_p(d_ptr), offset))

    def signal_device(self, d_flag, value, stream=None):
        ffi.check(self._lib.rtc_signal_device(self._ctx, C.c_void_p(d_flag), value, C.c_void_p(stream) if stream else None))

    def wait_device(self, d_flags, n, value, stream=None):
        ffi.check(self._lib.rtc_wait_device(self._ctx, C.c_void_p(d_flags), n, value, C.c_void_p(stream) if stream else None))

    def wait_signal_device(self, d_flags, n, value, d_then_flag, then_value, stream=None):
        ffi.check(self._lib.rtc_wait_signal_device(self._ctx, C.c_void_p(d_flags), n, value, C.c_void_p(d_then_flag), then_value,
                                                   C.c_void_p(stream) if stream else None))

    def set_wait_timeout(self, seconds):
        ffi.check(self._lib.rtc_set_wait_timeout(self._ctx, C.c_double(seconds)))

    def wait_result(self, stream=None):
        """rtc_wait_result: synchronises the stream; raises RtcError(RTC_ERR_TIMEOUT) naming the missing flag when an
        earlier rtc_wait_device ran out of time."""
        missing = C.c_int(-1)
        ffi.check(self._lib.rtc_wait_result(self._ctx, C.c_void_p(stream) if stream else None, C.byref(missing)))

    def register_host_buffer(self, array):
        """rtc_register_host_buffer for a numpy array the caller keeps alive (a long-lived Canvas frame): rtc_render then
        delivers into it by DMA, without staging or host threads."""
        assert array.flags["C_CONTIGUOUS"]
        ffi.check(self._lib.rtc_register_host_buffer(self._ctx, array.ctypes.data_as(C.c_void_p), array.nbytes))

    def unregister_host_buffer(self, array):
        ffi.check(self._lib.rtc_unregister_host_buffer(self._ctx, array.ctypes.data_as(C.c_void_p)))

    def delivery_info(self):
        """rtc_delivery_info: {"d2h_bytes", "bands", "bands_direct"} of the last render() on this context."""
        b, n, k = C.c_uint64(0), C.c_int(0), C.c_int(0)
        ffi.check(self._lib.rtc_delivery_info(self._ctx, C.byref(b), C.byref(n), C.byref(k)))
        return {"d2h_bytes": int(b.value), "bands": int(n.value), "bands_direct": int(k.value)}

    def shard_rows(self, cam_desc, opts):
        return self._lib.rtc_shard_rows(C.byref(cam_desc), C.byref(opts))

    def measure_peak_tflops(self, precision):
        v = C.c_double()
        ffi.check(self._lib.rtc_measure_peak_tflops(self._ctx, precision, C.byref(v)))
        return v.value


_default_renderer = None


def default_renderer():
    global _default_renderer
    if _default_renderer is None:
        _default_renderer = Renderer([0])
    return _default_renderer


class Camera:
    """src/camera.rs:16-58."""
    MAX_REFLECTIONS = 5  # :31

    def __init__(self, hsize, vsize, fov):
        self.hsize, self.vsize, self.fov = int(hsize), int(vsize), float(fov)
        half_view = math.tan(self.fov / 2.0)
        aspect = float(self.hsize) / float(self.vsize)
        if aspect < 1.0:
            self.half_height = half_view
            self.half_width = half_view * aspect
        else:
            self.half_width = half_view
            self.half_height = half_view / aspect
        self.pixel_size = (self.half_width * 2.) / float(self.hsize)
        self.view_transform_inv = Matrix.identity()
        self.antialiasing_samples = 1
        self.aa_seed = 0   # seed of the resample offsets (samples >= 2; include/rtc.h "Resample offsets")

    @staticmethod
    def new(hsize, vsize, fov):
        return Camera(hsize, vsize, fov)

    @staticmethod
    def new_with_transform(hsize, vsize, fov, m):  # :33-37
        c = Camera(hsize, vsize, fov)
        c.view_transform_inv = m.inverse()
        return c

    def set_samples(self, n):  # :60 (n: u8)
        if not 0 <= int(n) <= 255:
            raise ValueError("antialiasing_samples is a u8 (src/camera.rs:26)")
        self.antialiasing_samples = int(n)

    def flatten(self):
        d = ffi.RtcCamera()
        d.hsize, d.vsize = self.hsize, self.vsize
        d.half_width, d.half_height, d.pixel_size = self.half_width, self.half_height, self.pixel_size
        d.view_inv[:] = self.view_transform_inv.to_row_major()
        d.samples = self.antialiasing_samples
        d.aa_seed = self.aa_seed & 0xffffffff
        return d

    def _render(self, world, rect, precision, renderer, max_depth):
        r = renderer or default_renderer()
        r.upload(world)
        opts = Renderer.make_opts(precision=precision, rect=rect, max_depth=max_depth)
        rgb, mask, stats = r.render(self.flatten(), opts)
        return Canvas(self.hsize, self.vsize, rgb, mask, stats)

    def render(self, world, precision=ffi.PRECISION_F64_PARITY, renderer=None, max_depth=MAX_REFLECTIONS):
        """Camera::render (:116-126) — including its loop bounds: last row/column stay black."""
        return self._render(world, ffi.RECT_REFERENCE_RENDER, precision, renderer, max_depth)

    def render_async(self, world, precision=ffi.PRECISION_F64_PARITY, renderer=None, max_depth=MAX_REFLECTIONS):
        """Camera::render_async (:144-160) — the full image."""
        return self._render(world, ffi.RECT_FULL, precision, renderer, max_depth)

    render_async1 = render_async  # :128-142
