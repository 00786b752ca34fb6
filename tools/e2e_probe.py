"""Where the end-to-end time of rtc_render (host buffers) goes: stats of a few 4K frames per output format.
usage: python tools/e2e_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtc_b200 as rtc
from rtc_b200 import ffi
def _timed(fn):
    t = time.perf_counter(); fn(); return (time.perf_counter() - t) * 1e3


r = rtc.Renderer([0]); r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(3840, 2160).flatten()
bufs = {ffi.OUT_F64_RGB: np.empty((2160, 3840, 3), np.float64), ffi.OUT_F32_RGB: np.empty((2160, 3840, 3), np.float32), ffi.OUT_U8_RGB: np.empty((2160, 3840, 3), np.uint8)}
for name, prec, fmt in (("f32 kernel -> f64 Canvas", 1, ffi.OUT_F64_RGB), ("f32 kernel -> f32", 1, ffi.OUT_F32_RGB),
                        ("f32 kernel -> u8", 1, ffi.OUT_U8_RGB), ("f64 kernel -> f64 Canvas", 0, ffi.OUT_F64_RGB)):
    opts = rtc.Renderer.make_opts(precision=prec, out_format=fmt)
    best = None
    for _ in range(6):
        t = time.perf_counter()
        out, hit, st = r.render(cam, opts, want_mask=False, out=bufs[fmt])
        dt = (time.perf_counter() - t) * 1e3
        if best is None or dt < best[0]: best = (dt, st)
    plain = min(_timed(lambda: r.render(cam, opts, want_mask=False, out=bufs[fmt], want_stats=False)) for _ in range(6))
    dt, st = best
    print(f"{name:26s} wall {dt:7.3f} ms (no stats: {plain:7.3f}) | total {st.total_ms:7.3f} kernel {st.kernel_ms:6.3f} after-kernel {st.d2h_ms:6.3f} host-delivery {st.gather_ms:6.3f}")
flat = rtc.scenes.reflect_refract_scene().flatten()
ts = []
for _ in range(20):
    t = time.perf_counter(); r.upload_flat(*flat); ts.append((time.perf_counter() - t) * 1e3)
print(f"rtc_scene_upload (13 shapes): min {min(ts):.3f} ms  median {sorted(ts)[10]:.3f} ms")
host = bufs[ffi.OUT_F64_RGB]
oe = rtc.Renderer.make_opts(precision=1, out_format=ffi.OUT_F64_RGB)
ts = []
for _ in range(12):
    t = time.perf_counter(); r.upload_flat(*flat); r.render(cam, oe, want_mask=False, out=host, want_stats=False); ts.append((time.perf_counter() - t) * 1e3)
print(f"upload + render (the bench's e2e step): min {min(ts):.3f} ms  median {sorted(ts)[6]:.3f} ms")
