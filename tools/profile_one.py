"""Render a few frames of the reflect/refract scene through the C ABI (for ncu / sanitizer runs).
usage: python tools/profile_one.py [f32|f64] [width] [height] [frames] [samples] [room|c4]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # device memory only

import rtc_b200 as rtc
from rtc_b200 import ffi

prec = ffi.PRECISION_F32_FAST if (len(sys.argv) < 2 or sys.argv[1] == "f32") else ffi.PRECISION_F64_PARITY
w = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
h = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 3
samples = int(sys.argv[5]) if len(sys.argv) > 5 else 1
r = rtc.Renderer([0])
scene = sys.argv[6] if len(sys.argv) > 6 else "room"
if scene == "c4":
    r.upload(rtc.scenes.synthetic_scene())
    cam = rtc.scenes.synthetic_camera(w, h, samples=samples).flatten()
else:
    r.upload(rtc.scenes.reflect_refract_scene())
    cam = rtc.scenes.reflect_refract_camera(w, h, samples=samples).flatten()
fmt = ffi.OUT_F32_RGB if prec == ffi.PRECISION_F32_FAST else ffi.OUT_F64_RGB
opts = rtc.Renderer.make_opts(precision=prec, out_format=fmt)
buf = torch.empty((h, w, 3), dtype=torch.float32 if fmt == ffi.OUT_F32_RGB else torch.float64, device="cuda")
for _ in range(frames):
    st = ffi.RtcStats()
    r.render_device(cam, opts, buf.data_ptr(), None, None, st)
    print(f"{w}x{h} prec={prec} kernel_ms={st.kernel_ms:.3f} rays={st.rays} Mrays/s={st.rays / st.kernel_ms / 1e3:.1f}")
