"""Quick kernel timing of the C3 frame through rtc_render_device (small / general f32 kernel, f64).
usage: python tools/perf_quick.py [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 5
w, h = 3840, 2160
r = rtc.Renderer([0])
r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(w, h).flatten()
buf = torch.empty((h, w, 3), dtype=torch.float64, device="cuda")
for name, prec, fmt, flags in (("f32-small", 1, ffi.OUT_F32_RGB, 0), ("f32-general", 1, ffi.OUT_F32_RGB, ffi.OPT_F32_GENERAL_KERNEL),
                               ("f64", 0, ffi.OUT_F64_RGB, 0)):
    opts = rtc.Renderer.make_opts(precision=prec, out_format=fmt, flags=flags)
    ms = []
    for _ in range(frames):
        st = ffi.RtcStats()
        r.render_device(cam, opts, buf.data_ptr(), None, None, st)
        ms.append(st.kernel_ms)
    print(f"{name:12s} kernel_ms min={min(ms):.3f} med={sorted(ms)[len(ms)//2]:.3f} rays={st.rays} hits={st.shaded_hits} Mrays/s={st.rays / min(ms) / 1e3:.0f}")
