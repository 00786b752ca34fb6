"""Latency floor of one render launch: tiny frames (f32 small kernel), depth 0 and 5, events around the launch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
r = rtc.Renderer([0])
r.upload(rtc.scenes.reflect_refract_scene())
buf = torch.empty((2160, 3840, 3), dtype=torch.float32, device="cuda")
for (w, h) in ((8, 4), (64, 32), (256, 128), (960, 540), (1920, 1080)):
    cam = rtc.scenes.reflect_refract_camera(w, h).flatten()
    for depth in (0, 5):
        opts = rtc.Renderer.make_opts(precision=1, out_format=ffi.OUT_F32_RGB, max_depth=depth)
        ms = []
        for _ in range(10):
            st = ffi.RtcStats()
            r.render_device(cam, opts, buf.data_ptr(), None, None, st)
            ms.append(st.kernel_ms)
        print(f"{w}x{h} depth {depth}: kernel_ms min={min(ms)*1e3:.1f} us med={sorted(ms)[5]*1e3:.1f} us rays={st.rays}")
