"""(RTC_B200_LIB selects a kernel-variant library for A/B runs.)  Kernel timings of the BASELINE.json configs through rtc_render_device on one GPU:
C2 (1920x1080 room), C4 (10k shapes, 4K), C5 (8K, 2x2 supersample).  usage: python tools/perf_configs.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
r = rtc.Renderer([0])
room = rtc.scenes.reflect_refract_scene().flatten()
c4 = rtc.scenes.synthetic_scene().flatten()
two = rtc.scenes.reflect_refract_scene()
two.add_light(rtc.Light(rtc.Color(0.3, 0.5, 0.9), (3.5, 3.0, -3.0)))
cases = [("C3 room 3840x2160", room, rtc.scenes.reflect_refract_camera(3840, 2160), 5),
         ("C3 room 3840x2160 samples=4", room, rtc.scenes.reflect_refract_camera(3840, 2160, samples=4), 3),   # resample pass
         ("C5 room 7680x4320 samples=4", room, rtc.scenes.reflect_refract_camera(7680, 4320, samples=4), 2),
         ("C3 room 3840x2160 2 lights", (two.flatten_lights(), two.flatten()[1], two.flatten()[2]), rtc.scenes.reflect_refract_camera(3840, 2160), 3),
         ("C2 room 1920x1080", room, rtc.scenes.reflect_refract_camera(1920, 1080), 5),
         ("C5 room 7680x4320 ss2x2", room, rtc.scenes.reflect_refract_camera(7680, 4320, samples=0), 3),
         ("C4 10k shapes 3840x2160", c4, rtc.scenes.synthetic_camera(3840, 2160), 2)]
for name, flat, cam, reps in cases:
    if len(flat[0]) > 1 if hasattr(flat[0], "__len__") else False:
        ffi.check(ffi.lib().rtc_scene_upload_lights(r.handle, flat[0], len(flat[0]), flat[1], flat[2]))
    else:
        r.upload_flat(*flat)
    d = cam.flatten()
    buf = torch.empty((d.vsize, d.hsize, 3), dtype=torch.float64, device="cuda")
    for pname, prec, fmt in (("f32", 1, ffi.OUT_F32_RGB), ("f64", 0, ffi.OUT_F64_RGB)):
        ms = []
        for _ in range(reps):
            st = ffi.RtcStats()
            r.render_device(d, rtc.Renderer.make_opts(precision=prec, out_format=fmt), buf.data_ptr(), None, None, st)
            ms.append(st.kernel_ms)
        print(f"{name:28s} {pname} kernel_ms min={min(ms):10.3f} rays={st.rays:12d} Mrays/s={st.rays / min(ms) / 1e3:10.0f}", flush=True)
