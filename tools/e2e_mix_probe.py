"""Mixed delivery of rtc_render: end-to-end time of the C3 4K frame into a registered f64 Canvas buffer for
every split — bands of 8 / 16 and how many of them go by f64 DMA — with the measurement build of the library
(`make -C raytracer-challenge_b200/csrc probe`, which reads RTC_MIX_BANDS / RTC_MIX_K; the product library reads no environment)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["RTC_B200_LIB"] = os.path.join(ROOT, "raytracer-challenge_b200", "csrc", "librtc_b200_probe.so")
import numpy as np
import torch

import rtc_b200 as rtc
from rtc_b200 import ffi

W, H = 3840, 2160
r = rtc.Renderer([0])
world = rtc.scenes.reflect_refract_scene()
flat = world.flatten()
camd = rtc.scenes.reflect_refract_camera(W, H).flatten()
o = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, rect=ffi.RECT_FULL, max_depth=5, out_format=ffi.OUT_F64_RGB)
host = np.zeros((H, W, 3))
r.register_host_buffer(host)


def run(n=12):
    ts = []
    for i in range(n + 2):
        a = time.perf_counter()
        r.upload_flat(*flat)
        r.render(camd, o, want_mask=False, out=host, want_stats=False)
        ts.append(time.perf_counter() - a)
    ts = sorted(ts[2:])
    return ts[0] * 1e3, ts[len(ts) // 2] * 1e3


os.environ["RTC_MIX_TRACE"] = "1"
print("adaptive, the library's own plan:", run(20), r.delivery_info(), flush=True)
del os.environ["RTC_MIX_TRACE"]
want = None
for bands in (8,):
    os.environ["RTC_MIX_BANDS"] = str(bands)
    for k in range(0, 3):
        os.environ["RTC_MIX_K"] = str(k)
        lo, med = run()
        if want is None:
            want = host.copy()
        ok = np.array_equal(host, want)
        print(f"bands {bands:2d}  direct {k:2d}  min {lo:6.3f}  median {med:6.3f} ms  {r.delivery_info()}  frame identical: {ok}", flush=True)
del os.environ["RTC_MIX_K"]
del os.environ["RTC_MIX_BANDS"]
for plan in ("1,1,1,1,1,1,1,1", "1,2,2,2,2,2,2,2,1", "1,1,2,2,2,2,2,2,1,1", "1,3,3,3,3,2,1", "1,2,3,3,3,2,1,1", "2,3,3,3,3,1,1", "1,2,2,2,2,2,2,1,1,1"):
    os.environ["RTC_MIX_PLAN"] = plan
    for k in (0, 1, 2, 3):
        os.environ["RTC_MIX_K"] = str(k)
        lo, med = run()
        print(f"plan {plan:24s} direct bands {k}  min {lo:6.3f}  median {med:6.3f} ms  {r.delivery_info()}  frame identical: {np.array_equal(host, want)}", flush=True)
del os.environ["RTC_MIX_K"]
del os.environ["RTC_MIX_PLAN"]
