"""rtc_render of the 10k-shape scene at 4K into a host f64 frame: wall time vs kernel time.  usage: python tools/e2e_c4_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rtc_b200 as rtc
from rtc_b200 import ffi
r = rtc.Renderer([0]); r.upload(rtc.scenes.synthetic_scene())
cam = rtc.scenes.synthetic_camera(3840, 2160).flatten()
out = np.empty((2160, 3840, 3), np.float64)
for _ in range(3):
    t = time.perf_counter()
    _, _, st = r.render(cam, rtc.Renderer.make_opts(precision=1), want_mask=False, out=out)
    print(f"wall {(time.perf_counter() - t) * 1e3:.1f} ms  kernel {st.kernel_ms:.1f} ms  launches {st.kernel_launches}", flush=True)
