"""Kernel time of the C4 frame (10k shapes, 4K) in fast mode, for one library build (RTC_B200_LIB selects a variant).
usage: [RTC_B200_LIB=path/to/lib.so] python tools/perf_c4.py [width height]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
r = rtc.Renderer([0])
r.upload(rtc.scenes.synthetic_scene())
d = rtc.scenes.synthetic_camera(w, h).flatten()
buf = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
ms = []
for _ in range(3):
    st = ffi.RtcStats()
    r.render_device(d, rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB), buf.data_ptr(), None, None, st)
    ms.append(st.kernel_ms)
print(f"{os.path.basename(ffi.LIB_PATH):24s} C4 {w}x{h} f32 kernel_ms min={min(ms):9.3f} rays={st.rays} checksum={float(buf.double().sum()):.6f}", flush=True)
