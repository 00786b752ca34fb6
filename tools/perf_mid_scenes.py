"""Parity-kernel time of mid-size scenes (100 ... 3000 random spheres + planes, 1920x1080): where the streams should move from
shared memory (few resident blocks once they are large) to global memory (8 blocks per SM).  RTC_B200_LIB selects a variant."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
r = rtc.Renderer([0])
cam = rtc.scenes.synthetic_camera(1920, 1080).flatten()
buf = torch.empty((1080, 1920, 3), dtype=torch.float64, device="cuda")
for n in [int(a) for a in sys.argv[1:]] or (100, 200, 400, 800, 1500, 3000):
    r.upload_flat(*rtc.scenes.synthetic_scene(n_spheres=n, seed=0xB200).flatten())
    out = []
    for prec in (0, 1):
        ms = []
        for _ in range(3):
            st = ffi.RtcStats()
            r.render_device(cam, rtc.Renderer.make_opts(precision=prec, out_format=ffi.OUT_F64_RGB), buf.data_ptr(), None, None, st)
            ms.append(st.kernel_ms)
        out.append(min(ms))
    print(f"{n:5d} spheres: f64 {out[0]:8.3f} ms   fast {out[1]:8.3f} ms   checksum {float(buf.sum()):.6f}", flush=True)
