"""C4 (10k shapes) f32 general kernel: bounding-sphere filter on / off — kernel time and bit-equality of the frames.
usage: python tools/perf_c4.py [width] [height] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
w = int(sys.argv[1]) if len(sys.argv) > 1 else 3840
h = int(sys.argv[2]) if len(sys.argv) > 2 else 2160
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
r = rtc.Renderer([0])
r.upload(rtc.scenes.synthetic_scene())
cam = rtc.scenes.synthetic_camera(w, h).flatten()
frames = {}
for name, flags in (("filter", 0), ("exact-only", ffi.OPT_NO_BOUND_FILTER)):
    buf = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
    hit = torch.empty((h, w), dtype=torch.uint8, device="cuda")
    ms = []
    for _ in range(reps):
        st = ffi.RtcStats()
        r.render_device(cam, rtc.Renderer.make_opts(precision=1, out_format=ffi.OUT_F32_RGB, flags=flags), buf.data_ptr(), hit.data_ptr(), None, st)
        ms.append(st.kernel_ms)
    frames[name] = (buf, hit, st.rays)
    print(f"C4 {w}x{h} f32 {name:10s} kernel_ms min={min(ms):9.3f} rays={st.rays} Mrays/s={st.rays / min(ms) / 1e3:8.1f}", flush=True)
a, b = frames["filter"], frames["exact-only"]
print("frames bit-identical:", torch.equal(a[0], b[0]), " hit masks:", torch.equal(a[1], b[1]), " rays:", a[2] == b[2])
