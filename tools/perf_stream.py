"""C3 frame timing of the f32 small-scene kernel the way bench.py launches it: on a side stream,
CUDA events around each launch, no counters / hit mask (the kAux = false instantiation), L2 flushed
between frames.  Also prints the counted launch (stream = NULL, kAux = true) for comparison.
usage: python tools/perf_stream.py [frames] [flags]    time the library ffi.py resolves (RTC_B200_LIB honoured)
       python tools/perf_stream.py --variants [glob] [flags]  one child process per build/variants/<glob>"""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)

if len(sys.argv) > 1 and sys.argv[1] == "--variants":
    pat = sys.argv[2] if len(sys.argv) > 2 else "lib_*.so"
    for lib in sorted(glob.glob(os.path.join(root, "build", "variants", pat))):
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "7"] + sys.argv[3:4], env=dict(os.environ, RTC_B200_LIB=lib),
                             stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        print(os.path.basename(lib), out.stdout.strip() or out.stderr.strip()[-300:], flush=True)
    sys.exit(0)

import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 7
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0   # e.g. 4 = RTC_OPT_LINEAR_TILE_ORDER
w, h = 3840, 2160
r = rtc.Renderer([0])
r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(w, h).flatten()
buf = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
side = torch.cuda.Stream()
opts = rtc.Renderer.make_opts(precision=1, out_format=ffi.OUT_F32_RGB, flags=flags)
ms = []
with torch.cuda.stream(side):
    for i in range(frames + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        r.render_device(cam, opts, buf.data_ptr(), None, side.cuda_stream, None)
        e1.record(side)
        side.synchronize()
        if i >= 2:
            ms.append(e0.elapsed_time(e1))
st = ffi.RtcStats()
cms = []
for _ in range(3):
    r.render_device(cam, opts, buf.data_ptr(), None, None, st)
    cms.append(st.kernel_ms)
print(f"plain min={min(ms):.4f} med={sorted(ms)[len(ms) // 2]:.4f} ms | counted min={min(cms):.4f} ms rays={st.rays} "
      f"| checksum={float(buf.double().sum()):.6f}")
