"""Time kernel variants: for every build/variants/lib_*.so run a child process with RTC_B200_LIB set.
usage: python tools/perf_variants.py f64|f32 [glob]"""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
which = sys.argv[1] if len(sys.argv) > 1 else "f64"
pat = sys.argv[2] if len(sys.argv) > 2 else "lib_*.so"
child = r'''
import os, sys
sys.path.insert(0, %r)
import torch, rtc_b200 as rtc
from rtc_b200 import ffi
r = rtc.Renderer([0]); r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(3840, 2160).flatten()
buf = torch.empty((2160, 3840, 3), dtype=torch.float64, device="cuda")
prec, fmt = (0, ffi.OUT_F64_RGB) if %r == "f64" else (1, ffi.OUT_F32_RGB)
ms = []
for _ in range(5):
    st = ffi.RtcStats(); r.render_device(cam, rtc.Renderer.make_opts(precision=prec, out_format=fmt), buf.data_ptr(), None, None, st); ms.append(st.kernel_ms)
print("%%.3f %%d" %% (min(ms), st.rays))
''' % (root, which)
for lib in sorted(glob.glob(os.path.join(root, "build", "variants", pat))):
    env = dict(os.environ, RTC_B200_LIB=lib)
    out = subprocess.run([sys.executable, "-c", child], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    print(os.path.basename(lib), out.stdout.strip() or out.stderr.strip()[-300:], flush=True)
