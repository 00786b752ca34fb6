"""Where does the f32 two-light frame differ from the oracle by more than 1/255?  usage: python tools/diag_lights.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtc_b200 as rtc, orc
import test_lights as T
r = rtc.Renderer([0])
cam = rtc.scenes.reflect_refract_camera(640, 320)
for depth in (0, 1, 5):
    for name, w in (("one", rtc.scenes.reflect_refract_scene()), ("two", T._two_light_room(rtc))):
        r.upload(w)
        want, wh, cnt = orc.OracleWorld(w).render(cam, max_depth=depth, threads=os.cpu_count())
        for general in (0, rtc.ffi.OPT_F32_GENERAL_KERNEL):
            got, hit, st = r.render(cam.flatten(), rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F32_FAST, max_depth=depth, flags=general))
            bad = np.abs(orc.quantize_u8(got).astype(int) - orc.quantize_u8(want).astype(int)).max(axis=2) > 1
            ys, xs = np.nonzero(bad)
            print(name, "depth", depth, "general" if general else "auto", "bad", int(bad.sum()), "of", bad.size,
                  "rows", (ys.min(), ys.max()) if len(ys) else None, "cols", (xs.min(), xs.max()) if len(xs) else None, flush=True)
            if name == "two" and general and depth == 5 and len(ys):
                for k in range(0, len(ys), max(1, len(ys) // 12)):
                    y, x = ys[k], xs[k]
                    print("   ", (x, y), got[y, x], want[y, x])
